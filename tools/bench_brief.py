"""stdin: output of bench.py; prints the headline numbers and the stage table of its JSON line."""
import json
import sys

for line in sys.stdin:
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    if "unavailable" in d:
        print(d)
        continue
    r = d.get("roofline") or {}
    print(f"n_gpus {d['n_gpus']} ms/step {d['ms_per_step']:.3f} value {d['value'] / 1e9:.3f} G/s e2e "
          f"{d['e2e'].get('ms_per_step', 0):.2f} ms {d['e2e']['value'] / 1e9:.3f} G/s launches {d.get('gpu_launches')}")
    if r:
        print(f"roofline {r.get('stage')} [{r.get('kernel')}] {r.get('achieved', 0):.0f}/{r.get('peak', 0):.0f} GB/s frac {r.get('frac', 0):.3f} "
              f"kernel_ms {r.get('kernel_ms', 0):.3f} traffic {r.get('traffic')}")
        print(" ".join(f"{k}={v}" for k, v in (r.get("stages_ms") or {}).items()))
    if d.get("cpu_baseline"):
        print("cpu", d["cpu_baseline"])
    print("clocks", d.get("clocks"), "result", (d.get("detail") or d["config"]).get("result"))
    if r.get("kernels_8d"):
        print("8(d) families:", {k: (v["ms"], round(v["frac"], 3) if v["frac"] else None) for k, v in r["kernels_8d"].items()})
        print("pass_frac", round(r.get("pass_frac", 0), 3), "traffic_ratio", r.get("traffic_ratio"), "step", r.get("step"))
    if d.get("parity"):
        print("parity", d["parity"])
    for leg in d.get("cli_e2e") or []:
        print("cli", leg["config"], leg["tuples"], leg["seconds"], "s", round(leg["pairs_per_s"]), "pairs/s", leg["breakdown_s"],
              leg["outputs_sha256"][:12], leg.get("identical_to_other_arm"))
