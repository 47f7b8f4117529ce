set -x
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01d_bench.log 2>&1 || exit 1
tail -n 1 gpurun_out/r01d_bench.log | cut -c1-300
ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:^(pack_words|table_|radix_|scan_|rank_columns|filter_build|hamming|csr_|uf_|iota_kernel|big_|replay_|compact_clusters|keep_scatter|hist_values|run_starts)" --csv --log-file gpurun_out/r01d_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01d_ncu1.log 2>&1
echo launches rc=$?
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k "regex:^(hamming1_probe|table_insert|radix_pass|rank_columns|replay_kernel|uf_union|filter_build|pack_words)" -s 13 -c 13 -o gpurun_out/r01d_top python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r01d_ncu2.log 2>&1
echo full rc=$?
ls -la gpurun_out/
