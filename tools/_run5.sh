mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:^(edit_replay_tasks|edit_candidates_shift)" -c 2 -o gpurun_out/r01k_edit python tools/stage_table.py 3 20000000 1 > gpurun_out/r01k_edit_ncu.log 2>&1
echo rc=$?
tail -3 gpurun_out/r01k_edit_ncu.log
