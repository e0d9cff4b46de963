#!/bin/bash
# ncu evidence of round 2 (profiles/): launch list of the bench command, then ONE --set full capture of the top kernels
cd "$GRAFT_REPO_ROOT" || exit 1
out=gpurun_out
timeout 600 python bench.py --steps 2 --warmup 1 --skip-extras > $out/r2_prof_bench.json 2> $out/r2_prof_bench.err || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/r2_launches.csv \
  python bench.py --steps 2 --warmup 1 --skip-extras > $out/r2_ncu_launch.log 2>&1
echo "launch list rc=$?"; wc -l $out/r2_launches.csv
# exact passes (7 matching kernels each): skip the 3 warm-up passes of the probe, capture the 4th
TC_NO_SPECULATION=1 timeout 1500 ncu --set full --clock-control none --import-source on \
  -k regex:'k_fk|k_broad|k_gjk|k_epa_wp|k_finalize|k_grad_rows' --launch-skip 21 --launch-count 7 -f -o $out/r2_full \
  python tools/narrow_probe.py 4096 1 > $out/r2_ncu_full.log 2>&1
echo "full capture rc=$?"; ls -la $out/r2_full.ncu-rep
