#!/bin/bash
# Round-2 evidence pass on one B200: smoke, ncu launch list of the bench command, full captures of
# the GAE rows kernel and of the lazy PER pair.  Each ncu run only after the same command exited 0.
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r02_smoke.log
CMD="python bench.py --steps 20 --warmup 3 --skip-cpu --skip-e2e --skip-updates"
timeout 300 $CMD > gpurun_out/r02_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/r02_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gae_rows_tma -s 5 -c 1 -f -o gpurun_out/r02_gae_top $CMD > gpurun_out/r02_ncu_full.log 2>&1; echo "ncu gae rc=$?"
timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"ranged|per_weight_cluster" -s 60 -c 2 -f -o gpurun_out/r02_per_top $CMD > gpurun_out/r02_ncu_full_per.log 2>&1; echo "ncu per rc=$?"
ls -la gpurun_out/*.ncu-rep
