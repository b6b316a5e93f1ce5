#!/bin/bash
# One GPU-box pass: gpu tests, smoke, bench (both arms), ncu launch list + full capture of the top kernel.
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 50 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
CMD="python bench.py --steps 20 --warmup 3 --skip-cpu --skip-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gae_rows -s 5 -c 1 -f -o gpurun_out/gae_top $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 600 ncu --set full --clock-control none --cache-control none --import-source on -k regex:"ranged|cluster" -s 40 -c 2 -f -o gpurun_out/per_top $CMD > gpurun_out/ncu_full_per.log 2>&1; echo "ncu per rc=$?"
timeout 300 python scripts/bench_rows.py > gpurun_out/rows.jsonl 2> gpurun_out/rows.err; echo "rows rc=$?"
timeout 300 python scripts/per_bench.py > gpurun_out/per_bench.jsonl 2> gpurun_out/per_bench.err; echo "per_bench rc=$?"
