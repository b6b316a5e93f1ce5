#!/bin/bash
# ncu --set full capture of one launch of each chosen gae_rows variant (C2 shape).
mkdir -p gpurun_out
for v in "$@"; do
  cmd="python scripts/gae_sweep.py --variants $v --iters 2"
  $cmd > gpurun_out/ncu_plain_v$v.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:gae_rows -s 8 -c 1 -f \
      -o gpurun_out/gae_v$v $cmd > gpurun_out/ncu_full_v$v.log 2>&1
  echo "variant $v rc=$?"
done
