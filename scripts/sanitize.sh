#!/bin/bash
# compute-sanitizer passes over the GPU parity tests (the reference has no race
# detection; the CUDA path gets memcheck / racecheck / synccheck here).  Full-size cases are
# deselected: the tools slow kernels down by one to two orders of magnitude.
mkdir -p gpurun_out
SEL='not full_size and not large and not host_pipeline and not 40000 and not cuda_graph and not graphed'
for tool in memcheck synccheck racecheck; do
  timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool $tool --error-exitcode 7 --print-limit 30 \
      --log-file gpurun_out/sanitizer_$tool.log \
      python -m pytest tests -m gpu -x -q -k "$SEL" -p no:cacheprovider > gpurun_out/sanitizer_${tool}_pytest.log 2>&1
  echo "$tool rc=$?" | tee -a gpurun_out/sanitizer_summary.txt
  tail -3 gpurun_out/sanitizer_${tool}_pytest.log
  grep -c "=========     at\|Error:\|ERROR SUMMARY" gpurun_out/sanitizer_$tool.log | sed "s/^/$tool log lines of interest: /"
  grep "ERROR SUMMARY\|RACECHECK SUMMARY" gpurun_out/sanitizer_$tool.log | tail -2
done
