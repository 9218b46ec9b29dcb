#!/bin/bash
# compute-sanitizer over the small (C1-size) parity cases through the C ABI.  ONE tool per gpurun call (B200_PROFILING.md).
TOOL=${1:-memcheck}
mkdir -p gpurun_out
SEL='test_vs_reference_extension and (c1_cloud or cloud_ragged) or test_gather_and_streamed or test_edge or test_fused_depth_sil'
timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "$SEL" > gpurun_out/sanitizer_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitizer_plain.log; exit 1; }
tail -1 gpurun_out/sanitizer_plain.log
timeout 2400 compute-sanitizer --tool $TOOL --log-file gpurun_out/sanitizer_$TOOL.log --print-limit 20 python -m pytest tests/test_gpu_parity.py -q -x -k "$SEL" > gpurun_out/sanitizer_${TOOL}_pytest.log 2>&1; echo "sanitizer rc=$?"
tail -2 gpurun_out/sanitizer_${TOOL}_pytest.log
tail -5 gpurun_out/sanitizer_$TOOL.log
