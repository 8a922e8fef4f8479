#!/bin/bash
# One compute-sanitizer tool per GPU-box visit:  bash tools/gpu_sanitize.sh memcheck|racecheck|synccheck|initcheck [pytest -k expression]
# Target: the fused-halo tests on 4 virtual ranks (one GPU), the persistent/ring kernels, and, with 2+ GPUs, the peer-memory halo.
TOOL=${1:-memcheck}
SEL=${2:-"virtual_ranks_fused_halo or persistent_kernels_equal_classic or multi_gpu_direct_halo"}
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_persistent.py -m gpu -q -k "$SEL" 2>&1 | tail -3
timeout 1500 compute-sanitizer --tool $TOOL --target-processes all --log-file gpurun_out/sanitizer_$TOOL.log \
    python -m pytest tests/test_gpu_kernels.py tests/test_gpu_persistent.py -m gpu -q -x -k "$SEL" > gpurun_out/sanitizer_${TOOL}_pytest.log 2>&1
echo "exit $?"; tail -n 3 gpurun_out/sanitizer_${TOOL}_pytest.log; grep -c "ERROR SUMMARY" gpurun_out/sanitizer_$TOOL.log; grep "ERROR SUMMARY\|RACECHECK SUMMARY" gpurun_out/sanitizer_$TOOL.log | sort | uniq -c | tail -5
