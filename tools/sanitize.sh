#!/bin/bash
# compute-sanitizer targets for the kernel library. Run on a GPU box, one tool per call:
#   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck [gemm|kernels|bert]
# gemm    : the stand-alone GEMM harness (no Python), numerics cases only
# kernels : tests/test_gpu_kernels.py (every kernel against its PyTorch oracle)
# bert    : one end-to-end native BERT test
# NOTE: on the shared B200 pool used for this project compute-sanitizer is closed by the operators (the wrapper
# refuses to start it); the script is kept for boxes where it is available.
# Output goes to gpurun_out/sanitize_<tool>_<target>.log; the exit code is the sanitizer's (non-zero on findings).
set -e
TOOL=${1:-memcheck}
TARGET=${2:-gemm}
mkdir -p gpurun_out
LOG=gpurun_out/sanitize_${TOOL}_${TARGET}.log
case "$TARGET" in
  gemm)
    [ -x build/gemm_test ] || nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
        --expt-relaxed-constexpr csrc/tests/gemm_test.cu csrc/gemm_sm100.cu csrc/lib.cu -o build/gemm_test -lcuda
    CMD="build/gemm_test --no-perf" ;;
  kernels) CMD="python -m pytest tests/test_gpu_kernels.py -x -q -k 'not 4096'" ;;
  bert)    CMD="python -m pytest tests/test_gpu_bert.py -x -q -k test_native_loss_and_grads_match_torch" ;;
  *) echo "unknown target $TARGET"; exit 2 ;;
esac
echo "compute-sanitizer --tool $TOOL $CMD" | tee "$LOG"
# --error-exitcode makes findings fail the run; kernels here use bounded waits, so a protocol bug traps
eval compute-sanitizer --tool "$TOOL" --error-exitcode 9 --print-limit 50 $CMD >> "$LOG" 2>&1
RC=$?
tail -5 "$LOG"
exit $RC
