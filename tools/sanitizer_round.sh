#!/bin/bash
# compute-sanitizer pass, ONE tool per gpurun call (B200_PROFILING.md: several tools in one call have left
# a GPU unusable).  usage: bash tools/sanitizer_round.sh memcheck|racecheck|initcheck
# Selections are the small cases of the parity tests: the sanitizer slows kernels 10-100x.
#   memcheck : every stage, both render precisions incl. the tcgen05 kernels (global / shared accesses), Chamfer on tcgen05 and CUDA cores
#   racecheck: shared-memory hazards of the CUDA-core kernels (stage 1, losses, reprojection, CUDA-core Chamfer, fp32 render);
#              the tcgen05 kernels synchronise through mbarriers / async proxies that racecheck does not model, so they are left out
#   initcheck: the workspace users (render with save=1 on both precisions, Chamfer) -- reads of device memory never written
set -u
TOOL=${1:-memcheck}
OUT=gpurun_out/r2_sanitizer_${TOOL}.log
mkdir -p gpurun_out
case $TOOL in
  memcheck)
    SEL="tests/test_gpu_stage1.py tests/test_gpu_losses.py tests/test_gpu_render.py tests/test_gpu_render_tc.py tests/test_zz_no_alloc.py"
    K="not 28830 and not 60000 and not depth_clouds and not cfg2 and not cfg4 and not full_size and not psnr and not consecutive and not capacity"
    EXTRA="--leak-check no" ;;
  racecheck)
    export NNB_SANITIZER_CUDA_CORES=1
    SEL="tests/test_gpu_stage1.py tests/test_gpu_losses.py tests/test_gpu_render.py"
    K="not 28830 and not 60000 and not depth_clouds and not bf16 and not tcgen05 and not mixed"
    EXTRA="" ;;
  initcheck)
    SEL="tests/test_gpu_render_tc.py tests/test_gpu_render.py tests/test_gpu_losses.py"
    K="(tc_forward or tc_backward or renderer_vs_reference or chamfer_vs_reference or chamfer_indices) and not 28830 and not 60000"
    EXTRA="" ;;
  *) echo "unknown tool $TOOL"; exit 2 ;;
esac
python -m pytest $SEL -q -x -m gpu -k "$K" > gpurun_out/r2_sanitizer_${TOOL}_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/r2_sanitizer_${TOOL}_plain.log; exit 1; }
tail -2 gpurun_out/r2_sanitizer_${TOOL}_plain.log
timeout 1500 compute-sanitizer --tool $TOOL $EXTRA --print-limit 20 --error-exitcode 0 \
    python -m pytest $SEL -q -x -m gpu -k "$K" -p no:cacheprovider > $OUT 2>&1
echo "sanitizer exit code $?" >> $OUT
grep -E "ERROR SUMMARY|passed|failed|RACECHECK SUMMARY|sanitizer exit" $OUT | tail -8
