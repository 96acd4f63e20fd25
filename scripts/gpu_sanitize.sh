# compute-sanitizer memcheck + racecheck over the smoke path and the parity tests that drive every queue path
# of the point pass, the compaction kernels and the erode kernels; logs under gpurun_out/ (copied to profiles/)
mkdir -p gpurun_out
export PATH=/usr/local/cuda/bin:$PATH
K='point_pass_fp32 or device_compaction or degenerate or mask_bounding_boxes'
for tool in memcheck racecheck; do
  timeout 1200 compute-sanitizer --tool $tool --log-file gpurun_out/sanitizer_${tool}_smoke.log python __graft_entry__.py --smoke > gpurun_out/sanitizer_${tool}_smoke.out 2>&1
  echo "$tool smoke rc=$?"; tail -3 gpurun_out/sanitizer_${tool}_smoke.log
  timeout 2400 compute-sanitizer --tool $tool --log-file gpurun_out/sanitizer_${tool}_tests.log python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "$K" > gpurun_out/sanitizer_${tool}_tests.out 2>&1
  echo "$tool tests rc=$?"; tail -3 gpurun_out/sanitizer_${tool}_tests.log; tail -2 gpurun_out/sanitizer_${tool}_tests.out
done
