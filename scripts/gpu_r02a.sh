# r02 session: parity tests on the point pass, variant sweep (libraries from scripts/build_variants.sh),
# ncu capture of the point pass.  usage: bash scripts/gpu_r02a.sh [variant names...]
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/pytest_gpu.log
bash scripts/run_variants.sh "$@"
for v in "$@"; do
  LIDARFUSE_LIB=$PWD/build/variants/liblidarfuse_$v.so timeout 300 python bench.py --workload cfg5 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --no-packed > gpurun_out/var5_$v.json 2> gpurun_out/var5_$v.err || echo "$v cfg5 failed"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/var5_$v.json').read().strip().splitlines()[-1]); print('cfg5 %-10s step %.3f'%('$v',d['ms_per_step']), d['kernel_ms'])" || tail -5 gpurun_out/var5_$v.err
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'lf_fuse_kernel' -s 3 -c 1 -o gpurun_out/r02a_fuse -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --no-packed > gpurun_out/ncu_f.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_f.log | cut -c1-300
