# parity tests, then every variant with and without the prefilter (cfg3 and cfg5), then ncu of the default library
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
bash scripts/run_variants.sh "$@"

for extra in ""; do for v in "$@"; do
  LIDARFUSE_LIB=$PWD/build/variants/liblidarfuse_$v.so timeout 300 python bench.py --workload cfg5 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --no-packed $extra > gpurun_out/var5_$v.json 2> gpurun_out/var5_$v.err || echo "$v cfg5 failed"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/var5_$v.json').read().strip().splitlines()[-1]); print('cfg5 %-10s %s step %.3f'%('$v','$extra',d['ms_per_step']), d['kernel_ms'])" || tail -5 gpurun_out/var5_$v.err
done; done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'lf_fuse_kernel' -s 3 -c 1 -o gpurun_out/r02b_fuse -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --no-packed > gpurun_out/ncu_f.log 2>&1
echo "ncu rc=$?"
