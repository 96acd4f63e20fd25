# usage: bash scripts/run_variants.sh name1 name2 ...   (libraries from scripts/build_variants.sh)
mkdir -p gpurun_out
for v in "$@"; do
  LIDARFUSE_LIB=$PWD/build/variants/liblidarfuse_$v.so timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-parity --no-packed $BENCH_EXTRA > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err || echo "$v failed"
  python - "$v" <<'PY'
import json, sys
v = sys.argv[1]
try:
    d = json.loads(open('gpurun_out/var_%s.json' % v).read().strip().splitlines()[-1])
    print('%-12s step %.3f ms  %s' % (v, d['ms_per_step'], {k: round(x, 4) for k, x in d['kernel_ms'].items()}))
except Exception as e:
    print(v, 'no result', e)
PY
done
