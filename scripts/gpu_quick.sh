# quick GPU check: parity tests + short bench (no CPU baseline); extra bench args: $@
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
for mode in "" "--workload cfg5 --steps 10"; do
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --e2e-steps 1 $mode "$@" > gpurun_out/bench_quick${mode:0:3}.json 2> gpurun_out/bench_quick${mode:0:3}.err; echo "bench $mode rc=$?"
python - "$mode" <<'PY'
import json, sys
m = sys.argv[1]
try:
    d=json.load(open('gpurun_out/bench_quick%s.json' % m[:3]))
    print(m, 'ms/step', d['ms_per_step'], 'value %.4g'%d['value'], d['kernel_ms'], 'roofline frac', d['roofline']['frac'], 'e2e %.4g' % d['e2e']['value'])
except Exception as e:
    print('no bench line', e); print(open('gpurun_out/bench_quick%s.err' % m[:3]).read()[-2000:])
PY
done
