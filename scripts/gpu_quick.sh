# quick GPU check: parity tests + short bench (no CPU baseline)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --e2e-steps 1 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench_quick.json'))
    print('ms/step', d['ms_per_step'], 'value %.4g'%d['value'], d['kernel_ms'], 'roofline frac', d['roofline']['frac'])
except Exception as e:
    print('no bench line', e); print(open('gpurun_out/bench_quick.err').read()[-2000:])
PY
