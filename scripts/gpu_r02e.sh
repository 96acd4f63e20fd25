# stand-alone kernels reworked: parity tests, then the next-row bench lines
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench_next.py > gpurun_out/bench_next.jsonl 2> gpurun_out/bench_next.err; echo "bench_next rc=$?"
tail -3 gpurun_out/bench_next.err
python - <<'PY'
import json
for l in open('gpurun_out/bench_next.jsonl'):
    if l.startswith('{'):
        d = json.loads(l)
        print('%-60s %8.4f ms  frac %s  %s' % (d['kernel'][:60], d['ms'], round(d.get('frac', 0), 3), d.get('note', '')[:60]))
PY
