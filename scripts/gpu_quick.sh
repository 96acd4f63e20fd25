# quick GPU check: parity tests + next-row timings (+ short bench); extra bench args: $@
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench_next.py > gpurun_out/bench_next.jsonl 2> gpurun_out/bench_next.err; echo "bench_next rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/bench_next.jsonl'):
    if l.startswith('{'):
        d = json.loads(l)
        print('%-60s %8.4f ms  frac %s' % (d['kernel'][:60], d['ms'], round(d.get('frac', 0), 3)))
PY
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-cfg4 "$@" 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); print('cfg3 step %.3f' % d['ms_per_step'], d['kernel_ms'], 'parity', d['parity_checked_frames']); print('cfg5', d['cfg5'] and d['cfg5']['kernel_ms']); print('packed', d['packed_masks_in'] and d['packed_masks_in']['kernel_ms'])"
