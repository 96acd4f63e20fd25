mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-packed --no-cfg4 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); print('cfg3 step %.3f' % d['ms_per_step'], d['kernel_ms'], 'parity frames', d['parity_checked_frames']); print('cfg5', d['cfg5'] and d['cfg5']['kernel_ms'])"
timeout 300 python scripts/latency_cfg12.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(' ', d['config'], 'direct %.4f ms' % d['direct']['device_ms_per_call'], 'graph %.4f ms' % d['graph']['device_ms_per_call'])"
