# packed-masks-in check: parity tests of the new path, bench with the secondary measurement, next-row timings
mkdir -p gpurun_out
nproc; grep -m1 "model name" /proc/cpuinfo
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --e2e-steps 2 > gpurun_out/bench_packed.json 2> gpurun_out/bench_packed.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d = json.load(open('gpurun_out/bench_packed.json'))
    print('ms/step', d['ms_per_step'], d['kernel_ms'], 'e2e %.4g' % d['e2e']['value'], d['e2e']['ms_per_step'])
    print(json.dumps(d['packed_masks_in'], indent=1))
except Exception as e:
    print('no bench line', e); print(open('gpurun_out/bench_packed.err').read()[-3000:])
PY
timeout 600 python bench_next.py > gpurun_out/bench_next.jsonl 2> gpurun_out/bench_next.err; echo "next rc=$?"; grep -h "erode_pack" gpurun_out/bench_next.jsonl
