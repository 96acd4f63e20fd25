# tests + the driver's two bench commands at N = 1
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/pytest_gpu.log
( time timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err ) 2>&1 | grep real; echo "bench rc=$?"
tail -5 gpurun_out/bench.err
( time timeout 900 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_ref.err ) 2>&1 | grep real
python - <<'PY'
import json
for n in ('bench', 'bench_reference'):
    try:
        d = json.loads(open('gpurun_out/%s.json' % n).read().strip().splitlines()[-1])
        print(n, 'value %.4g' % d['value'], 'ms/step', d.get('ms_per_step'), 'steps', d['steps'], 'warmup', d['warmup'])
        if n == 'bench':
            print(' kernel_ms', d['kernel_ms'], 'roofline', round(d['roofline']['frac'], 4), 'launches', d['gpu_launches'], 'parity frames', d['parity_checked_frames'])
            e = d['e2e']
            print(' e2e', e['variant'], '%.4g' % e['value'], 'ms', round(e['ms_per_step'], 1), 'floor ms', e.get('host_dram_floor_ms'), 'pack GB/s', e.get('host_pack_gbs_alone'), e['host'])
            for k, v in e['other_variants'].items():
                print('   ', k, '%.4g' % v['value'], round(v['ms_per_step'], 1), v.get('host_pack_gbs_alone'))
            print(' cfg4', d['cfg4_job'] and {k: d['cfg4_job'][k] for k in ('job_ms', 'value', 'checks')})
            print(' cfg5', d['cfg5'] and {k: d['cfg5'][k] for k in ('ms_per_step', 'value', 'kernel_ms', 'step_frac_of_measured_peak')})
            print(' cpu', d['cpu_baseline'])
        else:
            print(' ', d['cpu_baseline'])
    except Exception as ex:
        print(n, 'no line', ex)
PY
