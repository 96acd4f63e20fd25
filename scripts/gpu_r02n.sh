# usage: bash scripts/gpu_r02n.sh N   -- the driver's multi-GPU commands (both arms) on N GPUs
N=$1
mkdir -p gpurun_out
( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err ) 2>&1 | grep real; echo "bench rc=$?"
tail -3 gpurun_out/bench_n$N.err
python - $N <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open('gpurun_out/bench_n%s.json' % n).read().strip().splitlines()[-1])
    print('N', n, 'value %.4g' % d['value'], 'ms/step', d['ms_per_step'], d['kernel_ms'], d['config']['exchange'][:40])
    e = d['e2e']
    print(' e2e', e['variant'], '%.4g' % e['value'], 'ms', round(e['ms_per_step'], 1), 'floor', e.get('host_dram_floor_ms'), 'pack', e.get('host_pack_gbs_alone'), e['host'])
    for k, v in e['other_variants'].items():
        print('   ', k, '%.4g' % v['value'], round(v['ms_per_step'], 1), v.get('host_pack_gbs_alone'))
    print(' cfg4', d['cfg4_job'] and {k: d['cfg4_job'][k] for k in ('job_ms', 'value', 'checks')})
    print(' cfg5', d['cfg5'] and {k: d['cfg5'][k] for k in ('ms_per_step', 'value', 'kernel_ms')})
except Exception as ex:
    print('no line', ex)
PY
