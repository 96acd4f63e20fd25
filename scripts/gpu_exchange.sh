# N-GPU bench with both counter exchanges (peer stores vs NCCL all-reduce), device numbers only.
N=${1:-2}
mkdir -p gpurun_out
for x in peer nccl; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 \
    bench.py --gpus $N --steps 50 --warmup 5 --exchange $x --no-e2e --no-packed > gpurun_out/bench_n${N}_$x.json 2> gpurun_out/bench_n${N}_$x.err; echo "$x rc=$?"
  tail -2 gpurun_out/bench_n${N}_$x.err
  python -c "
import json,sys
d=json.loads(open('gpurun_out/bench_n${N}_$x.json').read().strip().splitlines()[-1])
print('$x', d['n_gpus'], 'ms/step', round(d['ms_per_step'],4), 'value %.4g'%d['value'], d['kernel_ms'], d['config'].get('exchange'))"
done
