N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 \
  bench.py --gpus $N --workload cfg5 --steps 20 --warmup 3 > gpurun_out/bench_cfg5_n$N.json 2> gpurun_out/bench_cfg5_n$N.err; echo rc=$?
tail -3 gpurun_out/bench_cfg5_n$N.err; cut -c1-300 gpurun_out/bench_cfg5_n$N.json
