# Scaling check on N GPUs of one box (N from $1, default 8): host facts, then bench.py under torchrun.
N=${1:-8}
mkdir -p gpurun_out
nproc; free -g | head -2; nvidia-smi -L | wc -l
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$?"
tail -5 gpurun_out/bench_n$N.err; free -g | head -2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
  bench.py --impl reference --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n${N}_reference.json 2>> gpurun_out/bench_n$N.err; echo "ref rc=$?"
