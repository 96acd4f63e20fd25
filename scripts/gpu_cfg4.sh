N=${1:-8}
mkdir -p gpurun_out
timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 scripts/cfg4_job.py > gpurun_out/cfg4_n$N.json 2> gpurun_out/cfg4_n$N.err; echo rc=$?
tail -3 gpurun_out/cfg4_n$N.err; cat gpurun_out/cfg4_n$N.json | cut -c1-400
