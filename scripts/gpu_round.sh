# Round-end GPU pass: parity tests, smoke, bench (both arms), next-row timings, ncu launch list and full capture
# of the two hot kernels, latency of the real frames.  Outputs under gpurun_out/.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference > gpurun_out/bench_reference.json 2>> gpurun_out/bench.err; echo "ref rc=$?"
timeout 600 python bench.py --workload cfg5 --steps 20 --no-e2e --no-cpu-baseline > gpurun_out/bench_cfg5.json 2>> gpurun_out/bench.err; echo "cfg5 rc=$?"
timeout 900 python bench_next.py > gpurun_out/bench_next.jsonl 2>> gpurun_out/bench.err; echo "next rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --no-packed --no-cfg4 --no-cfg5 > gpurun_out/ncu_l.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'lf_erode_fast_kernel|lf_fuse_kernel' -s 4 -c 2 -o gpurun_out/full -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --no-packed --no-cfg4 --no-cfg5 > gpurun_out/ncu_f.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'lf_fuse_kernel' -s 2 -c 1 -o gpurun_out/full_cfg5 -f python bench.py --workload cfg5 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-parity --no-packed > gpurun_out/ncu_5.log 2>&1
timeout 300 python scripts/latency_cfg12.py > gpurun_out/latency_cfg12.json 2>> gpurun_out/bench.err; echo "latency rc=$?"
ls -la gpurun_out
