mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "packed or erosion" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python bench_next.py > gpurun_out/bench_next.jsonl 2> gpurun_out/bench_next.err; echo "next rc=$?"; grep -h "erode_pack" gpurun_out/bench_next.jsonl
timeout 600 python bench.py --steps 20 --no-cpu-baseline --no-e2e > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; python -c "
import json; d=json.load(open('gpurun_out/bench_q.json')); print(d['ms_per_step'], d['kernel_ms'], d['packed_masks_in']['ms_per_step'], d['packed_masks_in']['kernel_ms'])"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'lf_erode_packed_kernel' -s 1 -c 1 -o gpurun_out/full_packed -f python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_p.log 2>&1; tail -2 gpurun_out/ncu_p.log
