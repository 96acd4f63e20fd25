# usage: bash scripts/gpu_ncu_next.sh <out-prefix> kernel1 kernel2 ...   -- one ncu --set full capture (4th launch) per kernel of bench_next.py
mkdir -p gpurun_out
O="$1"; shift
for K in "$@"; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 3 -c 1 -o gpurun_out/${O}_$K -f python bench_next.py --reps 1 > gpurun_out/${O}_$K.log 2>&1
  echo "$K ncu rc=$?"
done
ls -la gpurun_out | head -30
