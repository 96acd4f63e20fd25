# usage: bash scripts/gpu_ncu_next.sh <kernel-regex> <out-name> [count]   -- ncu --set full of bench_next.py kernels
mkdir -p gpurun_out
K="$1"; O="$2"; C="${3:-8}"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$K" -c $C -o gpurun_out/$O -f python bench_next.py --reps 1 > gpurun_out/$O.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/$O.log
