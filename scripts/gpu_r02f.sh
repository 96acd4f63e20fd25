# point pass, 64-point phase B: bench lines (cfg3 and cfg5) and latency per variant
mkdir -p gpurun_out
bash scripts/run_variants.sh base g2 g2s
BENCH_EXTRA="--workload cfg5 --steps 10" bash scripts/run_variants.sh base g2 g2s
for v in base g2 g2s; do echo "== $v"; LIDARFUSE_LIB=$PWD/build/variants/liblidarfuse_$v.so timeout 300 python scripts/latency_cfg12.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['config'], 'direct %.4f ms' % d['direct']['device_ms_per_call'], 'graph %.4f ms' % d['graph']['device_ms_per_call'])"; done
