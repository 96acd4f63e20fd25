for v in "$@"; do echo "== $v"; LIDARFUSE_LIB=$PWD/build/variants/liblidarfuse_$v.so python scripts/latency_cfg12.py | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['config'], 'direct %.4f graph %.4f ms' % (d['direct']['device_ms_per_call'], d['graph']['device_ms_per_call']))"; done
