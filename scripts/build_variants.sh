# Kernel experiments: objects of every source once, then one library per set of -D flags for the
# point-pass sources.  usage: bash scripts/build_variants.sh name1 "-DLF_FUSE_CLAIM=1024" name2 "..."
# Libraries land in build/variants/liblidarfuse_<name>.so; run with LIDARFUSE_LIB=<path>.
set -e
cd "$(dirname "$0")/.."
C=lidar_object_detection_b200/csrc; O=build/variants; mkdir -p $O
FLAGS="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-O3 -I include -I $C -DLF_BUILDING=1"
for f in ctx project erode boxes2d resize colour csvfmt packhost points compact synth publish; do
  if [ ! -f $O/$f.o ] || [ $C/$f.cu -nt $O/$f.o ]; then nvcc $FLAGS -c $C/$f.cu -o $O/$f.o & fi
done
wait
while [ $# -ge 2 ]; do
  name=$1; defs=$2; shift 2
  nvcc $FLAGS $defs -c $C/fuse.cu -o $O/fuse_$name.o &
  wait
  nvcc -shared -o $O/liblidarfuse_$name.so $O/fuse_$name.o $O/ctx.o $O/project.o $O/erode.o $O/boxes2d.o \
    $O/resize.o $O/colour.o $O/csvfmt.o $O/packhost.o $O/points.o $O/compact.o $O/synth.o $O/publish.o
  echo built $name "($defs)"
done
