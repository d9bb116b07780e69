#!/bin/bash
# Build an experimental libkas with extra -D flags next to the product one:
#   profiles/tools/build_variant.sh <name> [-DKAS_...=...]   ->  build/variants/libkas_<name>.so
# Timed against each other with KAS_LIB=<.so> python profiles/tools/time_k1.py / time_k2.py.
set -e
cd "$(dirname "$0")/../.."
name=$1; shift
out=build/variants; mkdir -p $out/obj_$name
objs=""
for s in kas_match kas_replay kas_events kas_glob kas_ingest kas_comm kas_rdb; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
    -Xcompiler -fPIC,-fvisibility=hidden,-Wall --expt-relaxed-constexpr "$@" \
    -c kiosk-autoscaler_b200/csrc/$s.cu -o $out/obj_$name/$s.o &
  objs="$objs $out/obj_$name/$s.o"
done
wait
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $out/libkas_$name.so $objs -ldl
echo $out/libkas_$name.so
