#!/bin/bash
# Like build_variant.sh, for flags that only concern kas_replay.cu: compiles that one file with the
# extra -D flags and links it with the product build's other objects (build the product first):
#   profiles/tools/build_variant_replay.sh <name> [-DKAS_...=...]  ->  build/variants/libkas_<name>.so
set -e
cd "$(dirname "$0")/../.."
name=$1; shift
out=build/variants; mkdir -p $out/obj_$name
lib=kiosk-autoscaler_b200/lib
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
  -Xcompiler -fPIC,-fvisibility=hidden,-Wall --expt-relaxed-constexpr "$@" \
  -c kiosk-autoscaler_b200/csrc/kas_replay.cu -o $out/obj_$name/kas_replay.o
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $out/libkas_$name.so \
  $out/obj_$name/kas_replay.o $lib/kas_match.o $lib/kas_events.o $lib/kas_glob.o $lib/kas_ingest.o $lib/kas_comm.o $lib/kas_rdb.o -ldl
echo $out/libkas_$name.so
