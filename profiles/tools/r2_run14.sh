set -x
python -m pytest tests/test_gpu_replay.py -m gpu -x -q -k "sweep or uint16 or random_fleets or large_replay" > gpurun_out/r2_t14.log 2>&1; tail -2 gpurun_out/r2_t14.log
rm -f gpurun_out/r2_k2_14.log
for v in "" nokeepall; do
  if [ -n "$v" ]; then export KAS_LIB=$PWD/build/variants/libkas_$v.so; fi
  echo "== $v" >> gpurun_out/r2_k2_14.log
  python profiles/tools/time_k2.py >> gpurun_out/r2_k2_14.log 2>&1
  KAS_TRACE_DTYPE=uint16 python profiles/tools/time_k2.py >> gpurun_out/r2_k2_14.log 2>&1
done
cat gpurun_out/r2_k2_14.log
