python -m pytest tests -m gpu -x -q > gpurun_out/r2_t6.log 2>&1; tail -3 gpurun_out/r2_t6.log
python profiles/tools/time_grouped.py > gpurun_out/r2_grouped6.log 2>&1
KAS_TRACE_DTYPE=uint16 python profiles/tools/time_grouped.py >> gpurun_out/r2_grouped6.log 2>&1
cat gpurun_out/r2_grouped6.log
timeout 900 python bench.py --config c5 --steps 20 --warmup 3 --no-extras > gpurun_out/r2_bench_c5_n1.json 2> gpurun_out/r2_bench_c5_n1.err; echo c5 rc=$?; tail -3 gpurun_out/r2_bench_c5_n1.err
timeout 900 python bench.py --config c4 --steps 20 --warmup 3 --no-extras > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err; echo c4 rc=$?; tail -3 gpurun_out/r2_bench_c4_n1.err
# the sweep kernel's new packed path, as a variant library
KAS_LIB=$PWD/build/variants/libkas_sweepnext.so python -m pytest tests/test_gpu_replay.py -m gpu -x -q -k "sweep or uint16 or random_fleets or large_replay" > gpurun_out/r2_t6_sweepnext.log 2>&1; tail -2 gpurun_out/r2_t6_sweepnext.log
for v in "" sweepnext sweepnextfma; do
  if [ -n "$v" ]; then export KAS_LIB=$PWD/build/variants/libkas_$v.so; fi
  echo "== $v" >> gpurun_out/r2_k2_6.log
  python profiles/tools/time_k2.py >> gpurun_out/r2_k2_6.log 2>&1
  KAS_TRACE_DTYPE=uint16 python profiles/tools/time_k2.py >> gpurun_out/r2_k2_6.log 2>&1
done
unset KAS_LIB
cat gpurun_out/r2_k2_6.log
