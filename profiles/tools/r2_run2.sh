python -m pytest tests/test_gpu_multirank.py tests/test_gpu_replay.py -m gpu -x -q -k "lean or comm or offsets" > gpurun_out/r2_t2.log 2>&1; tail -3 gpurun_out/r2_t2.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo bench rc=$?; tail -5 gpurun_out/r2_bench_n1.err; head -c 600 gpurun_out/r2_bench_n1.json
for v in "" gs_s8 gs_s16 gs_s24 gs_t256s16 gs_t256s24 gs_t64s8; do
  if [ -n "$v" ]; then export KAS_LIB=$PWD/build/variants/libkas_$v.so; fi
  echo "== $v" >> gpurun_out/r2_grouped2.log
  python profiles/tools/time_grouped.py >> gpurun_out/r2_grouped2.log 2>&1
  KAS_TRACE_DTYPE=uint16 python profiles/tools/time_grouped.py >> gpurun_out/r2_grouped2.log 2>&1
done
unset KAS_LIB
cat gpurun_out/r2_grouped2.log
ncu --set full --clock-control none --import-source on -k regex:replay_grouped_sorted -s 2 -c 1 -o gpurun_out/r2_gs -f python profiles/tools/time_grouped.py > gpurun_out/r2_ncu_gs.log 2>&1; tail -2 gpurun_out/r2_ncu_gs.log
