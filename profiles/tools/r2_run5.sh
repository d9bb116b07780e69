python -m pytest tests/test_gpu_replay.py -m gpu -x -q -k "lean or offsets or uint16_trace" > gpurun_out/r2_t5.log 2>&1; tail -2 gpurun_out/r2_t5.log
rm -f gpurun_out/r2_grouped5.log
for v in "" gs_t256s32st2 gs_t512s48st2 gs_t512s64st2 gs_t512s32st3 gs_t1024s96st2 gs_t512s24st4; do
  if [ -n "$v" ]; then export KAS_LIB=$PWD/build/variants/libkas_$v.so; fi
  echo "== $v" >> gpurun_out/r2_grouped5.log
  python profiles/tools/time_grouped.py >> gpurun_out/r2_grouped5.log 2>&1
  KAS_TRACE_DTYPE=uint16 python profiles/tools/time_grouped.py >> gpurun_out/r2_grouped5.log 2>&1
done
cat gpurun_out/r2_grouped5.log
export KAS_LIB=$PWD/build/variants/libkas_gs_t512s48st2.so
KAS_TRACE_DTYPE=uint16 ncu --set full --clock-control none --import-source on -k regex:replay_grouped_sorted -s 2 -c 1 -o gpurun_out/r2_gs512u16 -f python profiles/tools/time_grouped.py > gpurun_out/r2_ncu_gs512u16.log 2>&1; tail -2 gpurun_out/r2_ncu_gs512u16.log
