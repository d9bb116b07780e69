python -m pytest tests -m gpu -x -q > gpurun_out/r2_t4.log 2>&1; tail -3 gpurun_out/r2_t4.log
rm -f gpurun_out/r2_grouped4.log
for v in "" gs_t256s24 gs_t256s32st2 gs_t256s16st2 gs_t512s36st2 gs_t512s48st2 gs_t256s32st2u gs_t512s48st2u; do
  if [ -n "$v" ]; then export KAS_LIB=$PWD/build/variants/libkas_$v.so; fi
  echo "== $v" >> gpurun_out/r2_grouped4.log
  python profiles/tools/time_grouped.py >> gpurun_out/r2_grouped4.log 2>&1
  KAS_TRACE_DTYPE=uint16 python profiles/tools/time_grouped.py >> gpurun_out/r2_grouped4.log 2>&1
done
cat gpurun_out/r2_grouped4.log
export KAS_LIB=$PWD/build/variants/libkas_gs_t256s32st2.so
ncu --set full --clock-control none --import-source on -k regex:replay_grouped_sorted -s 2 -c 1 -o gpurun_out/r2_gs256 -f python profiles/tools/time_grouped.py > gpurun_out/r2_ncu_gs256.log 2>&1; tail -2 gpurun_out/r2_ncu_gs256.log
