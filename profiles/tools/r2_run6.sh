python -m pytest tests -m gpu -x -q > gpurun_out/r2_t6.log 2>&1; tail -3 gpurun_out/r2_t6.log
python profiles/tools/time_grouped.py > gpurun_out/r2_grouped6.log 2>&1
KAS_TRACE_DTYPE=uint16 python profiles/tools/time_grouped.py >> gpurun_out/r2_grouped6.log 2>&1
cat gpurun_out/r2_grouped6.log
timeout 900 python bench.py --config c5 --steps 20 --warmup 3 --no-extras > gpurun_out/r2_bench_c5_n1.json 2> gpurun_out/r2_bench_c5_n1.err; echo c5 rc=$?; tail -3 gpurun_out/r2_bench_c5_n1.err
timeout 900 python bench.py --config c4 --steps 20 --warmup 3 --no-extras > gpurun_out/r2_bench_c4_n1.json 2> gpurun_out/r2_bench_c4_n1.err; echo c4 rc=$?; tail -3 gpurun_out/r2_bench_c4_n1.err
