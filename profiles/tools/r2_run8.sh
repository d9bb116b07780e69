set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2_t8.log 2>&1; tail -3 gpurun_out/r2_t8.log
python profiles/tools/time_k2.py > gpurun_out/r2_k2_8.log 2>&1
KAS_TRACE_DTYPE=uint16 python profiles/tools/time_k2.py >> gpurun_out/r2_k2_8.log 2>&1
cat gpurun_out/r2_k2_8.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo bench rc=$?; tail -3 gpurun_out/r2_bench_n1.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench_n1_reference.json 2> gpurun_out/r2_bench_ref.err; echo ref rc=$?
timeout 600 python bench.py --metric decisions --steps 20 --warmup 5 --no-extras > gpurun_out/r2_bench_n1_decisions.json 2> gpurun_out/r2_bench_dec.err; echo dec rc=$?
timeout 600 python bench.py --impl reference --metric decisions --steps 20 --warmup 5 > gpurun_out/r2_bench_n1_decisions_reference.json 2>> gpurun_out/r2_bench_ref.err; echo decref rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 20 --warmup 5 --no-cpu --no-extras > gpurun_out/r2_ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:replay_grouped_sorted -s 2 -c 1 -o gpurun_out/r2_gsfinal -f python profiles/tools/time_grouped.py > gpurun_out/r2_ncu_gsfinal.log 2>&1; tail -2 gpurun_out/r2_ncu_gsfinal.log
gzip -9 gpurun_out/r2_gsfinal.ncu-rep
ls -la gpurun_out
