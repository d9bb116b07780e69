set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2_t12.log 2>&1; tail -3 gpurun_out/r2_t12.log
python profiles/tools/time_k2.py > gpurun_out/r2_k2_12.log 2>&1
KAS_TRACE_DTYPE=uint16 python profiles/tools/time_k2.py >> gpurun_out/r2_k2_12.log 2>&1
cat gpurun_out/r2_k2_12.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-replay --no-cpu > gpurun_out/r2_bench_x.json 2> gpurun_out/r2_bench_x.err; echo rc=$?
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_x.json')); x=d['extras']['startup_delay_3_ticks']; print({k:(v['ms'], v['roofline']['frac']) for k,v in x.items() if isinstance(v,dict)}); print(d['extras']['patch_events'])"
