set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2_t10.log 2>&1; tail -3 gpurun_out/r2_t10.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo bench rc=$?; tail -3 gpurun_out/r2_bench_n1.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n1.json')); x=d['extras']; print(x['one_resource_c1']); print(x['glob']['ms']); print(d['value'], d['replay']['sweep']['ms'], d['replay']['sweep_uint16']['ms'])"
python profiles/tools/time_c1.py > gpurun_out/r2_c1.log 2>&1; cat gpurun_out/r2_c1.log
