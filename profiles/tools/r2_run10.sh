set -x
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo bench rc=$?; tail -3 gpurun_out/r2_bench_n1.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n1.json')); x=d['extras']; print(x['startup_delay_3_ticks']); print(d['replay']['e2e_uint16'])"
ncu --set full --clock-control none --import-source on -k regex:replay_sweep -s 2 -c 1 -o gpurun_out/r2_sweep -f python profiles/tools/time_k2.py > gpurun_out/r2_ncu_sweep.log 2>&1; tail -2 gpurun_out/r2_ncu_sweep.log
gzip -9 gpurun_out/r2_sweep.ncu-rep
