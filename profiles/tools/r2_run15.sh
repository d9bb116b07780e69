set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2_t15.log 2>&1; tail -3 gpurun_out/r2_t15.log
python profiles/tools/time_grouped.py > gpurun_out/r2_grouped15.log 2>&1
KAS_TRACE_DTYPE=uint16 python profiles/tools/time_grouped.py >> gpurun_out/r2_grouped15.log 2>&1
cat gpurun_out/r2_grouped15.log
KAS_BENCH_KEYS=20000000 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:pack_heads -c 3 --csv --log-file gpurun_out/r2_pack_traffic.csv python bench.py --config c5 --no-replay --no-cpu --no-extras --steps 3 --warmup 3 > gpurun_out/r2_pack_ncu.log 2>&1
tail -8 gpurun_out/r2_pack_traffic.csv
python -c "
import json
" 
