set -x
python -m pytest tests/test_gpu_glob.py tests/test_gpu_autoscaler.py -m gpu -x -q > gpurun_out/r2_t9.log 2>&1; tail -2 gpurun_out/r2_t9.log
for rs in 0 1 2; do
KAS_BENCH_RESERVED_SMS=$rs timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$rs bench.py --gpus 2 --steps 50 --warmup 5 --no-replay > gpurun_out/r2_bench_n2_rs$rs.json 2> gpurun_out/r2_bench_n2_rs$rs.err; echo rs$rs rc=$?; python -c "
import json; d=json.load(open('gpurun_out/r2_bench_n2_rs$rs.json')); print('reserved', $rs, d['ms_per_step'], d['roofline']['kernel_ms'], d['value'])"
done
python - <<'PY'
import sys
sys.path[:0]=['.','kiosk-autoscaler_b200']
import bench, torch
from autoscaler import engine
cx=type('C',(),{})()
PY
timeout 600 python bench.py --steps 5 --warmup 3 --no-replay --no-cpu > gpurun_out/r2_bench_extras.json 2> gpurun_out/r2_bench_extras.err; python -c "
import json; d=json.load(open('gpurun_out/r2_bench_extras.json')); print('glob', d['extras']['glob'])"
