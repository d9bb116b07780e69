# 8-GPU pass: the multi-rank parity test (2 / 4 / 8 ranks), then the bench at 8 GPUs for configs 2, 4 and 5.
set -x
nvidia-smi topo -m > gpurun_out/r2_topo_n8.txt 2>&1
python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/r2_multirank_n8.log 2>&1; tail -5 gpurun_out/r2_multirank_n8.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err; echo n8 rc=$?; tail -3 gpurun_out/r2_bench_n8.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 5 --config c4 > gpurun_out/r2_bench_c4_n8.json 2> gpurun_out/r2_bench_c4_n8.err; echo c4 rc=$?; tail -3 gpurun_out/r2_bench_c4_n8.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --steps 20 --warmup 5 --config c5 > gpurun_out/r2_bench_c5_n8.json 2> gpurun_out/r2_bench_c5_n8.err; echo c5 rc=$?; tail -3 gpurun_out/r2_bench_c5_n8.err
