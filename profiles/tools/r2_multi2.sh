set -x
nvidia-smi topo -m > gpurun_out/r2_topo_n2.txt 2>&1
python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/r2_multirank_n2.log 2>&1; tail -5 gpurun_out/r2_multirank_n2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo n2 rc=$?; tail -5 gpurun_out/r2_bench_n2.err; head -c 400 gpurun_out/r2_bench_n2.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --steps 20 --warmup 5 --config c5 > gpurun_out/r2_bench_c5_n2.json 2> gpurun_out/r2_bench_c5_n2.err; echo c5n2 rc=$?; tail -5 gpurun_out/r2_bench_c5_n2.err
