# The round's measurement pass on ONE GPU: GPU tests, bench (both arms, both metrics), ncu launch list, one
# --set full capture.  Run on the GPU box: bash profiles/tools/measure_pass.sh ; results land in gpurun_out/.
# (One .ncu-rep per pass and gzip it: gpurun returns at most 64 MiB.)  Multi-GPU: r2_multi2.sh / r2_multi8.sh.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/gputest.log 2>&1; tail -3 gpurun_out/gputest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err || exit 11
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_n1_reference.json 2> gpurun_out/bench_ref.err || exit 12
timeout 600 python bench.py --metric decisions --steps 20 --warmup 5 --no-extras > gpurun_out/bench_n1_decisions.json 2> gpurun_out/bench_dec.err || exit 13
timeout 600 python bench.py --impl reference --metric decisions --steps 20 --warmup 5 > gpurun_out/bench_n1_decisions_reference.json 2>> gpurun_out/bench_ref.err || exit 14
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 20 --warmup 5 --no-cpu --no-extras > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:${KAS_NCU_KERNEL:-replay_grouped_sorted} -s 2 -c 1 -o gpurun_out/capture -f python profiles/tools/${KAS_NCU_SCRIPT:-time_grouped.py} > gpurun_out/ncu_capture.log 2>&1
gzip -9 gpurun_out/capture.ncu-rep
