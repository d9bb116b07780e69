# The round's measurement pass: bench (both arms), ncu launch list, three --set full captures.
# Run on the GPU box: bash profiles/tools/measure_pass.sh ; results land in gpurun_out/off/.
set -x
mkdir -p gpurun_out/off
O=gpurun_out/off
python bench.py > $O/bench_n1.json 2> $O/bench_n1.err || exit 11
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err || exit 12
python bench.py --steps 20 --warmup 3 --no-cpu > $O/plain.json 2> $O/plain.err || exit 13
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 20 --warmup 3 --no-cpu > $O/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:match_count -s 5 -c 1 -o $O/match -f python bench.py --steps 20 --warmup 3 --no-cpu > $O/ncu_match.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:replay_identity -s 2 -c 1 -o $O/identity -f python bench.py --steps 20 --warmup 3 --no-cpu > $O/ncu_id.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:replay_sweep -s 2 -c 1 -o $O/sweep -f python bench.py --steps 20 --warmup 3 --no-cpu > $O/ncu_sweep.log 2>&1
cat $O/bench_n1.json
