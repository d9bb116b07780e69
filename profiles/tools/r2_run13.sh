set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r2_t13.log 2>&1; tail -15 gpurun_out/r2_t13.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; tail -2 gpurun_out/r2_smoke.log
