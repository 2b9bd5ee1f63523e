cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 400 python -m pytest tests -x -q -m gpu > gpurun_out/final1_pytest.log 2>&1; tail -3 gpurun_out/final1_pytest.log | cut -c1-300
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" > gpurun_out/final1_smoke.log 2>&1; tail -2 gpurun_out/final1_smoke.log | cut -c1-300
timeout 300 python bench.py > gpurun_out/final1_bench.json 2> gpurun_out/final1_bench.err; cut -c1-300 gpurun_out/final1_bench.json
