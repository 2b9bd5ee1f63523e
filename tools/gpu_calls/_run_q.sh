set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -x -q -m gpu > gpurun_out/q_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/q_pytest.log; tail -6 gpurun_out/q_pytest.log | cut -c1-300
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/q_bench_n1.json 2> gpurun_out/q_bench_n1.err; cut -c1-200 gpurun_out/q_bench_n1.json
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 120 python tools/timeline.py > gpurun_out/q_timeline.log 2>&1; head -2 gpurun_out/timeline.txt
