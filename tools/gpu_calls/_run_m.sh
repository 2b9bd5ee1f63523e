set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -x -q -m gpu > gpurun_out/m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/m_pytest.log; tail -6 gpurun_out/m_pytest.log | cut -c1-300
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/m_bench_n1.json 2> gpurun_out/m_bench_n1.err; cut -c1-200 gpurun_out/m_bench_n1.json
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/m_smoke.log 2>&1; tail -2 gpurun_out/m_smoke.log
timeout 300 python bench.py --impl reference --steps 50 --warmup 5 > gpurun_out/m_bench_ref_n1.json 2> gpurun_out/m_bench_ref.err; cut -c1-200 gpurun_out/m_bench_ref_n1.json
