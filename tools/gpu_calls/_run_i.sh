set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/i_bench_n1.json 2> gpurun_out/i_bench_n1.err; cut -c1-220 gpurun_out/i_bench_n1.json
timeout 900 python -m pytest tests/ -x -q -m gpu > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/i_pytest.log; tail -5 gpurun_out/i_pytest.log | cut -c1-200
timeout 200 python tools/ncu_all_kernels.py 1 > gpurun_out/i_ncu_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -o gpurun_out/prof_all_kernels -f python tools/ncu_all_kernels.py 1 > gpurun_out/i_ncu.log 2>&1
tail -2 gpurun_out/i_ncu_plain.log; tail -3 gpurun_out/i_ncu.log | cut -c1-200; ls -la gpurun_out/*.ncu-rep
