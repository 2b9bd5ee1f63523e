set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_x_fullsize_gpu.py tests/test_engine_gpu.py -q -m gpu > gpurun_out/e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e_pytest.log
tail -25 gpurun_out/e_pytest.log
timeout 200 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/e_bench_xhat1.json 2> gpurun_out/e_bench_xhat1.err; cut -c1-200 gpurun_out/e_bench_xhat1.json
DCAD_SAVE_XHAT=0 timeout 200 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/e_bench_xhat0.json 2> gpurun_out/e_bench_xhat0.err; cut -c1-200 gpurun_out/e_bench_xhat0.json
timeout 300 python baseline/socket_round_reference.py > gpurun_out/e_socket_round_b200_box.json 2> gpurun_out/e_socket_round.err; cat gpurun_out/e_socket_round_b200_box.json
timeout 200 python tools/ncu_all_kernels.py 1 > gpurun_out/e_ncu_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dcad -o gpurun_out/prof_all_kernels -f python tools/ncu_all_kernels.py 1 > gpurun_out/e_ncu.log 2>&1
tail -3 gpurun_out/e_ncu_plain.log; tail -3 gpurun_out/e_ncu.log; ls -la gpurun_out/*.ncu-rep
