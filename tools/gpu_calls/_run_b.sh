set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_x_gemm_epilogue_gpu.py tests/test_x_fullsize_gpu.py tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -25 gpurun_out/b_pytest.log
timeout 300 python tools/fullsize_parity.py > gpurun_out/b_parity.log 2>&1; tail -16 gpurun_out/b_parity.log
GEMM_EPI=2 timeout 600 python tools/tune_gemm.py > gpurun_out/b_tune_epi2.log 2>&1; tail -14 gpurun_out/b_tune_epi2.log
DCAD_GEMM_EPI=2 timeout 300 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/b_bench_epi2.json 2> gpurun_out/b_bench_epi2.err; cut -c1-300 gpurun_out/b_bench_epi2.json
