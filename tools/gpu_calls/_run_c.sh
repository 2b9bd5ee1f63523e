set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_x_gemm_epilogue_gpu.py tests/test_x_fullsize_gpu.py -q -m gpu > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c_pytest.log
tail -8 gpurun_out/c_pytest.log
TIMELINE_MODES=0,1,2 timeout 300 python tools/gemm_timeline.py > gpurun_out/c_timeline.log 2>&1; cat gpurun_out/c_timeline.log
