set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L
timeout 1200 python -m pytest tests/ -x -q -m gpu > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -15 gpurun_out/a_pytest.log
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/a_bench_epi0.json 2> gpurun_out/a_bench_epi0.err; cut -c1-400 gpurun_out/a_bench_epi0.json
DCAD_GEMM_EPI=1 timeout 300 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/a_bench_epi1.json 2> gpurun_out/a_bench_epi1.err; cut -c1-300 gpurun_out/a_bench_epi1.json
DCAD_GEMM_EPI=1 DCAD_GEMM_WIDE=1 DCAD_LIN2_DGRAD_CTAS=0 timeout 300 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/a_bench_epi1_wide.json 2> gpurun_out/a_bench_epi1_wide.err; cut -c1-300 gpurun_out/a_bench_epi1_wide.json
timeout 600 python tools/tune_gemm.py > gpurun_out/a_tune_epi0.log 2>&1; tail -14 gpurun_out/a_tune_epi0.log
GEMM_EPI=1 timeout 600 python tools/tune_gemm.py > gpurun_out/a_tune_epi1.log 2>&1; tail -14 gpurun_out/a_tune_epi1.log
