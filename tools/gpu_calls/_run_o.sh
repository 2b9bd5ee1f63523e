set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 200 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/o_bench_trim.json 2> gpurun_out/o_bench_trim.err; cut -c1-200 gpurun_out/o_bench_trim.json; tail -3 gpurun_out/o_bench_trim.err
DCAD_LIN1_FWD_BN=256 DCAD_LIN2_DGRAD_BN=256 DCAD_LIN2_DGRAD_CTAS=0 timeout 200 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/o_bench_trim_256.json 2> gpurun_out/o_bench_trim_256.err; cut -c1-200 gpurun_out/o_bench_trim_256.json
timeout 200 python bench.py --steps 100 --warmup 10 --no-amp-arm --no-trim-padding > gpurun_out/o_bench_notrim.json 2> gpurun_out/o_bench_notrim.err; cut -c1-200 gpurun_out/o_bench_notrim.json
TOKENS=1536 timeout 300 python tools/tune_gemm.py > gpurun_out/o_tune_T1536.log 2>&1; tail -14 gpurun_out/o_tune_T1536.log | cut -c1-330
