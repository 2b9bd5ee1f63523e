set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_x_fullsize_gpu.py -q -m gpu -k "full_size" > gpurun_out/j_pytest_fullsize.log 2>&1; tail -12 gpurun_out/j_pytest_fullsize.log | cut -c1-400
timeout 100 python tools/fullsize_parity.py > gpurun_out/j_parity.log 2>&1; tail -14 gpurun_out/j_parity.log | cut -c1-200
timeout 150 python tools/ncu_all_kernels.py 1 > gpurun_out/j_ncu_plain.log 2>&1 && {
timeout 500 ncu --section SpeedOfLight --section LaunchStats --section Occupancy --section MemoryWorkloadAnalysis --section ComputeWorkloadAnalysis --clock-control none -o /tmp/prof_all_kernels -f python tools/ncu_all_kernels.py 1 > gpurun_out/j_ncu_all.log 2>&1
timeout 200 python tools/ncu_summary.py /tmp/prof_all_kernels.ncu-rep gpurun_out/ncu_summary_all_kernels.json > gpurun_out/j_ncu_summary.txt 2>&1
SZ=$(stat -c %s /tmp/prof_all_kernels.ncu-rep); echo "rep size $SZ"; if [ "$SZ" -lt 25000000 ]; then cp /tmp/prof_all_kernels.ncu-rep gpurun_out/; fi
timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_tcgen05 -c 3 -o gpurun_out/prof_gemm_top -f python tools/ncu_all_kernels.py 1 > gpurun_out/j_ncu_gemm.log 2>&1
}
tail -3 gpurun_out/j_ncu_all.log | cut -c1-200; head -30 gpurun_out/j_ncu_summary.txt | cut -c1-160; ls -la gpurun_out/*.ncu-rep; du -sh gpurun_out
