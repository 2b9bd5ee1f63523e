cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 120 python tools/ncu_all_kernels.py 1 > gpurun_out/attn_plain.log 2>&1 || { tail -5 gpurun_out/attn_plain.log; exit 1; }
timeout 300 ncu --set full --import-source on --clock-control none -k 'regex:attn_(fwd|bwd)_kernel' --launch-skip 5 -c 2 -o /tmp/attn python tools/ncu_all_kernels.py 1 > gpurun_out/attn_ncu.log 2>&1
tail -2 gpurun_out/attn_ncu.log | cut -c1-200
ncu -i /tmp/attn.ncu-rep --page raw --csv > gpurun_out/attn_raw.csv 2>/dev/null
ncu -i /tmp/attn.ncu-rep --page source --csv > gpurun_out/attn_source.csv 2>/dev/null
ls -la /tmp/attn.ncu-rep gpurun_out/attn_*
