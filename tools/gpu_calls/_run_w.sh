set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
PROBE_ROWS=0 timeout 200 ncu --set full --import-source on --clock-control none -k regex:adam_flat --launch-skip 3 -c 1 -o /tmp/adam0 python tools/adam_region2_probe.py > gpurun_out/w_ncu.log 2>&1
tail -3 gpurun_out/w_ncu.log
ncu -i /tmp/adam0.ncu-rep --page raw --csv > gpurun_out/w_adam0_raw.csv 2>/dev/null
ncu -i /tmp/adam0.ncu-rep --page source --csv > gpurun_out/w_adam0_source.csv 2>/dev/null
ls -la /tmp/adam0.ncu-rep gpurun_out/
