cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
run() { env "$@" timeout 200 python bench.py --steps 200 --warmup 20 --no-amp-arm > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err; python -c "
import json,sys;d=json.load(open('gpurun_out/z_bench.json'));print(sys.argv[1:], round(d['ms_per_step'],4), round(d['value'],1))" "$@"; }
run X=0
run DCAD_WGRAD_CTAS=128
run DCAD_WGRAD_CTAS=112
run DCAD_PLAN=lin2_fwd=256/4
run DCAD_PLAN=lin1_dgrad=256/4
run DCAD_QKV_DGRAD_SPLIT=2
