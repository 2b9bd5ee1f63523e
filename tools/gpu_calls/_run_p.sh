set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" timeout 120 python bench.py --steps 100 --warmup 10 --no-amp-arm > gpurun_out/p_$name.json 2> gpurun_out/p_$name.err; python -c "
import json; d=json.load(open('gpurun_out/p_$name.json')); print('$name', d['value'], d['ms_per_step'])" || tail -3 gpurun_out/p_$name.err; }
run base X=1
run qkv192 DCAD_PLAN="qkv_fwd=192"
run split4 DCAD_PLAN="lin2_fwd=256/4,lin1_dgrad=256/4"
run out64 DCAD_PLAN="out_fwd=64,out_dgrad=64"
run all DCAD_PLAN="qkv_fwd=192,lin2_fwd=256/4,lin1_dgrad=256/4,out_fwd=64,out_dgrad=64"
run all_ctas128 DCAD_PLAN="qkv_fwd=192,lin2_fwd=256/4,lin1_dgrad=256/4,lin2_dgrad_ctas=128"
run qkvdg4 DCAD_PLAN="qkv_fwd=192,lin2_fwd=256/4,lin1_dgrad=256/4,qkv_dgrad=256/4"
