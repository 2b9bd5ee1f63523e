cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for b in 32 48 64 32; do
DCAD_ADAM_BLOCKS=$b timeout 200 python bench.py --steps 200 --warmup 20 --no-amp-arm > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err; python -c "
import json;d=json.load(open('gpurun_out/z_bench.json'));print('adam_blocks $b', round(d['ms_per_step'],4), round(d['value'],1))"
done
