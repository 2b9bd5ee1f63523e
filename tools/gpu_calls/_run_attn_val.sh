cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_x_fullsize_gpu.py tests/test_engine_gpu.py -x -q -m gpu -k "attn or attention or parity or full or training or engine" > gpurun_out/av_pytest.log 2>&1; tail -3 gpurun_out/av_pytest.log | cut -c1-300
for i in 1 2; do
timeout 200 python bench.py --steps 200 --warmup 20 --no-amp-arm > gpurun_out/av_bench.json 2> gpurun_out/av_bench.err; python -c "
import json;d=json.load(open('gpurun_out/av_bench.json'));print('new', round(d['ms_per_step'],4), round(d['value'],1))"
done
