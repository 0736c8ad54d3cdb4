timeout 300 python -m pytest tests/test_gpu_collectives.py -m gpu -x -q -k "allgather or reduce_scatter" > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
for op in allgather reduce_scatter; do
timeout 200 bin/dlnb_launch -n 2 python tools/sweep.py --op $op --dtype f16 --min-bytes 220480334 --max-bytes 220480334 --ctas 64,128 --out gpurun_out/sweep_w2_${op}_odd.json > /dev/null 2>&1
done
python tools/summarize_sweep.py gpurun_out/sweep_w2_allgather_odd.json gpurun_out/sweep_w2_reduce_scatter_odd.json
