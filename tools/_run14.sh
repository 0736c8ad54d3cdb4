timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
timeout 600 bin/dlnb_launch -n 2 python tests/mg_worker.py > gpurun_out/mg2.log 2>&1; tail -4 gpurun_out/mg2.log | cut -c1-300
for op in allgather reduce_scatter; do
timeout 200 bin/dlnb_launch -n 2 python tools/sweep.py --op $op --dtype f16 --min-bytes 220480334 --max-bytes 220480334 --ctas 64,128 --out gpurun_out/sweep_w2_${op}_odd.json > /dev/null 2>&1
done
python tools/summarize_sweep.py gpurun_out/sweep_w2_allgather_odd.json gpurun_out/sweep_w2_reduce_scatter_odd.json
