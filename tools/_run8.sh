timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
for op in allreduce reduce_scatter allgather; do
bin/dlnb_launch -n 2 python tools/sweep.py --op $op --dtype f16 --min-bytes 4096 --max-bytes 67108864 --ctas 32,128 --out gpurun_out/sweep_w2_$op.json > gpurun_out/sweep_w2_$op.log 2>&1 || tail -5 gpurun_out/sweep_w2_$op.log
done
bin/dlnb_launch -n 2 python tools/sweep.py --op allreduce --dtype f16 --algo 1 --min-bytes 4096 --max-bytes 4194304 --ctas 8,32 --out gpurun_out/sweep_w2_ar_oneshot.json > /dev/null 2>&1
python tools/summarize_sweep.py gpurun_out/sweep_w2_*.json
