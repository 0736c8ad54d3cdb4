timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
python tools/spin_probe.py > gpurun_out/spin_probe2.log 2>&1; head -10 gpurun_out/spin_probe2.log
python tools/sweep.py --op allreduce --dtype f16 --min-bytes 262144 --max-bytes 1073741824 --out gpurun_out/sweep_w1_b2b.json > /dev/null 2>&1
for op in allreduce allgather reduce_scatter; do
bin/dlnb_launch -n 2 python tools/sweep.py --op $op --dtype f16 --min-bytes 4096 --max-bytes 1073741824 --ctas 32,64,128,148 --out gpurun_out/sweep_w2_$op.json > gpurun_out/sweep_w2_$op.log 2>&1 || tail -5 gpurun_out/sweep_w2_$op.log
done
bin/dlnb_launch -n 2 python tools/sweep.py --op allreduce --dtype f16 --algo 1 --min-bytes 4096 --max-bytes 4194304 --ctas 8,32 --out gpurun_out/sweep_w2_ar_oneshot.json > /dev/null 2>&1
python tools/summarize_sweep.py gpurun_out/sweep_w1_b2b.json gpurun_out/sweep_w2_*.json
