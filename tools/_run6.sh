set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python tools/spin_probe.py 2>&1 | head -12
python tools/sweep.py --op allreduce --dtype f16 --min-bytes 262144 --max-bytes 1073741824 --out gpurun_out/sweep_w1_b2b.json 2>&1 | cut -c1-200
for op in allreduce allgather reduce_scatter; do
bin/dlnb_launch -n 2 python tools/sweep.py --op $op --dtype f16 --min-bytes 16384 --max-bytes 1073741824 --ctas 16,32,64,128 --out gpurun_out/sweep_w2_$op.json 2>&1 | cut -c1-220
done
bin/dlnb_launch -n 2 python tools/sweep.py --op allreduce --dtype f16 --algo 1 --min-bytes 4096 --max-bytes 4194304 --ctas 8,32 2>&1 | cut -c1-220
