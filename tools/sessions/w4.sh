#!/bin/bash
# One 4-GPU session (gpurun --gpus 4): parity, in-switch grid sweep, two-shot comparison, RS / AG both ways, NCCL line.
O=gpurun_out/r2w4; mkdir -p $O
L="bin/dlnb_launch -n 4"
timeout 400 $L python tests/mg_worker.py --quick > $O/mg_worker_quick.txt 2>&1; echo "mg_worker rc=$?" | tee -a $O/summary.txt
timeout 60 bin/peer_copy_peak --gpus 4 > $O/peer_copy_peak.jsonl 2>&1
S="--min-bytes 67108864 --max-bytes 1073741824 --iters 10"
timeout 200 $L python tools/sweep.py --op allreduce --dtype f16 --algo 3 --mc-ctas 16,32,48 $S --out $O/sweep_arm.json > $O/sweep_arm.log 2>&1; echo "arm rc=$?" | tee -a $O/summary.txt
timeout 200 $L python tools/sweep.py --op allreduce --dtype f16 --algo 2 --ctas 64,128 $S --out $O/sweep_ar2.json > $O/sweep_ar2.log 2>&1; echo "ar2 rc=$?" | tee -a $O/summary.txt
timeout 200 $L python tools/sweep.py --op allreduce --dtype f16 --ctas 0 --min-bytes 4096 --max-bytes 16777216 --out $O/sweep_ar_auto_small.json > $O/sweep_ar_auto_small.log 2>&1
timeout 200 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 128 $S --out $O/sweep_rs.json > $O/sweep_rs.log 2>&1
timeout 200 $L python tools/sweep.py --op reduce_scatter --dtype f16 --algo 3 --mc-ctas 32 $S --out $O/sweep_rsm.json > $O/sweep_rsm.log 2>&1
timeout 200 $L python tools/sweep.py --op allgather --dtype f16 --ctas 128 $S --out $O/sweep_ag.json > $O/sweep_ag.log 2>&1
timeout 200 $L python tools/sweep.py --op allgather --dtype f16 --algo 3 --mc-ctas 32 $S --out $O/sweep_agm.json > $O/sweep_agm.log 2>&1
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29520 tools/nccl_compare.py --op allreduce --min-bytes 4096 --max-bytes 1073741824 --out $O/nccl_w4_allreduce.json > $O/nccl_w4_allreduce.log 2>&1
cat $O/summary.txt; tail -2 $O/mg_worker_quick.txt
