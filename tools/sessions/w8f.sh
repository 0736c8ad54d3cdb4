#!/bin/bash
# The one 8-GPU session of the round: the driver's bench command (1 timed step), then the sweeps, link traffic and stress.
O=gpurun_out/r2w8f; mkdir -p $O
L="bin/dlnb_launch -n 8"
DLNB_BENCH_C5_ITERS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 1 --warmup 3 --no-e2e > $O/bench_n8.json 2> $O/bench_n8.err; echo "bench rc=$?" | tee -a $O/summary.txt
S="--min-bytes 67108864 --max-bytes 1073741824 --iters 8"
timeout 150 $L python tools/sweep.py --op allreduce --dtype f16 --ctas 0 --min-bytes 4096 --max-bytes 1073741824 --iters 10 --out $O/sweep_ar_auto.json > $O/sweep_ar_auto.log 2>&1; echo "auto rc=$?" | tee -a $O/summary.txt
timeout 150 $L python tools/sweep.py --op allreduce --dtype f16 --algo 3 --mc-ctas 16,32 $S --out $O/sweep_arm.json > $O/sweep_arm.log 2>&1; echo "arm rc=$?" | tee -a $O/summary.txt
timeout 150 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 128 $S --out $O/sweep_rs.json > $O/sweep_rs.log 2>&1
timeout 150 $L python tools/sweep.py --op allgather --dtype f16 --ctas 128 $S --out $O/sweep_ag.json > $O/sweep_ag.log 2>&1
timeout 150 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 128 --iters 8 --counts 110240167,31368208 --out $O/sweep_rs_shards.json > $O/sweep_rs_shards.log 2>&1
timeout 150 $L python tools/sweep.py --op allgather --dtype f16 --ctas 128 --iters 8 --counts 110240167,31368208 --out $O/sweep_ag_shards.json > $O/sweep_ag_shards.log 2>&1
timeout 200 $L python tools/stress.py --launches 8000 --out $O/stress.json > $O/stress.log 2>&1; echo "stress rc=$?" | tee -a $O/summary.txt
cat $O/summary.txt; tail -c 600 $O/bench_n8.err
