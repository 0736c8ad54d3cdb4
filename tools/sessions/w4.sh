#!/bin/bash
# One 4-GPU session (gpurun --gpus 4): parity, in-switch grid sweep, two-shot comparison, RS / AG, link counters.
O=gpurun_out/r2w4; mkdir -p $O
L="bin/dlnb_launch -n 4"
(nvidia-smi nvlink -gt d -i 0 | head -30; nvidia-smi topo -m | head -12) > $O/nvlink_smi.txt 2>&1
timeout 600 $L python tests/mg_worker.py --quick > $O/mg_worker_quick.txt 2>&1; echo "mg_worker rc=$?" | tee -a $O/summary.txt
timeout 120 bin/peer_copy_peak --gpus 4 > $O/peer_copy_peak.jsonl 2>&1
S="--min-bytes 67108864 --max-bytes 1073741824 --nvlink"
timeout 300 $L python tools/sweep.py --op allreduce --dtype f16 --algo 3 --mc-ctas 16,24,32,48,64 $S --out $O/sweep_arm.json > $O/sweep_arm.log 2>&1; echo "arm rc=$?" | tee -a $O/summary.txt
timeout 300 $L python tools/sweep.py --op allreduce --dtype f16 --algo 2 --ctas 64,128 $S --out $O/sweep_ar2.json > $O/sweep_ar2.log 2>&1; echo "ar2 rc=$?" | tee -a $O/summary.txt
timeout 300 $L python tools/sweep.py --op allreduce --dtype f16 --ctas 0 --min-bytes 4096 --max-bytes 1073741824 --out $O/sweep_ar_auto.json > $O/sweep_ar_auto.log 2>&1; echo "auto rc=$?" | tee -a $O/summary.txt
timeout 300 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 64,128 $S --out $O/sweep_rs.json > $O/sweep_rs.log 2>&1
timeout 300 $L python tools/sweep.py --op reduce_scatter --dtype f16 --algo 3 --mc-ctas 16,32,64 $S --out $O/sweep_rsm.json > $O/sweep_rsm.log 2>&1
timeout 300 $L python tools/sweep.py --op allgather --dtype f16 --ctas 64,128 $S --out $O/sweep_ag.json > $O/sweep_ag.log 2>&1
timeout 300 $L python tools/sweep.py --op allgather --dtype f16 --algo 3 --mc-ctas 16,32,64 $S --out $O/sweep_agm.json > $O/sweep_agm.log 2>&1
timeout 300 $L python tools/stress.py --launches 30000 --out $O/stress.json > $O/stress.log 2>&1; echo "stress rc=$?" | tee -a $O/summary.txt
(ncu --query-metrics 2>/dev/null | grep -i "nvl" | head -40) > $O/ncu_nvl_metrics.txt 2>&1
timeout 200 ncu --metrics nvlrx__bytes.sum,nvltx__bytes.sum,gpu__time_duration.sum -k regex:copy_kernel -c 6 --csv --log-file $O/ncu_nvl_peer_copy.csv bin/peer_copy_peak --gpus 2 --reps 1 > $O/ncu_nvl_peer_copy.log 2>&1; echo "ncu nvl rc=$?" | tee -a $O/summary.txt
for op in allreduce allgather reduce_scatter; do timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29520 tools/nccl_compare.py --op $op --min-bytes 4096 --max-bytes 1073741824 --out $O/nccl_w4_$op.json > $O/nccl_w4_$op.log 2>&1; done
timeout 300 $L bin/nccl/cpp/data_parallel/dp llama3_8b_64_bfloat16 240 . -t fp16 -w 1 -r 1 > $O/proxy_nccl_dp_llama8b.txt 2>&1; echo "nccl dp rc=$?" | tee -a $O/summary.txt
timeout 300 $L bin/nccl/cpp/data_parallel/fsdp llama3_8b_64_bfloat16 32 4 . -t fp16 -w 1 -r 1 > $O/proxy_nccl_fsdp_llama8b.txt 2>&1; echo "nccl fsdp rc=$?" | tee -a $O/summary.txt
DLNB_LOOP_ITERS=50 timeout 300 $L bin/nccl/cpp/data_parallel/dp_loop vit_h_128_bfloat16 1000 . -t fp32 -w 3 > $O/proxy_nccl_dploop_vith1000.txt 2>&1; echo "nccl dp_loop rc=$?" | tee -a $O/summary.txt
cat $O/summary.txt; tail -2 $O/mg_worker_quick.txt
