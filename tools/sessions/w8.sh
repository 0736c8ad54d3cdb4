#!/bin/bash
# One 8-GPU tuning session (gpurun --gpus 8): parity, link peak, sweeps (in-switch grids, RS / AG both ways, odd shards), stress.
O=gpurun_out/r2w8; mkdir -p $O
L="bin/dlnb_launch -n 8"
(nvidia-smi nvlink -gt d -i 0 | head -30) > $O/nvlink_smi.txt 2>&1
timeout 600 $L python tests/mg_worker.py --quick > $O/mg_worker_quick.txt 2>&1; echo "mg_worker rc=$?" | tee -a $O/summary.txt
timeout 120 bin/peer_copy_peak --gpus 8 > $O/peer_copy_peak.jsonl 2>&1
S="--min-bytes 67108864 --max-bytes 1073741824 --nvlink"
timeout 300 $L python tools/sweep.py --op allreduce --dtype f16 --ctas 0 --min-bytes 4096 --max-bytes 1073741824 --nvlink --out $O/sweep_ar_auto.json > $O/sweep_ar_auto.log 2>&1; echo "auto rc=$?" | tee -a $O/summary.txt
timeout 300 $L python tools/sweep.py --op allreduce --dtype f16 --algo 3 --mc-ctas 8,16,24,32 $S --out $O/sweep_arm.json > $O/sweep_arm.log 2>&1; echo "arm rc=$?" | tee -a $O/summary.txt
timeout 300 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 64,128 $S --out $O/sweep_rs.json > $O/sweep_rs.log 2>&1
timeout 300 $L python tools/sweep.py --op reduce_scatter --dtype f16 --algo 3 --mc-ctas 16,32,64 $S --out $O/sweep_rsm.json > $O/sweep_rsm.log 2>&1
timeout 300 $L python tools/sweep.py --op allgather --dtype f16 --ctas 64,128 $S --out $O/sweep_ag.json > $O/sweep_ag.log 2>&1
timeout 300 $L python tools/sweep.py --op allgather --dtype f16 --algo 3 --mc-ctas 16,32,64 $S --out $O/sweep_agm.json > $O/sweep_agm.log 2>&1
for ov in 1 2 8; do DLNB_SHIFT_OVERLAP=$ov timeout 200 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 128 --counts 110240167,31368208 --out $O/sweep_rs_shards_ov$ov.json > $O/sweep_rs_shards_ov$ov.log 2>&1; done
timeout 300 $L python tools/sweep.py --op allreduce --dtype f16 --ctas 0 --min-bytes 4096 --max-bytes 1048576 --strict 1 --out $O/sweep_ar_strict.json > $O/sweep_ar_strict.log 2>&1
timeout 400 $L python tools/stress.py --launches 50000 --out $O/stress.json > $O/stress.log 2>&1; echo "stress rc=$?" | tee -a $O/summary.txt
for op in allreduce allgather reduce_scatter; do timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 tools/nccl_compare.py --op $op --min-bytes 4096 --max-bytes 1073741824 --out $O/nccl_w8_$op.json > $O/nccl_w8_$op.log 2>&1; done
timeout 300 $L bin/nccl/cpp/data_parallel/dp llama3_8b_64_bfloat16 240 . -t fp16 -w 1 -r 1 > $O/proxy_nccl_dp_llama8b.txt 2>&1; echo "nccl dp rc=$?" | tee -a $O/summary.txt
timeout 300 $L bin/nccl/cpp/data_parallel/fsdp llama3_8b_64_bfloat16 32 8 . -t fp16 -w 1 -r 1 > $O/proxy_nccl_fsdp_llama8b.txt 2>&1; echo "nccl fsdp rc=$?" | tee -a $O/summary.txt
DLNB_LOOP_ITERS=50 timeout 300 $L bin/nccl/cpp/data_parallel/dp_loop vit_h_128_bfloat16 1000 . -t fp32 -w 3 > $O/proxy_nccl_dploop_vith1000.txt 2>&1; echo "nccl dp_loop rc=$?" | tee -a $O/summary.txt
cat $O/summary.txt; tail -2 $O/mg_worker_quick.txt
