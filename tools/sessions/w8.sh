#!/bin/bash
# One 8-GPU tuning session (gpurun --gpus 8): parity, sweeps (in-switch grids, RS / AG both ways, odd shards), stress, NCCL line.
O=gpurun_out/r2w8; mkdir -p $O
L="bin/dlnb_launch -n 8"
(nvidia-smi nvlink -gt d -i 0 | head -12; python -c "
import sys; sys.path.insert(0,'.')
from dlnetbench_b200 import nvlink
l=nvlink.LinkCounters(0); print('LinkCounters ok =', l.ok, '|', l.why)") > $O/nvlink_counters_not_exposed.txt 2>&1
timeout 400 $L python tests/mg_worker.py --quick > $O/mg_worker_quick.txt 2>&1; echo "mg_worker rc=$?" | tee -a $O/summary.txt
timeout 60 bin/peer_copy_peak --gpus 8 > $O/peer_copy_peak.jsonl 2>&1
S="--min-bytes 67108864 --max-bytes 1073741824 --iters 10"
timeout 200 $L python tools/sweep.py --op allreduce --dtype f16 --ctas 0 --min-bytes 4096 --max-bytes 1073741824 --out $O/sweep_ar_auto.json > $O/sweep_ar_auto.log 2>&1; echo "auto rc=$?" | tee -a $O/summary.txt
timeout 200 $L python tools/sweep.py --op allreduce --dtype f16 --algo 3 --mc-ctas 8,16,24,32 $S --out $O/sweep_arm.json > $O/sweep_arm.log 2>&1; echo "arm rc=$?" | tee -a $O/summary.txt
timeout 200 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 128 $S --out $O/sweep_rs.json > $O/sweep_rs.log 2>&1
timeout 200 $L python tools/sweep.py --op reduce_scatter --dtype f16 --algo 3 --mc-ctas 16 --min-bytes 268435456 --max-bytes 268435456 --iters 10 --out $O/sweep_rsm.json > $O/sweep_rsm.log 2>&1
timeout 200 $L python tools/sweep.py --op allgather --dtype f16 --ctas 128 $S --out $O/sweep_ag.json > $O/sweep_ag.log 2>&1
timeout 200 $L python tools/sweep.py --op allgather --dtype f16 --algo 3 --mc-ctas 16 --min-bytes 268435456 --max-bytes 268435456 --iters 10 --out $O/sweep_agm.json > $O/sweep_agm.log 2>&1
for ov in 1 8; do DLNB_SHIFT_OVERLAP=$ov timeout 200 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 128 --iters 10 --counts 110240167,31368208 --out $O/sweep_rs_shards_ov$ov.json > $O/sweep_rs_shards_ov$ov.log 2>&1; done
timeout 200 $L python tools/sweep.py --op allreduce --dtype f16 --ctas 0 --min-bytes 4096 --max-bytes 262144 --strict 1 --out $O/sweep_ar_strict.json > $O/sweep_ar_strict.log 2>&1
timeout 300 $L python tools/stress.py --launches 30000 --out $O/stress.json > $O/stress.log 2>&1; echo "stress rc=$?" | tee -a $O/summary.txt
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 tools/nccl_compare.py --op allreduce --min-bytes 4096 --max-bytes 1073741824 --out $O/nccl_w8_allreduce.json > $O/nccl_w8_allreduce.log 2>&1
for cfg in b200 nccl; do
  timeout 120 $L bin/$cfg/cpp/data_parallel/dp gpt2_l_16_bfloat16 10 . -t fp16 -w 2 -r 5 > $O/proxy_${cfg}_dp_gpt2l.txt 2>&1; echo "$cfg dp rc=$?" | tee -a $O/summary.txt
  DLNB_LOOP_ITERS=40 timeout 120 $L bin/$cfg/cpp/data_parallel/dp_loop vit_h_128_bfloat16 1000 . -t fp32 -w 3 > $O/proxy_${cfg}_dploop_vith1000.txt 2>&1; echo "$cfg dp_loop rc=$?" | tee -a $O/summary.txt
  timeout 120 $L bin/$cfg/cpp/data_parallel/fsdp gpt2_l_16_bfloat16 36 8 . -t fp16 -w 2 -r 5 > $O/proxy_${cfg}_fsdp_gpt2l.txt 2>&1; echo "$cfg fsdp rc=$?" | tee -a $O/summary.txt
done
(ncu --query-metrics 2>/dev/null | grep -i "nvl" | head -40) > $O/ncu_nvl_metrics.txt 2>&1
timeout 120 ncu --metrics nvlrx__bytes.sum,nvltx__bytes.sum,gpu__time_duration.sum -k regex:copy_kernel -c 6 --csv --log-file $O/ncu_nvl_peer_copy.csv bin/peer_copy_peak --gpus 2 --reps 1 > $O/ncu_nvl_peer_copy.log 2>&1; echo "ncu nvl rc=$?" | tee -a $O/summary.txt
cat $O/summary.txt; tail -2 $O/mg_worker_quick.txt
