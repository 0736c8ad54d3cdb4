#!/bin/bash
O=gpurun_out/r2w2b; mkdir -p $O
L="bin/dlnb_launch -n 2"
(nvidia-smi nvlink -gt d -i 0 | head -40; nvidia-smi nvlink -h | head -60) > $O/nvlink_smi.txt 2>&1
timeout 120 bin/peer_copy_peak --gpus 2 > $O/peer_copy_peak.jsonl 2>&1; echo "peer rc=$?" | tee -a $O/summary.txt
timeout 120 bin/pdl_copybench > $O/pdl_copybench.txt 2>&1; echo "pdl rc=$?" | tee -a $O/summary.txt
for ov in 1 2 8; do DLNB_SHIFT_OVERLAP=$ov timeout 200 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 128 --counts 110240167,125472832 --nvlink --out $O/sweep_rs_shards_ov$ov.json > $O/sweep_rs_shards_ov$ov.log 2>&1; done
for ov in 1 8; do DLNB_SHIFT_OVERLAP=$ov timeout 200 $L python tools/sweep.py --op allgather --dtype f16 --ctas 128 --counts 110240167,125472832 --out $O/sweep_ag_shards_ov$ov.json > $O/sweep_ag_shards_ov$ov.log 2>&1; done
timeout 200 $L python tools/sweep.py --op sendrecv --dtype f16 --ctas 64 --min-bytes 4096 --max-bytes 268435456 --out $O/sweep_sendrecv.json > $O/sweep_sendrecv.log 2>&1; echo "sendrecv rc=$?" | tee -a $O/summary.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --model gpt2_l_16_bfloat16 --buckets 10 --steps 2 --no-e2e --no-cpu-baseline > $O/bench_small_n2.json 2> $O/bench_small_n2.err; echo "bench rc=$?" | tee -a $O/summary.txt
cat $O/summary.txt
