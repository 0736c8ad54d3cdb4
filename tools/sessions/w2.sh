#!/bin/bash
# One 2-GPU measurement session (gpurun --gpus 2): parity matrix, link peak, sweeps with link counters, stress, bench.
O=gpurun_out/r2w2; mkdir -p $O
L="bin/dlnb_launch -n 2"
timeout 600 $L python tests/mg_worker.py > $O/mg_worker_full.txt 2>&1; echo "mg_worker rc=$?" | tee -a $O/summary.txt
timeout 120 bin/peer_copy_peak --gpus 2 > $O/peer_copy_peak.jsonl 2>&1; echo "peer rc=$?" | tee -a $O/summary.txt
for ctas in 32,64,128; do :; done
timeout 300 $L python tools/sweep.py --op allreduce --dtype f16 --algo 2 --ctas 32,64,128 --min-bytes 67108864 --max-bytes 1073741824 --nvlink --out $O/sweep_ar2.json > $O/sweep_ar2.log 2>&1; echo "sweep ar2 rc=$?" | tee -a $O/summary.txt
timeout 300 $L python tools/sweep.py --op allreduce --dtype f16 --ctas 0 --min-bytes 4096 --max-bytes 1073741824 --out $O/sweep_ar_auto.json > $O/sweep_ar_auto.log 2>&1; echo "sweep auto rc=$?" | tee -a $O/summary.txt
timeout 200 $L python tools/sweep.py --op allreduce --dtype f16 --ctas 0 --min-bytes 4096 --max-bytes 1048576 --strict 1 --out $O/sweep_ar_strict.json > $O/sweep_ar_strict.log 2>&1
timeout 300 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 64,128 --min-bytes 67108864 --max-bytes 1073741824 --nvlink --out $O/sweep_rs.json > $O/sweep_rs.log 2>&1; echo "sweep rs rc=$?" | tee -a $O/summary.txt
timeout 300 $L python tools/sweep.py --op allgather --dtype f16 --ctas 64,128 --min-bytes 67108864 --max-bytes 1073741824 --nvlink --out $O/sweep_ag.json > $O/sweep_ag.log 2>&1; echo "sweep ag rc=$?" | tee -a $O/summary.txt
timeout 300 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 128 --counts 110240167,125472832 --nvlink --out $O/sweep_rs_shards.json > $O/sweep_rs_shards.log 2>&1; echo "sweep rs shards rc=$?" | tee -a $O/summary.txt
timeout 300 $L python tools/sweep.py --op sendrecv --dtype f16 --ctas 64 --min-bytes 4096 --max-bytes 268435456 --nvlink --out $O/sweep_sendrecv.json > $O/sweep_sendrecv.log 2>&1
timeout 400 $L python tools/stress.py --launches 30000 --out $O/stress.json > $O/stress.log 2>&1; echo "stress rc=$?" | tee -a $O/summary.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 1 --warmup 3 > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench rc=$?" | tee -a $O/summary.txt
tail -c 1500 $O/bench_n2.err
cat $O/summary.txt
