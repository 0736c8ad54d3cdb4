#!/bin/bash
O=gpurun_out/r2w2d; mkdir -p $O
timeout 300 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_1of2.txt 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt; tail -2 $O/pytest_gpu_1of2.txt
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "smoke rc=$?" | tee -a $O/summary.txt; tail -1 $O/smoke.txt
L="bin/dlnb_launch -n 2"
for op in allreduce reduce_scatter allgather; do
  NCU_SKIP=8 NCU_COUNT=4 timeout 150 $L tools/ncu_rank0.sh $O/ncu_nvl_w2_$op.csv python tools/sweep.py --op $op --dtype f16 --min-bytes 67108864 --max-bytes 67108864 --iters 4 > $O/ncu_nvl_w2_$op.log 2>&1; echo "ncu $op rc=$?" | tee -a $O/summary.txt
done
timeout 200 $L python tools/sweep.py --op reduce_scatter --dtype f16 --ctas 128 --iters 10 --counts 110240167,125472832,31368208 --out $O/sweep_rs_shards.json > $O/sweep_rs_shards.log 2>&1
timeout 200 $L python tools/sweep.py --op allgather --dtype f16 --ctas 128 --iters 10 --counts 110240167,125472832,31368208 --out $O/sweep_ag_shards.json > $O/sweep_ag_shards.log 2>&1
timeout 200 $L python tools/sweep.py --op sendrecv --dtype f16 --ctas 64 --min-bytes 4096 --max-bytes 268435456 --out $O/sweep_sendrecv.json > $O/sweep_sendrecv.log 2>&1; echo "sendrecv rc=$?" | tee -a $O/summary.txt
cat $O/summary.txt
