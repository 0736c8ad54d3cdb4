#!/bin/bash
# 1-GPU session: tests, copy probes (bulk-async, PDL), footprints, bench with / without PDL, sweep, ncu captures.
O=gpurun_out/r2w1; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt; tail -3 $O/pytest_gpu.txt
timeout 200 bin/bulk_copybench > $O/bulk_copybench.jsonl 2>&1; echo "bulk rc=$?" | tee -a $O/summary.txt
timeout 100 bin/pdl_copybench > $O/pdl_copybench.txt 2>&1
timeout 120 python tools/footprint_probe.py > $O/footprints.jsonl 2>&1; echo "footprint rc=$?" | tee -a $O/summary.txt
timeout 120 python tools/emu_probe.py --nranks 2 8 > $O/emu_probe.jsonl 2>&1
timeout 120 python tools/emu_probe.py --nranks 2 8 --ops reduce_scatter allgather --odd > $O/emu_probe_odd.jsonl 2>&1
DLNB_PDL=0 timeout 300 python bench.py --steps 1 --warmup 3 --no-sub --no-e2e --no-cpu-baseline --no-contention > $O/bench_n1_nopdl.json 2> $O/bench_n1_nopdl.err; echo "bench nopdl rc=$?" | tee -a $O/summary.txt
timeout 600 python bench.py --steps 1 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?" | tee -a $O/summary.txt
timeout 200 python tools/sweep.py --op allreduce --dtype f16 --min-bytes 4096 --max-bytes 1073741824 --out $O/sweep_w1_allreduce.json > $O/sweep_w1.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:coll_kernel -c 3 -o $O/ar_w1_67mb python tools/sweep.py --op allreduce --dtype f16 --min-bytes 66918844 --max-bytes 66918844 --iters 4 > $O/ncu_w1.log 2>&1; echo "ncu rc=$?" | tee -a $O/summary.txt
timeout 300 ncu --set full --clock-control none --import-source on -k regex:coll_kernel_emu -c 4 -o $O/emu_odd_full python tools/emu_probe.py --nranks 8 --ops reduce_scatter allgather --odd --reps 1 > $O/ncu_emu_odd.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_small.csv python bench.py --model gpt2_l_16_bfloat16 --buckets 10 --steps 1 --warmup 3 --no-sub --no-e2e --no-cpu-baseline --no-contention --no-parity > $O/ncu_launches.log 2>&1; echo "launch list rc=$?" | tee -a $O/summary.txt
cat $O/summary.txt
