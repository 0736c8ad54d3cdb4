set -x
python bench.py > gpurun_out/bench_r01_n1.json 2> gpurun_out/bench_r01_n1.err; tail -c 3000 gpurun_out/bench_r01_n1.json; tail -5 gpurun_out/bench_r01_n1.err
python tools/spin_probe.py > gpurun_out/spin_probe.log 2>&1; cat gpurun_out/spin_probe.log
python bench.py --steps 1 --no-e2e --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 1827 -c 481 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --no-e2e --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
tail -3 gpurun_out/ncu_bench.log
python tools/sweep.py --op allreduce --dtype f16 --min-bytes 66918844 --max-bytes 66918844 --iters 6 > gpurun_out/plain_sweep.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:coll_kernel -s 6 -c 3 -o gpurun_out/prof_ar_w1 python tools/sweep.py --op allreduce --dtype f16 --min-bytes 66918844 --max-bytes 66918844 --iters 6 > gpurun_out/ncu_sweep.log 2>&1
tail -3 gpurun_out/ncu_sweep.log; ls -la gpurun_out
