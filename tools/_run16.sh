python tools/sweep.py --op allreduce --dtype f16 --min-bytes 66918844 --max-bytes 66918844 --iters 6 > gpurun_out/plain_sweep.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:coll_kernel -s 12 -c 3 -o gpurun_out/prof_ar_w1_tuned python tools/sweep.py --op allreduce --dtype f16 --min-bytes 66918844 --max-bytes 66918844 --iters 6 > gpurun_out/ncu_sweep.log 2>&1
tail -2 gpurun_out/ncu_sweep.log
python tools/sweep.py --op allreduce --dtype f16 --min-bytes 1073741824 --max-bytes 1073741824 --iters 4 > gpurun_out/plain_sweep2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:coll_kernel -s 8 -c 2 -o gpurun_out/prof_ar_w1_1g python tools/sweep.py --op allreduce --dtype f16 --min-bytes 1073741824 --max-bytes 1073741824 --iters 4 > gpurun_out/ncu_sweep2.log 2>&1
tail -2 gpurun_out/ncu_sweep2.log
python tools/sweep.py --op allreduce --dtype f16 --min-bytes 262144 --max-bytes 1073741824 --out gpurun_out/sweep_w1_b2b.json > /dev/null 2>&1; python tools/summarize_sweep.py gpurun_out/sweep_w1_b2b.json
