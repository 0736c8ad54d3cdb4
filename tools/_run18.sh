timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
timeout 300 bin/dlnb_launch -n 2 python tests/mg_worker.py --quick > gpurun_out/mg2.log 2>&1; tail -3 gpurun_out/mg2.log | cut -c1-300
for a in 1 4; do
timeout 100 bin/dlnb_launch -n 2 python tools/sweep.py --op allreduce --dtype f16 --algo $a --min-bytes 4096 --max-bytes 1048576 --ctas 16 --out gpurun_out/sweep_w2_ar_algo$a.json > /dev/null 2>&1
done
python tools/summarize_sweep.py gpurun_out/sweep_w2_ar_algo1.json gpurun_out/sweep_w2_ar_algo4.json
B=bin/b200/cpp/data_parallel
DLNB_LOOP_ITERS=20 timeout 100 bin/dlnb_launch -n 2 $B/dp_loop vit_h_128_bfloat16 4000 . -t fp32 -g > gpurun_out/dploop4000_w2-g.log 2>&1; grep -o '"device_iter_ms_median":[0-9.]*\|"lower_bound_us":[0-9]*\|"cuda_graph":[a-z]*' gpurun_out/dploop4000_w2-g.log | tr '\n' ' '; echo
