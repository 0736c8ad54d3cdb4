B=bin/b200/cpp/data_parallel
for l in 1 2; do
DLNB_LANES=$l DLNB_LOOP_ITERS=20 timeout 100 bin/dlnb_launch -n 2 $B/dp_loop vit_h_128_bfloat16 4000 . -t fp32 -g > gpurun_out/dploop4000_w2-g.log 2>&1; grep -o '"device_iter_ms_median":[0-9.]*\|"lower_bound_us":[0-9]*\|"cuda_graph":[a-z]*' gpurun_out/dploop4000_w2-g.log | tr '\n' ' '; echo lanes=$l
done
timeout 300 bin/dlnb_launch -n 2 python tests/mg_worker.py --quick > gpurun_out/mg2.log 2>&1; tail -2 gpurun_out/mg2.log | cut -c1-300
DLNB_LOOP_ITERS=20 timeout 100 bin/dlnb_launch -n 2 $B/dp_loop vit_h_128_bfloat16 1000 . -t fp32 > gpurun_out/dploop1000_w2.log 2>&1; grep -o '"device_iter_ms_median":[0-9.]*\|"lower_bound_us":[0-9]*' gpurun_out/dploop1000_w2.log | tr '\n' ' '; echo
