timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
timeout 300 bin/dlnb_launch -n 2 python tests/mg_worker.py --quick > gpurun_out/mg2.log 2>&1; tail -3 gpurun_out/mg2.log | cut -c1-300
B=bin/b200/cpp/data_parallel
for g in "" "-g"; do
DLNB_LOOP_ITERS=20 timeout 100 bin/dlnb_launch -n 2 $B/dp_loop vit_h_128_bfloat16 4000 . -t fp32 $g > gpurun_out/dploop4000_w2$g.log 2>&1; grep -o '"device_iter_ms_median":[0-9.]*\|"lower_bound_us":[0-9]*\|"cuda_graph":[a-z]*' gpurun_out/dploop4000_w2$g.log | tr '\n' ' '; echo
done
timeout 100 bin/dlnb_launch -n 2 $B/fsdp gpt2_l_16_bfloat16 36 2 . -t fp16 -r 3 -g > gpurun_out/fsdp_graph_w2.log 2>&1; grep -o '"device_iter_ms_median":[0-9.]*\|"lower_bound_us":[0-9]*\|"cuda_graph":[a-z]*' gpurun_out/fsdp_graph_w2.log | tr '\n' ' '; tail -2 gpurun_out/fsdp_graph_w2.log | cut -c1-200
