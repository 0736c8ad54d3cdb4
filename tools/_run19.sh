cd /root/repo
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/pytest_multigpu.log 2>&1; tail -3 gpurun_out/pytest_multigpu.log
timeout 200 bin/dlnb_launch -n 8 python tools/sweep.py --op allreduce --dtype f16 --min-bytes 4096 --max-bytes 1073741824 --ctas 0 --out gpurun_out/sweep_w8_allreduce_auto.json > gpurun_out/sweep_w8_allreduce_auto.log 2>&1 || tail -5 gpurun_out/sweep_w8_allreduce_auto.log
python tools/summarize_sweep.py gpurun_out/sweep_w8_allreduce_auto.json
B=bin/b200/cpp/data_parallel
P='"device_iter_ms_median":[0-9.]*\|"lower_bound_us":[0-9]*\|"ratio_to_lower_bound":[0-9.]*\|"allreduce_busbw_GBps_avg":[0-9.]*\|"reduce_scatter_last_ms":[0-9.]*\|"reduce_scatter_busbw_GBps":[0-9.]*'
timeout 120 bin/dlnb_launch -n 8 $B/dp llama3_8b_64_bfloat16 240 . -t fp16 -w 1 -r 1 > gpurun_out/proxy_dp_llama8b_w8.log 2>&1; grep -o "$P" gpurun_out/proxy_dp_llama8b_w8.log | tr '\n' ' '; echo
timeout 120 bin/dlnb_launch -n 8 $B/fsdp llama3_8b_64_bfloat16 32 8 . -t fp16 -w 1 -r 1 > gpurun_out/proxy_fsdp_llama8b_w8.log 2>&1; grep -o "$P" gpurun_out/proxy_fsdp_llama8b_w8.log | tr '\n' ' '; echo
DLNB_LOOP_ITERS=30 timeout 60 bin/dlnb_launch -n 8 $B/dp_loop vit_h_128_bfloat16 1000 . -t fp32 > gpurun_out/proxy_dploop_vith_w8.log 2>&1; grep -o "$P" gpurun_out/proxy_dploop_vith_w8.log | tr '\n' ' '; echo
DLNB_LOOP_ITERS=30 timeout 60 bin/dlnb_launch -n 8 $B/dp_loop vit_h_128_bfloat16 4000 . -t fp32 -g > gpurun_out/proxy_dploop_vith4000_graph_w8.log 2>&1; grep -o "$P" gpurun_out/proxy_dploop_vith4000_graph_w8.log | tr '\n' ' '; echo
DLNB_LOOP_ITERS=1 timeout 150 bin/dlnb_launch -n 8 $B/fsdp_loop llama3_70b_16_bfloat16 80 8 . -t fp16 -w 0 > gpurun_out/proxy_fsdploop_70b_w8.log 2>&1; grep -o "$P" gpurun_out/proxy_fsdploop_70b_w8.log | tr '\n' ' '; echo
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 1 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; tail -c 1800 gpurun_out/bench_n8.json; tail -2 gpurun_out/bench_n8.err | cut -c1-300
