cd /root/repo
timeout 400 bin/dlnb_launch -n 8 python tests/mg_worker.py --quick > gpurun_out/mg8.log 2>&1; tail -8 gpurun_out/mg8.log | cut -c1-120
B=bin/b200/cpp/data_parallel
P='"device_iter_ms_median":[0-9.]*\|"lower_bound_us":[0-9]*\|"ratio_to_lower_bound":[0-9.]*'
for l in 1 2; do
DLNB_LANES=$l DLNB_LOOP_ITERS=30 timeout 60 bin/dlnb_launch -n 8 $B/dp_loop vit_h_128_bfloat16 4000 . -t fp32 -g > gpurun_out/proxy_dploop_vith4000_graph_w8_lanes$l.log 2>&1; grep -o "$P" gpurun_out/proxy_dploop_vith4000_graph_w8_lanes$l.log | tr '\n' ' '; echo lanes=$l
done
DLNB_LOOP_ITERS=30 timeout 60 bin/dlnb_launch -n 8 $B/dp_loop vit_h_128_bfloat16 4000 . -t fp32 > gpurun_out/proxy_dploop_vith4000_nograph_w8.log 2>&1; grep -o "$P" gpurun_out/proxy_dploop_vith4000_nograph_w8.log | tr '\n' ' '; echo nograph
timeout 100 bin/dlnb_launch -n 8 python tools/sweep.py --op allreduce --dtype f16 --min-bytes 4096 --max-bytes 16777216 --ctas 0 --out gpurun_out/sweep_w8_allreduce_auto_lanes.json > /dev/null 2>&1
python tools/summarize_sweep.py gpurun_out/sweep_w8_allreduce_auto_lanes.json
