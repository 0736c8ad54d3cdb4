cd /root/repo
timeout 400 bin/dlnb_launch -n 8 python tests/mg_worker.py --quick > gpurun_out/mg8.log 2>&1; tail -9 gpurun_out/mg8.log | cut -c1-200
timeout 200 bin/dlnb_launch -n 8 python tools/sweep.py --op allreduce --dtype f16 --algo 3 --min-bytes 262144 --max-bytes 1073741824 --ctas 16,32,64,128 --out gpurun_out/sweep_w8_ar_multimem.json > gpurun_out/sweep_w8_ar_multimem.log 2>&1 || tail -5 gpurun_out/sweep_w8_ar_multimem.log
timeout 200 bin/dlnb_launch -n 8 python tools/sweep.py --op allreduce --dtype f32 --algo 3 --min-bytes 4194304 --max-bytes 268435456 --ctas 32,64 --out gpurun_out/sweep_w8_ar_multimem_f32.json > /dev/null 2>&1
timeout 200 bin/dlnb_launch -n 4 python tools/sweep.py --op allreduce --dtype f16 --algo 3 --min-bytes 4194304 --max-bytes 1073741824 --ctas 32,64 --out gpurun_out/sweep_w4_ar_multimem.json > /dev/null 2>&1
timeout 200 bin/dlnb_launch -n 4 python tools/sweep.py --op allreduce --dtype f16 --algo 2 --min-bytes 4194304 --max-bytes 1073741824 --ctas 64,128 --out gpurun_out/sweep_w4_ar_twoshot.json > /dev/null 2>&1
B=bin/b200/cpp/data_parallel
for slack in 4000 8000 16000; do
DLNB_SPIN_SLACK_NS=$slack DLNB_LOOP_ITERS=30 timeout 60 bin/dlnb_launch -n 8 $B/dp_loop vit_h_128_bfloat16 1000 . -t fp32 > gpurun_out/proxy_dploop_vith_w8_slack$slack.log 2>&1; grep -o '"device_iter_ms_median":[0-9.]*' gpurun_out/proxy_dploop_vith_w8_slack$slack.log
done
DLNB_LOOP_ITERS=30 timeout 60 bin/dlnb_launch -n 8 $B/dp_loop vit_h_128_bfloat16 4000 . -t fp32 > gpurun_out/proxy_dploop_vith4000_w8.log 2>&1; grep -o '"device_iter_ms_median":[0-9.]*\|"lower_bound_us":[0-9]*' gpurun_out/proxy_dploop_vith4000_w8.log
python tools/summarize_sweep.py gpurun_out/sweep_w8_ar_multimem.json gpurun_out/sweep_w8_ar_multimem_f32.json gpurun_out/sweep_w4_ar_multimem.json gpurun_out/sweep_w4_ar_twoshot.json
