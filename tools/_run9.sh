# 8-GPU session: parity first, then sweeps, NCCL line, proxies. Everything logs to gpurun_out/.
cd /root/repo
nvidia-smi -L | wc -l; nproc; free -g | sed -n 2p
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/pytest_multigpu.log 2>&1; tail -5 gpurun_out/pytest_multigpu.log
for op in allreduce allgather reduce_scatter; do
timeout 200 bin/dlnb_launch -n 8 python tools/sweep.py --op $op --dtype f16 --min-bytes 4096 --max-bytes 1073741824 --ctas 32,64,128 --out gpurun_out/sweep_w8_$op.json > gpurun_out/sweep_w8_$op.log 2>&1 || tail -5 gpurun_out/sweep_w8_$op.log
done
timeout 100 bin/dlnb_launch -n 8 python tools/sweep.py --op allreduce --dtype f16 --algo 1 --min-bytes 4096 --max-bytes 4194304 --ctas 16,64 --out gpurun_out/sweep_w8_ar_oneshot.json > /dev/null 2>&1
for op in allreduce allgather reduce_scatter; do
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 tools/nccl_compare.py --op $op --out gpurun_out/nccl_w8_$op.json > gpurun_out/nccl_w8_$op.log 2>&1 || tail -5 gpurun_out/nccl_w8_$op.log
done
B=bin/b200/cpp/data_parallel
timeout 120 bin/dlnb_launch -n 8 $B/dp llama3_8b_64_bfloat16 240 . -t fp16 -w 1 -r 1 > gpurun_out/proxy_dp_llama8b_w8.log 2>&1; tail -c 1500 gpurun_out/proxy_dp_llama8b_w8.log | head -c 900
timeout 120 bin/dlnb_launch -n 8 $B/fsdp llama3_8b_64_bfloat16 32 8 . -t fp16 -w 1 -r 1 > gpurun_out/proxy_fsdp_llama8b_w8.log 2>&1; grep GLOBAL gpurun_out/proxy_fsdp_llama8b_w8.log | cut -c1-900
DLNB_LOOP_ITERS=50 timeout 60 bin/dlnb_launch -n 8 $B/dp_loop vit_h_128_bfloat16 1000 . -t fp32 > gpurun_out/proxy_dploop_vith_w8.log 2>&1; grep GLOBAL gpurun_out/proxy_dploop_vith_w8.log | cut -c1-900
DLNB_LOOP_ITERS=1 timeout 150 bin/dlnb_launch -n 8 $B/fsdp_loop llama3_70b_16_bfloat16 80 8 . -t fp16 -w 0 > gpurun_out/proxy_fsdploop_70b_w8.log 2>&1; grep GLOBAL gpurun_out/proxy_fsdploop_70b_w8.log | cut -c1-900
python tools/summarize_sweep.py gpurun_out/sweep_w8_allreduce.json | grep "ctas=128"
