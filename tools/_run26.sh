timeout 300 python -m pytest tests/test_multigpu.py -m gpu -x -q -k "4_gpus or n_gpus" > gpurun_out/pytest_multigpu4.log 2>&1; tail -3 gpurun_out/pytest_multigpu4.log
B=bin/b200/cpp/hybrid_parallel
P='"device_iter_ms_median":[0-9.]*\|"lower_bound_us":[0-9]*\|"ratio_to_lower_bound":[0-9.]*\|"dp_allreduce_busbw_GBps":[0-9.]*\|"pipe_msg_size_bytes":[0-9]*'
timeout 100 bin/dlnb_launch -n 4 $B/hybrid_2d llama3_8b_64_bfloat16 2 8 . -t fp16 -w 1 -r 1 > gpurun_out/hybrid2d_llama8b_w4.log 2>&1; grep -o "$P" gpurun_out/hybrid2d_llama8b_w4.log | tr '\n' ' '; echo
timeout 100 bin/dlnb_launch -n 4 $B/hybrid_2d gpt2_l_16_bfloat16 4 8 . -t fp16 -r 5 > gpurun_out/hybrid2d_gpt2l_w4.log 2>&1; grep -o "$P" gpurun_out/hybrid2d_gpt2l_w4.log | tr '\n' ' '; echo
