timeout 300 bin/dlnb_launch -n 2 python tests/mg_worker.py --quick > gpurun_out/mg2.log 2>&1; tail -4 gpurun_out/mg2.log | cut -c1-300
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
timeout 200 bin/dlnb_launch -n 2 python tools/sweep.py --op sendrecv --dtype f16 --min-bytes 4096 --max-bytes 1073741824 --ctas 0 --out gpurun_out/sweep_w2_sendrecv.json > gpurun_out/sweep_w2_sendrecv.log 2>&1 || tail -5 gpurun_out/sweep_w2_sendrecv.log
python tools/summarize_sweep.py gpurun_out/sweep_w2_sendrecv.json
B=bin/b200/cpp/hybrid_parallel
timeout 100 bin/dlnb_launch -n 2 $B/hybrid_2d gpt2_l_16_bfloat16 2 4 . -t fp16 -r 3 > gpurun_out/hybrid2d_w2.log 2>&1; grep GLOBAL gpurun_out/hybrid2d_w2.log | cut -c1-1100
