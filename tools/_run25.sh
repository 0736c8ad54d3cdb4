timeout 200 bin/dlnb_launch -n 2 python tests/mg_worker.py --quick > gpurun_out/mg2.log 2>&1; tail -3 gpurun_out/mg2.log | cut -c1-300
timeout 100 bin/dlnb_launch -n 2 python tools/sweep.py --op sendrecv --dtype f16 --iters 8 --min-bytes 4096 --max-bytes 268435456 --ctas 0 --out gpurun_out/sweep_w2_sendrecv.json > gpurun_out/sweep_w2_sendrecv.log 2>&1 || tail -5 gpurun_out/sweep_w2_sendrecv.log
python tools/summarize_sweep.py gpurun_out/sweep_w2_sendrecv.json
