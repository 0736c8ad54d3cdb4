DLNB_DEBUG=1 timeout 600 bin/dlnb_launch -n 2 python tests/mg_worker.py --quick > gpurun_out/mg2.log 2>&1; tail -6 gpurun_out/mg2.log
timeout 200 bin/dlnb_launch -n 2 python tools/sweep.py --op allreduce --dtype f16 --algo 3 --min-bytes 65536 --max-bytes 1073741824 --ctas 16,32,64,128 --out gpurun_out/sweep_w2_ar_multimem.json > gpurun_out/sweep_w2_ar_multimem.log 2>&1 || tail -5 gpurun_out/sweep_w2_ar_multimem.log
python tools/summarize_sweep.py gpurun_out/sweep_w2_ar_multimem.json
