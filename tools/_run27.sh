timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -4 gpurun_out/pytest_gpu.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 200 python bench.py --steps 1 --no-cpu-baseline > gpurun_out/bench_n1_final.json 2> gpurun_out/bench_n1_final.err; python -c "
import json; d=json.loads(open('gpurun_out/bench_n1_final.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ['value','ratio_to_lower_bound','gpu_launches']}, d['e2e']['value'], d['roofline']['frac'], d['roofline']['avg_launch_ms'], d['clocks'])"
