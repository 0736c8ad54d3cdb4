#!/bin/bash
# One 1-GPU session: tests, copy probes, footprints, sanitizer, ncu.
O=gpurun_out/r2w1; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "pytest rc=$?" | tee -a $O/summary.txt; tail -3 $O/pytest_gpu.txt
timeout 300 bin/bulk_copybench > $O/bulk_copybench.jsonl 2>&1; echo "bulk rc=$?" | tee -a $O/summary.txt
timeout 120 python tools/footprint_probe.py > $O/footprints.jsonl 2>&1; echo "footprint rc=$?" | tee -a $O/summary.txt
for ov in 1 2 8; do timeout 120 python tools/emu_probe.py --nranks 2 8 --ops reduce_scatter allgather --odd --shift-overlap $ov >> $O/emu_probe_odd.jsonl 2>&1; done
timeout 120 python tools/emu_probe.py --nranks 2 8 > $O/emu_probe.jsonl 2>&1
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_probe.py > $O/sanitizer_memcheck.txt 2>&1; echo "memcheck rc=$?" | tee -a $O/summary.txt
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python tools/sanitize_probe.py > $O/sanitizer_racecheck.txt 2>&1; echo "racecheck rc=$?" | tee -a $O/summary.txt
timeout 600 compute-sanitizer --tool synccheck --error-exitcode 9 python tools/sanitize_probe.py > $O/sanitizer_synccheck.txt 2>&1; echo "synccheck rc=$?" | tee -a $O/summary.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:coll_kernel_emu -c 6 -o $O/emu_odd_full python tools/emu_probe.py --nranks 8 --ops reduce_scatter allgather --odd --reps 1 > $O/ncu_emu_odd.log 2>&1; echo "ncu rc=$?" | tee -a $O/summary.txt
cat $O/summary.txt
