#!/bin/bash
# Run the command on every rank, rank 0 under ncu collecting ONLY the NVLink byte counters of the collective kernels
# (one pass: no kernel replay, so the flag barriers with the un-profiled peers are not disturbed). NVML / nvidia-smi do not
# expose the link counters on this pool; ncu's nvlrx__bytes / nvltx__bytes do.
#   bin/dlnb_launch -n 8 tools/ncu_rank0.sh out.csv python tools/sweep.py --op allreduce ...
OUT=$1; shift
if [ "${DLNB_RANK:-${RANK:-0}}" = "0" ]; then
  exec ncu --metrics nvlrx__bytes.sum,nvltx__bytes.sum,nvlrx__bytes_data_user.sum,nvltx__bytes_data_user.sum -k regex:coll_kernel -s ${NCU_SKIP:-0} -c ${NCU_COUNT:-6} --csv --log-file "$OUT" "$@"
else
  exec "$@"
fi
