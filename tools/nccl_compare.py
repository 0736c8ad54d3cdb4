"""NCCL comparison line (never the product): the same collectives, sizes and busBW convention
as tools/sweep.py, through torch.distributed's NCCL backend — what the reference's
CCLCommunicator would call (cpp/proxy_classes.hpp:149-209).

  torchrun --nproc-per-node N tools/nccl_compare.py --op allreduce --out gpurun_out/nccl_w8_allreduce.json
"""
import argparse
import json
import os

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--op", default="allreduce", choices=["allreduce", "allgather", "reduce_scatter"])
    ap.add_argument("--dtype", default="f16", choices=["f32", "f16", "bf16"])
    ap.add_argument("--min-bytes", type=int, default=1 << 12)
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dt = {"f32": torch.float32, "f16": torch.float16, "bf16": torch.bfloat16}[a.dtype]
    esz = torch.tensor([], dtype=dt).element_size()
    rows, size = [], a.min_bytes
    while size <= a.max_bytes:
        count = size // esz
        per = max(count // world, 1)
        nbuf = max(2, min(a.iters, (300 << 20) // max(size, 1) + 1))
        if a.op == "allreduce":
            ins = [torch.ones(count, dtype=dt, device="cuda") for _ in range(nbuf)]
            outs = ins
            wire = 2 * (world - 1) / world * count * esz
        elif a.op == "allgather":
            ins = [torch.ones(per, dtype=dt, device="cuda") for _ in range(nbuf)]
            outs = [torch.empty(per * world, dtype=dt, device="cuda") for _ in range(nbuf)]
            wire = (world - 1) * per * esz
        else:
            ins = [torch.ones(per * world, dtype=dt, device="cuda") for _ in range(nbuf)]
            outs = [torch.empty(per, dtype=dt, device="cuda") for _ in range(nbuf)]
            wire = (world - 1) * per * esz
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for rep in range(2):
            torch.cuda.synchronize()
            dist.barrier()
            e0.record()
            for i in range(a.iters):
                if a.op == "allreduce":
                    dist.all_reduce(ins[i % nbuf])
                elif a.op == "allgather":
                    dist.all_gather_into_tensor(outs[i % nbuf], ins[i % nbuf])
                else:
                    dist.reduce_scatter_tensor(outs[i % nbuf], ins[i % nbuf])
            e1.record()
            torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / a.iters
        t = torch.tensor([us], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        us = float(t.item())
        row = dict(op=a.op, dtype=a.dtype, world=world, bytes=(count if a.op == "allreduce" else per * world) * esz, ctas=0,
                   us_b2b=us, us_slot_med=us, busbw_GBps=wire / (us * 1e-6) / 1e9, impl="nccl " + ".".join(map(str, torch.cuda.nccl.version())))
        rows.append(row)
        del ins, outs
        size *= 4
    if rank == 0 and a.out:
        json.dump(rows, open(a.out, "w"), indent=1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
