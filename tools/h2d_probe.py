"""python tools/h2d_probe.py            |  torchrun --nproc-per-node N tools/h2d_probe.py
Copy-only probe behind bench.py's e2e numbers: every rank copies a pinned host buffer of the size one step of the
headline workload ships (bf16 q_img + q_txt + mask of its 32768/N rows; --fp32 for the reference's dtype) to its GPU,
all ranks at once, no compute. Prints GB/s per GPU (slowest rank) and in aggregate: the host->device ceiling of the box
that any end-to-end number at this input size sits under."""
import json
import os
import sys

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    esz = 4 if "--fp32" in sys.argv else 2
    b = 32768 // world
    nbytes = b * (49 + 77) * 512 * esz + b * 77 * 4
    host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    host.fill_(1)
    devb = [torch.empty(nbytes, dtype=torch.uint8, device=dev) for _ in range(2)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    out = {}
    for tag, reps in (("warm", 2), ("timed", 8)):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0.record()
        for i in range(reps):
            devb[i % 2].copy_(host, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        out[tag] = e0.elapsed_time(e1) / reps
    ms = torch.tensor([out["timed"]], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = ms.item()
        print(json.dumps({"n_gpus": world, "bytes_per_gpu": nbytes, "host_dtype": "fp32" if esz == 4 else "bf16",
                          "ms_per_copy_slowest_rank": ms, "gb_per_s_per_gpu": nbytes / ms / 1e6,
                          "gb_per_s_aggregate": world * nbytes / ms / 1e6,
                          "pairs_per_s_ceiling": 32768 / (ms * 1e-3), "cpus_visible": len(os.sched_getaffinity(0))}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
