"""Latency of the slab step's small collective (8 floats, MAX) + read-back, several ways.
Run under torchrun on 2+ GPUs."""
import os
import time

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cpu_group = dist.new_group(backend="gloo")
    s = torch.cuda.Stream()
    t = torch.zeros(8, device="cuda")
    pinned = torch.zeros(8).pin_memory()
    big = torch.zeros(64 << 20, device="cuda")
    res = {}

    def timed(name, fn, iters=200):
        with torch.cuda.stream(s):
            for _ in range(20):
                fn()
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            for _ in range(iters):
                fn()
            torch.cuda.synchronize()
            res[name] = (time.perf_counter() - t0) / iters * 1e6

    def a():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.cpu()

    def b():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        pinned.copy_(t, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def c():
        pinned.copy_(t, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        dist.all_reduce(pinned, op=dist.ReduceOp.MAX, group=cpu_group)

    def d():
        return t.cpu()

    def e():
        pinned.copy_(t, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def f():  # with a busy GPU in front, as in the step (a ~0.3 ms kernel)
        big.add_(1.0)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        pinned.copy_(t, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def g():
        big.add_(1.0)
        pinned.copy_(t, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for name, fn in [("nccl+cpu()", a), ("nccl+pinned", b), ("pinned+gloo", c), ("cpu() only", d),
                     ("pinned only", e), ("kernel+nccl+pinned", f), ("kernel+pinned", g)]:
        timed(name, fn)
    if rank == 0:
        for k, v in res.items():
            print(f"{k:24s} {v:8.1f} us", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
