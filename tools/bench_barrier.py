"""What a barrier between the network pass and the local pass costs on N GPUs (torchrun):
NCCL all-reduce of one int32 on idle GPUs and right after a kernel that stored into the peers' memory.
  python -m torch.distributed.run --nproc-per-node 2 tools/bench_barrier.py
"""
import json
import os

import torch
import torch.distributed as dist


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    token = torch.zeros(1, dtype=torch.int32, device=dev)
    big = torch.empty(1 << 28, dtype=torch.int64, device=dev)  # 2 GiB
    out = {}

    def timed(fn, reps=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(2 * reps)]
        for i in range(reps):
            e[2 * i].record()
            fn()
            e[2 * i + 1].record()
        torch.cuda.synchronize()
        ts = sorted(e[2 * i].elapsed_time(e[2 * i + 1]) * 1e3 for i in range(reps))
        return {"median_us": ts[len(ts) // 2], "min_us": ts[0], "max_us": ts[-1]}

    out["all_reduce_idle"] = timed(lambda: dist.all_reduce(token))

    def after_fill():
        big.fill_(1)  # 2 GiB of local stores, ~0.6 ms
        dist.all_reduce(token)
    out["fill_2GiB_then_all_reduce"] = timed(after_fill)
    out["fill_2GiB_alone"] = timed(lambda: big.fill_(1))
    if rank == 0:
        print(json.dumps({"world": world, **out}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
