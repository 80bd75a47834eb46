"""Developer tool: cudaEvent timeline of one multi-rank join on rank 0 (torchrun, N GPUs).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/timeline.py

Prints the device time between the launcher's steps (histograms + plan, network scatter of both
relations, per relation: wait for the own pushes, barrier, local pass; build/probe).
"""
import importlib
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")

RADIX_BITS = {2: (8, 9), 4: (9, 9), 8: (9, 10)}  # as bench.py


def main():
    world, rank = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"])
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    n = 1 << 28
    G = n * world
    net, loc = RADIX_BITS[world]
    cfg = rhj.make_config(net, loc, 32 - net, True)
    di = torch.empty(2 * n, dtype=torch.int64, device=dev)
    do = torch.empty(2 * n, dtype=torch.int64, device=dev)
    rhj.fill_perm(di, 16, n, rank * n, G, seed=1, start_rid=rank * n)
    rhj.fill_perm(do, 16, n, (world - 1 - rank) * n, G, seed=2, start_rid=rank * n)
    ctx = rhj.Context(cfg, rank, world, rank)
    cap = int(n * 1.1) + 4096
    ctx.plan(n, n, cap, cap)
    rhj.hashjoin.exchange_windows(ctx)
    inner, outer = rhj.Relation(di, n, G), rhj.Relation(do, n, G)
    for _ in range(3):
        rhj.join_multi_rank(ctx, inner, outer)
    torch.cuda.synchronize()
    dist.barrier()

    ev = [torch.cuda.Event(enable_timing=True) for _ in range(9)]
    token = torch.zeros(1, dtype=torch.int32, device=dev)
    host_group = rhj.hashjoin._host_barrier_group(None)

    def barrier(which):
        if host_group is not None:
            ctx.exchange_sync(which)
            dist.barrier(group=host_group)
        else:
            ctx.exchange_wait(which)
            dist.all_reduce(token)

    ev[0].record()
    ctx.histogram(di, n, do, n)
    local = torch.as_tensor(rhj.hashjoin._DeviceArray(ctx.local_histogram_ptr(), 2 * ctx.P), device=dev)
    gathered = torch.empty(world * 2 * ctx.P, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(gathered, local)
    ctx.compute_offsets(gathered)
    ev[1].record()
    ctx.network_partition()
    ev[2].record()
    ev[3].record()
    barrier(0)
    ev[4].record()
    ctx.local_partition_rel(0)
    ev[5].record()
    ev[6].record()
    barrier(1)
    ctx.local_partition_rel(1)
    ev[7].record()
    ctx.build_probe()
    ev[8].record()
    torch.cuda.synchronize()
    names = ["hist+plan", "K2 R,S", "wait push R", "barrier R", "K3 R", "wait push S", "barrier+K3 S", "K4"]
    if rank == 0:
        print(" | ".join(f"{nm} {ev[i].elapsed_time(ev[i + 1]):.2f}" for i, nm in enumerate(names)),
              "| total", round(ev[0].elapsed_time(ev[8]), 2))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
