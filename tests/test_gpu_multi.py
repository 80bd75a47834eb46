"""Real multi-GPU path: one process per GPU (NCCL), CUDA-IPC windows, peer stores from the
scatter kernel.  Needs at least two GPUs (gpurun --gpus 2); skipped otherwise."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    """A port nobody listens on right now, away from the ephemeral range the OS hands out
    (a port obtained with bind(0) can be re-used by an outgoing connection before the
    rendezvous store binds it)."""
    import random
    for _ in range(64):
        p = random.randint(20000, 29999)
        s = socket.socket()
        try:
            s.bind(("127.0.0.1", p))
            return p
        except OSError:
            continue
        finally:
            s.close()
    raise RuntimeError("no free port")


def _worker(rank, world, port, per, knobs, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    inner, outer = O.generate_reference(world, per, per)
    cfg = rhj.make_config(knobs[0], knobs[1], knobs[2], bool(knobs[3]))
    ri = rhj.Relation.from_numpy(inner[rank], per * world, device=rank)
    ro = rhj.Relation.from_numpy(outer[rank], per * world, device=rank)
    hj = rhj.HashJoin(world, rank, ri, ro, config=cfg, device=rank, window_caps=(per * world, per * world))
    res = [hj.join().astuple() for _ in range(2)]  # twice: windows are reused
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), np.array(res, dtype=np.uint64))
    # per-rank histogram getters must agree with the all-gathered view
    np.save(os.path.join(out_dir, f"ghist{rank}.npy"), hj.ctx.get_global_histogram(0))
    dist.barrier()
    hj.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("knobs", [(5, 5, 27, 1), (8, 8, 24, 1), (5, 5, 27, 0)])
def test_two_gpus_nccl_ipc(oracle, tmp_path, knobs):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    world, per = 2, 150000 if knobs[3] else 20000
    mp.spawn(_worker, args=(world, _free_port(), per, knobs, str(tmp_path)), nprocs=world, join=True)
    cfg = oracle.Config(*knobs)
    inner, outer = oracle.generate_reference(world, per, per)
    want = oracle.join(inner, outer, cfg)
    glob = sum(oracle.local_histogram(t, cfg) for t in inner)
    for r in range(world):
        got = np.load(tmp_path / f"rank{r}.npy")
        for rep in range(2):
            assert tuple(int(x) for x in got[rep]) == want[r].astuple()
        assert np.array_equal(np.load(tmp_path / f"ghist{r}.npy"), glob)
