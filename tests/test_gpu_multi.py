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
    if len(knobs) > 4:  # the exchange of three and more ranks (send buffers + pushes) on two
        cfg.flags |= rhj.FLAG_STAGED_EXCHANGE
        os.environ["RHJ_PER_RELATION_EXCHANGE"] = "1"  # and with the inner relation's histogram exchanged first
        cfg.chunk_tuples = 8192  # many chunks: the relation is scattered and pushed in several groups
    ri = rhj.Relation.from_numpy(inner[rank], per * world, device=rank)
    ro = rhj.Relation.from_numpy(outer[rank], per * world, device=rank)
    cfg.flags |= rhj.FLAG_STAGE_HOST_INPUTS
    hj = rhj.HashJoin(world, rank, ri, ro, config=cfg, device=rank)  # windows sized from a histogram, as the reference does
    # host buffers through the C ABI (rhj_stage_host_inputs): the multi-rank form of rhj_join_host
    rhj.join_multi_rank(hj.ctx, ri, ro, host=(np.ascontiguousarray(inner[rank]), np.ascontiguousarray(outer[rank])))
    host_res = hj.ctx.result().astuple()
    assert hj.ctx.timings().h2d_us > 0
    res = [hj.join().astuple() for _ in range(2)]  # twice: windows are reused
    # the same join with the outer network pass beside the inner local pass (two streams, split SMs)
    rhj.join_multi_rank(hj.ctx, ri, ro, overlap=hj.ctx.sm_count() // 2)
    assert hj.ctx.result().astuple() == res[0]
    assert hj.ctx.timings().snetcompl_us > 0
    # third time with the output materialised: same result, pairs of this rank
    pairs = torch.full((2 * (res[0][0] + 8),), -1, dtype=torch.int64, device=f"cuda:{rank}")
    assert hj.join(pairs=pairs).astuple() == res[0]
    np.save(os.path.join(out_dir, f"pairs{rank}.npy"), pairs.cpu().numpy().view(np.uint64)[: 2 * res[0][0]].reshape(-1, 2))
    assert int((pairs[2 * res[0][0]:] != -1).sum()) == 0
    assert host_res == res[0]
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), np.array(res, dtype=np.uint64))
    # per-rank histogram getters must agree with the all-gathered view
    np.save(os.path.join(out_dir, f"ghist{rank}.npy"), hj.ctx.get_global_histogram(0))
    dist.barrier()
    hj.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("knobs", [(5, 5, 27, 1), (8, 8, 24, 1), (5, 5, 27, 0), (8, 8, 24, 1, "staged")])
def test_two_gpus_nccl_ipc(oracle, tmp_path, knobs):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    world, per = 2, (150000 if len(knobs) == 4 else 4600000) if knobs[3] else 20000
    mp.spawn(_worker, args=(world, _free_port(), per, knobs, str(tmp_path)), nprocs=world, join=True)
    cfg = oracle.Config(*knobs[:4])
    inner, outer = oracle.generate_reference(world, per, per)
    want, want_pairs = oracle.join_pairs(inner, outer, cfg, world * per)
    glob = sum(oracle.local_histogram(t, cfg) for t in inner)
    for r in range(world):
        gp, wp = np.load(tmp_path / f"pairs{r}.npy"), want_pairs[r]
        assert np.array_equal(gp[np.lexsort((gp[:, 1], gp[:, 0]))], wp[np.lexsort((wp[:, 1], wp[:, 0]))])
        got = np.load(tmp_path / f"rank{r}.npy")
        for rep in range(2):
            assert tuple(int(x) for x in got[rep]) == want[r].astuple()
        assert np.array_equal(np.load(tmp_path / f"ghist{r}.npy"), glob)


RADIX_BITS = {2: (8, 9), 4: (9, 9), 8: (9, 10)}  # as bench.py: sub-partitions stay ~4 K tuples


def _worker_config(rank, world, port, shape, log2n, out_dir):
    """BASELINE configs 4 and 5 at their full per-GPU size, generated on the device (no host copy of
    the relations exists), checked through closed forms that do not depend on the size."""
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")
    from tools import checksum as CK
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    net, loc = RADIX_BITS[world]
    m64 = (1 << 64) - 1
    if shape == "c5":  # 1:16, packed 8-byte tuples, every build key matched 16 times
        ns = 1 << log2n
        nr = ns >> 4
        NR, NS = nr * world, ns * world
        cfg = rhj.make_config(net, loc - 2, 32 - net, True, tuple_bytes=8)
        dr = torch.empty(nr, dtype=torch.int64, device=dev)
        ds = torch.empty(ns, dtype=torch.int64, device=dev)
        rhj.fill_perm(dr, 8, nr, rank * nr, NR, seed=31, start_rid=rank * nr)
        rhj.fill_perm(ds, 8, ns, rank * ns, NS, seed=32, modulo=NR, start_rid=rank * ns)
        closed = (NS, 16 * (NR * (NR - 1) // 2) & m64, 16 * (NR * (NR - 1) // 2) & m64, NS * (NS - 1) // 2 & m64)
        # all five words, pairing included: unpack to {key, rid} pairs for the device-side checker
        un = lambda t: torch.stack([t >> 32, t & 0xFFFFFFFF], dim=1).reshape(-1)
        want = CK.expected_result(torch, dist, un(dr), nr, un(ds), ns, NR, world, rank, dev)
        assert tuple(want[:4]) == closed, (want, closed)
        tb, caps = 8, None
    else:  # "c4": Zipf(1)-like probe keys (P(k) ~ 1/k), unique build keys covering the whole domain;
        # "c4_balanced": the same with RHJ_FLAG_BALANCED_ASSIGNMENT
        n = 1 << log2n
        nr = ns = n
        NR = NS = n * world
        cfg = rhj.make_config(net, loc, 32 - net, True,
                              flags=rhj.FLAG_BALANCED_ASSIGNMENT if shape == "c4_balanced" else 0)
        dr = torch.empty(2 * n, dtype=torch.int64, device=dev)
        ds = torch.empty(2 * n, dtype=torch.int64, device=dev)
        rhj.fill_perm(dr, 16, n, rank * n, NR, seed=41, start_rid=rank * n)
        g = torch.Generator(device=dev)
        g.manual_seed(1000 + rank)
        keys = ds.view(-1, 2)[:, 0]
        step = 1 << 24
        for b in range(0, n, step):
            u = torch.rand(min(step, n - b), device=dev, dtype=torch.float64, generator=g)
            keys[b:b + step] = torch.clamp(torch.floor(torch.exp(u * float(np.log(NR)))).to(torch.int64) - 1, 0, NR - 1)
        ds.view(-1, 2)[:, 1] = torch.arange(rank * n, (rank + 1) * n, device=dev)
        # the ranks go through a fixed bijection, so that the hot keys are spread over the
        # partitions and ranks instead of being keys 0, 1, 2, ... (oracle orc_fill_zipf does the same)
        rhj.apply_perm(ds, n, 2, NR, seed=4177)
        want = CK.expected_result(torch, dist, dr, n, ds, n, NR, world, rank, dev)
        assert want[0] == NS and want[3] == NS * (NS - 1) // 2 & m64
        tb, caps = 16, None  # exact window sizes from a histogram: skew needs no guess
    ri = rhj.Relation(dr, nr, NR, tb)
    ro = rhj.Relation(ds, ns, NS, tb)
    hj = rhj.HashJoin(world, rank, ri, ro, config=cfg, device=rank, window_caps=caps)
    hj.join()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    e0.record()
    r = hj.join()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    got = CK.reduce_result(torch, dist, r, world, dev)
    share = torch.tensor([r.matches], dtype=torch.int64, device=dev)
    shares = [torch.zeros_like(share) for _ in range(world)]
    dist.all_gather(shares, share)
    tm = hj.ctx.timings().asdict()
    if rank == 0:
        import json
        with open(os.path.join(out_dir, "result.json"), "w") as f:
            json.dump({"shape": shape, "world": world, "got": list(got), "want": list(want), "ms": float(ms.item()),
                       "words": ["matches", "sum_key", "sum_rid_r", "sum_rid_s", "sum_mix"], "stages_us_rank0": tm,
                       "tuples_per_s": (NR + NS) / float(ms.item()) * 1e3,
                       "matches_per_rank": [int(s.item()) for s in shares]}, f)
    dist.barrier()
    hj.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("shape", ["c5", "c4", "c4_balanced"])
def test_baseline_configs_4_and_5_full_size_all_gpus(tmp_path, shape):
    """256 M probe tuples per GPU on every GPU of the box (2, 4 or 8)."""
    import json
    import torch
    world = torch.cuda.device_count()
    world = 8 if world >= 8 else 4 if world >= 4 else world
    if world < 2:
        pytest.skip("needs at least two GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_worker_config, args=(world, _free_port(), shape, 28, str(tmp_path)), nprocs=world, join=True)
    res = json.load(open(tmp_path / "result.json"))
    print("\n[config]", json.dumps(res))
    keep = os.environ.get("RHJ_KEEP_RESULTS")
    if keep:
        os.makedirs(keep, exist_ok=True)
        json.dump(res, open(os.path.join(keep, f"config_{shape}_n{world}.json"), "w"))
    assert len(res["want"]) == 5 and None not in res["want"]
    assert res["got"] == res["want"]


def _worker_distribute(rank, world, port, per, seed, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    t = torch.empty(2 * per, dtype=torch.int64, device=dev)
    rhj.fill_perm(t, 16, per, rank * per, world * per, seed=11, start_rid=rank * per)  # generated on the device
    rhj.distribute(rhj.Relation(t, per, world * per), world, rank, seed)               # exchanged over NVLink
    torch.cuda.synchronize()
    np.save(os.path.join(out_dir, f"dist{rank}.npy"), t.cpu().numpy().view(np.uint64).reshape(-1, 2))
    dist.barrier()
    dist.destroy_process_group()


def test_device_distribute_two_gpus(oracle, tmp_path):
    """generator + Relation::distribute without a host copy of the relation, equal to
    orc_distribute_perm bit for bit (the same launcher code is checked over gloo in test_multirank_cpu.py)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    world, per, seed = 2, 1000003, 77
    mp.spawn(_worker_distribute, args=(world, _free_port(), per, seed, str(tmp_path)), nprocs=world, join=True)
    rels = [oracle.fill_perm(per, r * per, world * per, seed=11, start_rid=r * per) for r in range(world)]
    oracle.distribute_perm(rels, seed)
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / f"dist{r}.npy"), rels[r].view(np.uint64).reshape(-1, 2))


def _worker_pieces(rank, world, port, per_r, per_s, piece, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    inner, outer = O.generate_reference(world, per_r, per_s, gen="modulo")
    cfg = rhj.make_config(6, 5, 26, True, chunk_tuples=8192)
    ctx = rhj.Context(cfg, rank, world, rank)
    ctx.plan(per_r, piece, per_r * world, piece * world)  # windows for one piece per rank, whatever its keys
    rhj.hashjoin.exchange_windows(ctx)
    rel = rhj.Relation.from_numpy(inner[rank], per_r * world, device=rank)
    res = [rhj.join_probe_pieces(ctx, rel, rhj.host_pieces(outer[rank], piece, f"cuda:{rank}")).astuple()
           for _ in range(2)]
    np.save(os.path.join(out_dir, f"pieces{rank}.npy"), np.array(res, dtype=np.uint64))
    dist.barrier()
    ctx.close()
    dist.destroy_process_group()


def test_probe_side_in_pieces_two_gpus(oracle, tmp_path):
    """SURVEY.md 8(f4) on two ranks: every rank streams its share of an outer relation eight times the inner
    one from host memory in pieces; each rank's result equals the oracle's for the whole relations."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    world, per_r, per_s, piece = 2, 30000, 240000, 32768
    mp.spawn(_worker_pieces, args=(world, _free_port(), per_r, per_s, piece, str(tmp_path)), nprocs=world, join=True)
    inner, outer = oracle.generate_reference(world, per_r, per_s, gen="modulo")
    want = oracle.join(inner, outer, oracle.Config(6, 5, 26, 1))
    for r in range(world):
        got = np.load(tmp_path / f"pieces{r}.npy")
        for rep in range(2):
            assert tuple(int(x) for x in got[rep]) == want[r].astuple()


def _worker_hot_split(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    inner = O.fill_perm(60000, rank * 60000, 60000 * world, seed=9)
    outer = O.fill_zipf(300000, rank * 300000, 60000 * world, 1.0, seed=20 + rank)
    cfg = rhj.make_config(6, 4, 26, True)
    ri = rhj.Relation.from_numpy(inner, 60000 * world, device=rank)
    ro = rhj.Relation.from_numpy(outer, 300000 * world, device=rank)
    hj = rhj.HotSplitJoin(world, rank, ri, ro, config=cfg, device=rank, inv_share=16)
    res = [hj.join().astuple() for _ in range(2)]
    plain = rhj.HashJoin(world, rank, ri, ro, config=cfg, device=rank)
    res.append(plain.join().astuple())
    np.save(os.path.join(out_dir, f"hot{rank}.npy"), np.array(res, dtype=np.uint64))
    dist.barrier()
    hj.close()
    plain.close()
    dist.destroy_process_group()


def test_hot_partition_splitting_two_gpus(oracle, tmp_path):
    """SURVEY.md 8(f3), second half: per-rank results under hot-partition splitting equal the oracle-side
    statement of the rule; the sum equals the plain join's; the busiest rank's matches drop."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker_hot_split, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    inner = [oracle.fill_perm(60000, r * 60000, 60000 * world, seed=9) for r in range(world)]
    outer = [oracle.fill_zipf(300000, r * 300000, 60000 * world, 1.0, seed=20 + r) for r in range(world)]
    cfg = oracle.Config(6, 4, 26, 1)
    want, hot = oracle.join_hot_split(inner, outer, cfg, 16)
    assert hot.any() and not hot.all()
    plain = oracle.join(inner, outer, cfg)
    got = [np.load(tmp_path / f"hot{r}.npy") for r in range(world)]
    for r in range(world):
        for rep in range(2):
            assert tuple(int(x) for x in got[r][rep]) == want[r]
        assert tuple(int(x) for x in got[r][2]) == plain[r].astuple()
    assert max(w[0] for w in want) < max(p.matches for p in plain)
