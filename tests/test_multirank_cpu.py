"""N>1 host logic on CPU: two processes over torch.distributed/gloo do what the ranks of a
multi-GPU join do between kernels -- all-gather the per-rank histograms, derive global
histogram / assignment / offsets from the gathered matrix -- and must agree with the oracle's
(and hence the reference's MPI_Allreduce + MPI_Exscan) result."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

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
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = O.Config(*knobs)
    P = 1 << cfg.net_bits
    inner, outer = O.generate_reference(world, per, per)  # every rank regenerates; uses only its share
    local = np.concatenate([O.local_histogram(inner[rank], cfg), O.local_histogram(outer[rank], cfg)])
    t = torch.from_numpy(local.view(np.int64).copy())
    gathered = torch.empty(world * 2 * P, dtype=torch.int64)
    dist.all_gather_into_tensor(gathered, t)
    mat = gathered.numpy().view(np.uint64).reshape(world, 2, P)
    L = rhj.lib()
    res = {}
    for which, name in ((0, "inner"), (1, "outer")):
        loc = np.ascontiguousarray(mat[:, which, :])
        g, b, rel, ab = (np.zeros(P, np.uint64) for _ in range(4))
        ws = np.zeros(world, np.uint64)
        L.rhj_host_offsets(loc.ctypes.data, world, rank, P, g.ctypes.data, b.ctypes.data, rel.ctypes.data,
                           ab.ctypes.data, ws.ctypes.data)
        res[name + "_global"], res[name + "_base"], res[name + "_abs"], res[name + "_ws"] = g, b, ab, ws
    # barrier + "result aggregation" as HashJoin/Measurements do (sum of per-rank counts)
    mine = O.join(inner, outer, cfg)[rank].matches
    tot = torch.tensor([mine], dtype=torch.int64)
    dist.all_reduce(tot)
    res["total"] = np.array([int(tot.item())], np.uint64)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), **res)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("knobs", [(5, 5, 27, 1), (8, 8, 24, 1)])
def test_two_ranks_gloo_offsets(oracle, tmp_path, knobs):
    import torch.multiprocessing as mp
    world, per = 2, 30000
    port = _free_port()
    mp.spawn(_worker, args=(world, port, per, knobs, str(tmp_path)), nprocs=world, join=True)
    cfg = oracle.Config(*knobs)
    inner, outer = oracle.generate_reference(world, per, per)
    for name, rel in (("inner", inner), ("outer", outer)):
        pl = oracle.plan(np.stack([oracle.local_histogram(t, cfg) for t in rel]), world)
        for r in range(world):
            z = np.load(tmp_path / f"rank{r}.npz")
            assert np.array_equal(z[name + "_global"], pl.global_)
            assert np.array_equal(z[name + "_base"], pl.base)
            assert np.array_equal(z[name + "_abs"], pl.abs_[r])
            assert np.array_equal(z[name + "_ws"], pl.window_size)
            assert int(z["total"][0]) == world * per


def _worker_distribute(rank, world, port, per, seed, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # main.cpp:94-98 with the harness generator: rank r fills its own key range, then distribute()
    mine = O.fill_perm(per, rank * per, world * per, seed=11, start_rid=rank * per)
    t = torch.from_numpy(mine.view(np.int64).copy())
    rel = rhj.Relation(t, per, world * per)
    rhj.distribute(rel, world, rank, seed, None)
    np.save(os.path.join(out_dir, f"dist{rank}.npy"), t.numpy().view(np.uint64).reshape(-1, 2))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,per", [(2, 10001), (3, 7000), (4, 4099)])
def test_device_agnostic_distribute_equals_oracle(oracle, tmp_path, world, per):
    """The launchers' Relation::distribute (torch point-to-point exchange + position bijection; the same code
    runs on NCCL with device tensors) against orc_distribute_perm, bit for bit, ragged last section included."""
    import torch.multiprocessing as mp
    seed = 77
    mp.spawn(_worker_distribute, args=(world, _free_port(), per, seed, str(tmp_path)), nprocs=world, join=True)
    rels = [oracle.fill_perm(per, r * per, world * per, seed=11, start_rid=r * per) for r in range(world)]
    before = np.sort(np.concatenate(rels).view(np.uint64).reshape(-1, 2)[:, 0])
    oracle.distribute_perm(rels, seed)
    for r in range(world):
        got = np.load(tmp_path / f"dist{r}.npy")
        assert np.array_equal(got, rels[r].view(np.uint64).reshape(-1, 2)), f"rank {r}"
    after = np.sort(np.concatenate([np.load(tmp_path / f"dist{r}.npy")[:, 0] for r in range(world)]))
    assert np.array_equal(before, after)  # a redistribution: the global key multiset is unchanged


def test_perm_positions_equals_oracle(oracle):
    import torch
    rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")
    for n, seed in ((1, 3), (2, 5), (1000, 7), (65537, 99), (100000, 123456789)):
        got = rhj.hashjoin.perm_positions(n, seed, "cpu").numpy()
        want = np.array([oracle.lib().orc_perm(i, n, seed) for i in range(n)], np.int64)
        assert np.array_equal(got, want), (n, seed)
        assert np.array_equal(np.sort(got), np.arange(n))


def _worker_hot(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")
    H = rhj.hashjoin
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    P = 64
    inner = O.fill_perm(6000, rank * 6000, 6000 * world, seed=9)
    outer = O.fill_zipf(20000, rank * 20000, 6000 * world, 1.0, seed=10 + rank)
    ri = rhj.Relation(torch.from_numpy(inner.view(np.int64).copy()), len(inner), 6000 * world)
    ro = rhj.Relation(torch.from_numpy(outer.view(np.int64).copy()), len(outer), 20000 * world)
    h = torch.bincount(ro.data[::2] & (P - 1), minlength=P)
    dist.all_reduce(h)
    hot = H.hot_partitions(h, world, 8)
    s_hot, s_cold = H.split_hot(ro, hot, P)
    r_hot, r_cold = H.split_hot(ri, hot, P)
    r_all = H.gather_rows(r_hot, world)
    np.savez(os.path.join(out_dir, f"hot{rank}.npz"), hot=hot.numpy(), s_hot=s_hot.numpy(), s_cold=s_cold.numpy(),
             r_cold=r_cold.numpy(), r_all=r_all.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_hot_partition_rule_and_split_plumbing(oracle, tmp_path):
    """hot_partitions / split_hot / gather_rows (launcher plumbing of HotSplitJoin) over gloo against the
    oracle-side statement of the rule (oracle.join_hot_split)."""
    import torch.multiprocessing as mp
    world, P = 3, 64
    mp.spawn(_worker_hot, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    inner = [oracle.fill_perm(6000, r * 6000, 6000 * world, seed=9) for r in range(world)]
    outer = [oracle.fill_zipf(20000, r * 20000, 6000 * world, 1.0, seed=10 + r) for r in range(world)]
    _, hot = oracle.join_hot_split(inner, outer, oracle.Config(6, 3, 26, 1), 8)
    assert hot.any() and not hot.all()
    pid = lambda t: (t["key"] & np.uint64(P - 1)).astype(np.int64)
    want_r_all = np.concatenate([t[hot[pid(t)]] for t in inner]).view(np.int64).reshape(-1, 2)
    for r in range(world):
        z = np.load(tmp_path / f"hot{r}.npz")
        assert np.array_equal(z["hot"], hot)
        assert np.array_equal(z["r_all"], want_r_all)
        assert np.array_equal(z["s_hot"], outer[r][hot[pid(outer[r])]].view(np.int64).reshape(-1, 2))
        assert np.array_equal(z["s_cold"], outer[r][~hot[pid(outer[r])]].view(np.int64).reshape(-1, 2))
        assert np.array_equal(z["r_cold"], inner[r][~hot[pid(inner[r])]].view(np.int64).reshape(-1, 2))


def _worker_caps(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    rhj = importlib.import_module("distributed-radxi-hash-join-on-gpus_b200")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    inner = O.fill_perm(5000 + 100 * rank, rank * 6000, 6000 * world, seed=9)
    outer = O.fill_zipf(20000, rank * 20000, 6000 * world, 1.0, seed=10 + rank)
    out = {}
    for name, flags in (("plain", 0), ("balanced", rhj.FLAG_BALANCED_ASSIGNMENT)):
        cfg = rhj.make_config(5, 4, 27, True, flags=flags)
        ri = rhj.Relation(torch.from_numpy(inner.view(np.int64).copy()), len(inner), 6000 * world)
        ro = rhj.Relation(torch.from_numpy(outer.view(np.int64).copy()), len(outer), 20000 * world)
        out[name] = np.array(rhj.hashjoin.exact_window_caps(cfg, world, rank, ri, ro), np.uint64)
    np.savez(os.path.join(out_dir, f"caps{rank}.npz"), **out)
    dist.barrier()
    dist.destroy_process_group()


def test_exact_window_caps_equal_the_reference_window_sizes(oracle, tmp_path):
    """The launchers size the receive windows as the reference does (data/Window.cpp:160-178): from the
    histograms.  exact_window_caps over gloo against the oracle's plan, p % n and size-balanced assignment."""
    import torch.multiprocessing as mp
    world = 3
    mp.spawn(_worker_caps, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    inner = [oracle.fill_perm(5000 + 100 * r, r * 6000, 6000 * world, seed=9) for r in range(world)]
    outer = [oracle.fill_zipf(20000, r * 20000, 6000 * world, 1.0, seed=10 + r) for r in range(world)]
    cfg = oracle.Config(5, 4, 27, 1)
    li = np.stack([oracle.local_histogram(t, cfg) for t in inner])
    lo = np.stack([oracle.local_histogram(t, cfg) for t in outer])
    bal = oracle.balanced_assignment(li.sum(axis=0, dtype=np.uint64), lo.sum(axis=0, dtype=np.uint64), world)
    for name, assign in (("plain", None), ("balanced", bal)):
        want = (int(oracle.plan(li, world, assign).window_size.max()) + 64,
                int(oracle.plan(lo, world, assign).window_size.max()) + 64)
        for r in range(world):
            got = np.load(tmp_path / f"caps{r}.npz")[name]
            assert (int(got[0]), int(got[1])) == want, (name, r)
