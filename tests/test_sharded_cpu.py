"""World-size-2 gloo test of the host-side logic of the multi-GPU path, runnable without a GPU:
every rank builds its sharded handle (host plan only), the ranks' slot ranges tile [0, N), the per-rank operator bytes
add up to the single-device total plus the replicated ancestor M2L rows, and the exchange step the device path
performs per matvec (all-gather of unequal owned shards into a full slot-order vector) is replayed over gloo with the
CPU oracle standing in for the per-rank row computation (the oracle is the checker here, not the product)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import cases

WORLD = 2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, port, name, out_q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        from aniso_b200 import Aniso
        from oracle import oracleapi

        sz, deg, K, g, sr, np_, ml, kind = cases.CASES[name]
        uid = bytes(128)  # NCCL is only initialised on first device use; the host plan does not need it
        a = Aniso(sz, deg, K, g, sr, np_, ml, nranks=WORLD, rank=rank, nccl_uid=uid, device=0)
        single = Aniso(sz, deg, K, g, sr, np_, ml)
        n = a.numberOfNodes
        b, e = a.owned_range()
        assert (b, e) == a.partition(WORLD, rank)[:2] == single.partition(WORLD, rank)[:2]
        # ranges of all ranks tile [0, n)
        rng = torch.tensor([b, e])
        allr = [torch.zeros(2, dtype=torch.long) for _ in range(WORLD)]
        dist.all_gather(allr, rng)
        assert int(allr[0][0]) == 0 and int(allr[-1][1]) == n
        for r in range(WORLD - 1):
            assert int(allr[r][1]) == int(allr[r + 1][0])
        # operator bytes: every rank stores the blocks its own targets need.  Under symmetric storage a pair whose two
        # leaves sit on different ranks is kept by both (neither can write into the other's scratch rows), so the
        # sum lies between the single-device (half-stored) total and the unsymmetrised one; M2L rows of shared
        # ancestors are replicated on top of that
        mine = a.apply_bytes()
        tot = torch.tensor([mine["p2p"], mine["m2l"]], dtype=torch.long)
        dist.all_reduce(tot)
        full = single.apply_bytes()
        assert full["p2p"] <= int(tot[0]) <= 2 * full["p2p"]
        assert full["m2l"] <= int(tot[1]) <= 2.04 * full["m2l"] + 64 * 8 * (np_ ** 4) * 30
        # replay of the per-matvec exchange: rows owned by this rank -> all-gather (unequal counts) -> full vector
        perm = a.slot_permutation()
        assert np.array_equal(perm, single.slot_permutation())
        o = oracleapi.OracleAniso(sz, deg, K, g, sr, np_, ml)
        x, y, w = o.nodes()
        ss, st = cases.sigma_fields(kind, x, y, sz, deg)
        o.set_coeff(ss, st)
        f = cases.random_source(n, 11)
        y_full_slots = o.mapping(f, 0)[perm]
        shard = torch.from_numpy(y_full_slots[b:e].copy())
        sizes = [int(r[1] - r[0]) for r in allr]
        pad = max(sizes)
        buf = torch.zeros(pad, dtype=torch.float64)
        buf[: e - b] = shard
        got = [torch.zeros(pad, dtype=torch.float64) for _ in range(WORLD)]
        dist.all_gather(got, buf)
        rebuilt = np.concatenate([got[r][: sizes[r]].numpy() for r in range(WORLD)])
        assert np.array_equal(rebuilt, y_full_slots)
        # a dot product reduced in fixed rank order gives identical bits on every rank
        part = torch.tensor([float(np.dot(y_full_slots[b:e], y_full_slots[b:e]))], dtype=torch.float64)
        parts = [torch.zeros(1, dtype=torch.float64) for _ in range(WORLD)]
        dist.all_gather(parts, part)
        total = sum(float(p) for p in parts)
        tt = torch.tensor([total], dtype=torch.float64)
        tts = [torch.zeros(1, dtype=torch.float64) for _ in range(WORLD)]
        dist.all_gather(tts, tt)
        assert all(float(t) == total for t in tts)
        a.delete()
        single.delete()
        out_q.put((rank, "ok"))
    except Exception as ex:  # surface the failure in the parent
        out_q.put((rank, repr(ex)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["a16", "nu6d5"])
def test_world2_partition_and_exchange(name):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, port, name, q)) for r in range(WORLD)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(WORLD)]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
