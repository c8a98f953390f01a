"""Worker of tests/test_sharded_gpu.py: one process per GPU under torchrun.  Every rank builds the sharded handle and, on
its own GPU, a one-rank handle of the same problem; the sharded results (owned rows) must equal the one-rank results.
Covers what only runs with nranks > 1: the NCCL all-gather / grouped broadcast of the source, the sharded upward pass with
its grouped all-gather of multipoles, cross-rank operator pairs under symmetric storage, the all-reduces of the fused CGS2
passes inside the pipelined GMRES, the host API (1/nranks upload + all-gather, strided read-back of owned rows) and the
block operator on d_blk_full."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402


def main():
    import ctypes

    import torch
    import torch.distributed as dist

    from aniso_b200 import Aniso, _lib

    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = _lib.load()

    def new_uid():
        buf = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            raw = ctypes.create_string_buffer(128)
            _lib.check(L.aniso_nccl_unique_id(raw))
            buf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
        buf = buf.cuda()
        dist.broadcast(buf, 0)
        return bytes(buf.cpu().numpy().tobytes())

    def rel(a, b):
        return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))

    names = sys.argv[1:] or ["c2", "nu6d5", "h32", "k3h16"]
    for name in names:
        cfg = cases.CASES.get(name) or cases.BLOCK_CASES[name]
        sz, deg, K, g, sr, np_, ml, kind = cfg
        a = Aniso(sz, deg, K, g, sr, np_, ml, nranks=world, rank=rank, nccl_uid=new_uid(), device=local)
        s = Aniso(sz, deg, K, g, sr, np_, ml)
        x, y, w = a.nodes()
        n = x.size
        ss, st = cases.sigma_fields(kind, x, y, sz, deg)
        for o in (a, s):
            o.setCoeff(ss, st)
        slot0, slot1 = a.owned_range()
        perm = a.slot_permutation()
        own = perm[slot0:slot1]  # reference indices of the owned rows
        f = cases.random_source(n, 11) * cases.gaussian_source(x, y) + 0.1 * cases.tilted_source(x, y)
        kids = sorted({0, min(1, a.kernelSize - 1), a.kernelSize - 1})
        for kid in kids:
            a.cache(kid)
            s.cache(kid)
            got, want = a.mapping(f, kid), s.mapping(f, kid)  # host API: owned rows filled, the others untouched (zero)
            assert rel(got[own], want[own]) < 1e-12, (name, kid, "mapping", rel(got[own], want[own]))
            other = np.setdiff1d(np.arange(n), own)
            assert not got[other].any(), (name, kid, "rows of other ranks were written")
        # device-resident operator on owned shards (the GMRES operator; sharded upward pass where the tree allows)
        xs = torch.from_numpy(f[perm][slot0:slot1].copy()).cuda()
        ys = torch.empty_like(xs)
        _lib.check(L.aniso_forward_dev(a.h, xs.data_ptr(), ys.data_ptr()))
        a.synchronize()
        want = f - s.mapping(ss * f, 0)
        assert rel(ys.cpu().numpy(), want[own]) < 1e-12, (name, "forward_dev", rel(ys.cpu().numpy(), want[own]))
        # GMRES (pipelined Arnoldi, all-reduced CGS2 projections): same iteration count, same solution
        b = s.mapping(cases.gaussian_source(x, y), 0)
        x1, rc1, it1, h1 = s.gmres(b)
        x2, rc2, it2, h2 = a.gmres(b)
        assert rc1 == rc2 == 0 and it1 == it2, (name, "gmres iterations", it1, it2)
        assert rel(x2[own], x1[own]) < 1e-9, (name, "gmres solution", rel(x2[own], x1[own]))
        np.testing.assert_allclose(h2[: it1 - 2], h1[: it1 - 2], rtol=1e-3, atol=1e-10)
        for form in ("mgs",):  # the reference's MGS: i+2 dependent all-reduces per step
            os.environ["ANISO_GMRES_ORTHO"] = form
            x3, rc3, it3, _ = a.gmres(b)
            os.environ.pop("ANISO_GMRES_ORTHO")
            assert rc3 == 0 and it3 == it1 and rel(x3[own], x1[own]) < 1e-9, (name, form, it3, it1)
        if K > 1:  # multi-mode block operator (aniso.m:121-156) on the ranks
            for kid in range(a.kernelSize):
                a.cache(kid)
                s.cache(kid)
            u = cases.random_source(K * n, 5)
            idx = np.concatenate([own + m * n for m in range(K)])
            for fn in ("forward", "mforward"):
                got, want = getattr(a, fn)(u), getattr(s, fn)(u)
                assert rel(got[idx], want[idx]) < 1e-11, (name, fn, rel(got[idx], want[idx]))
        cfgline = a.config()
        a.delete()
        s.delete()
        dist.barrier()
        if rank == 0:
            print("case %s ok on %d ranks (%s)" % (name, world, cfgline), flush=True)
    print("SHARDED_OK rank %d" % rank, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
