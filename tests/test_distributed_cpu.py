"""World-size-2 gloo run of the prefix-sharded path (dekupl-run_b200/distributed.py) on CPU.
The per-rank compute is an ORACLE-backed backend (test infrastructure), so what is under test is
the sharding and collective logic: R/U all-reduce, NF broadcast (reference mode), radix-select
histogram all-reduces (full mode), BH candidate all-gather, survivor all-reduce. The sharded result
must equal the oracle run on the unsharded input."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def _image(v):
    b = v.view(np.uint64)
    return np.where(b >> np.uint64(63), ~b, b | np.uint64(1 << 63))


def _unimage(i):
    i = np.uint64(i)
    b = (i & np.uint64((1 << 63) - 1)) if (i >> np.uint64(63)) else ~i
    return np.array([b], dtype=np.uint64).view(np.float64)[0]


class OracleBackend:
    """same interface as distributed.GpuBackend, computed by the CPU oracle"""

    def __init__(self):
        import oracle_lib as orc
        self.orc = orc

    def join(self, keys, cnts, n, r, a):
        self.mk, self.mc, U = self.orc.join(keys, cnts, r, a)
        return len(self.mk), U

    def nf_reference(self, step_rows=0):
        nf, used, uz = self.orc.nf_reference(self.mc, step_rows or 1000000)
        return nf, dict(consumed=used, steps=0, undefined_zone=uz, n_pos=0)

    def nf_begin(self):
        pos = (self.mc > 0).all(axis=1)
        c = self.mc[pos].astype(np.float64)
        S = self.mc.shape[1]
        lg = np.log(c)
        m = np.zeros(len(c))
        for j in range(S):                      # sequential sum in sample order (computeNF.c:21-23)
            m = m + lg[:, j]
        m = m / S
        self.img = _image(np.ascontiguousarray((lg - m[:, None]).T))   # [S][npos]
        self.S = S
        return int(pos.sum())

    def nf_set_rank(self, k):
        self.k = np.full(self.S, k, dtype=np.uint64)
        self.prefix = np.zeros(self.S, dtype=np.uint64)
        self.shift = 56

    def nf_hist(self):
        h = np.zeros((self.S, 256), dtype=np.int64)
        for s in range(self.S):
            x = self.img[s]
            if self.shift + 8 < 64:
                x = x[(x >> np.uint64(self.shift + 8)) == (self.prefix[s] >> np.uint64(self.shift + 8))]
            h[s] = np.bincount(((x >> np.uint64(self.shift)) & np.uint64(255)).astype(np.int64), minlength=256)
        return torch.from_numpy(h.reshape(-1))

    def nf_advance(self, hist):
        h = hist.numpy().reshape(self.S, 256)
        for s in range(self.S):
            acc, d = 0, 0
            while d < 255 and acc + h[s, d] <= int(self.k[s]):
                acc += int(h[s, d]); d += 1
            self.prefix[s] |= np.uint64(d << self.shift)
            self.k[s] -= np.uint64(acc)
        self.shift -= 8

    def nf_result(self):
        return np.exp(np.array([_unimage(p) for p in self.prefix]))

    def ttest_rows(self, nf, idxA, idxB):
        self.p, _, _, self.fc = self.orc.ttest_rows(self.mc, nf, idxA, idxB)

    device = torch.device("cpu")

    def count_le(self, cut):
        return int((self.p <= cut).sum())

    def next_cut(self, p_cut, n_global, R_global):                 # the rule of dk_bh_next_cut (csrc/dk_bh.cu)
        if not (p_cut < 1.0) or R_global == 0:
            return p_cut
        return p_cut * (n_global / R_global) * (1.0 + 1e-8) + 1e-300

    def candidates(self, p_cut, cut=None):
        self.cand = np.nonzero(self.p <= (p_cut if cut is None else min(cut, p_cut)))[0]
        self.narrowed_to = cut
        return torch.from_numpy(self.p[self.cand].copy())

    def adjust_global(self, all_p, R_global):
        self.adj = np.full(len(self.p), np.nan)
        if all_p.numel() == 0:
            return
        sp = np.sort(all_p.numpy())
        j = np.arange(len(sp))
        last = np.r_[sp[1:] != sp[:-1], True]
        v = np.where(last, (float(R_global) / (j + 1.0)) * sp, np.inf)          # TtestFilter.c:467
        F = np.minimum.accumulate(v[::-1])[::-1]
        pos = np.searchsorted(sp, self.p[self.cand], side="left")
        self.adj[self.cand] = np.minimum(1.0, F[pos])

    def bh_rank_slice(self, p, n_below, R_global):
        sp = np.sort(p.numpy())
        j = np.arange(len(sp))
        last = np.r_[sp[1:] != sp[:-1], True] if len(sp) else np.zeros(0, bool)
        v = np.where(last, (float(R_global) / (n_below + j + 1.0)) * sp, np.inf)
        F = np.minimum.accumulate(v[::-1])[::-1].copy() if len(sp) else v
        return torch.from_numpy(sp), torch.from_numpy(F)

    def apply_adjusted(self, sp, F):
        self.adj = np.full(len(self.p), np.nan)
        if sp.numel():
            pos = np.searchsorted(sp.numpy(), self.p[self.cand], side="left")
            self.adj[self.cand] = np.minimum(1.0, F.numpy()[pos])

    def filter(self, pmax, fmin):
        keep = (self.adj <= pmax) & (np.abs(self.fc) >= fmin)
        self.keep = np.nonzero(keep)[0]
        return int(keep.sum()), (0, 0, 0)

    def free(self):
        pass


def _worker(rank, world, port, case, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from dekupl_run_b200 import distributed as D
        keys, cnts, edges = case["keys"], case["cnts"], case["edges"]
        lo, hi = edges[rank], edges[rank + 1]
        sk, sc = [], []
        for k, c in zip(keys, cnts):
            a, b = np.searchsorted(k, lo), np.searchsorted(k, hi)
            sk.append(k[a:b]); sc.append(c[a:b])
        be = OracleBackend()
        st = D.run_sharded(be, sk, sc, [len(k) for k in sk], case["r"], case["a"], case["idxA"], case["idxB"], case["pmax"], 0.0,
                           nf_mode=case["nf_mode"], nf_step=case["step"], device=torch.device("cpu"))
        q.put((rank, st["U"], st["R"], st["D"], np.asarray(st["nf"]).tolist(), be.mk[be.keep].tolist(), be.adj[be.keep].tolist()))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


@pytest.mark.parametrize("nf_mode", ["reference", "full"])
def test_two_rank_sharded_path_equals_unsharded_oracle(nf_mode):
    import oracle_lib as orc
    from common import random_dumps
    rng = np.random.default_rng(7)
    S, nA = 6, 3
    keys, cnts = random_dumps(rng, S, 60_000, 40, 0.15, 80, p_core=0.4)
    # counts = per-k-mer abundance x per-sample depth (+ small jitter): the size factors converge within a few
    # steps, so computeNF's early stop (computeNF.c:145-152) fires inside rank 0's shard in reference mode
    depth = [1.0, 1.1, 0.9, 1.3, 0.8, 1.2]
    for s in range(S):
        b = (keys[s] % np.uint64(2000)).astype(np.int64) + 2000
        cnts[s] = (b * depth[s] + rng.integers(0, 40, size=len(b))).astype(np.int32)
    for c in cnts[nA:]:                       # an effect on group B for the first 2% of each dump
        c[: len(c) // 50] *= 8
    idxA, idxB = list(range(nA)), list(range(nA, S))
    r, a, pmax, step = 2, 3, 0.2, 2000
    mk, mc, U = orc.join(keys, cnts, r, a)
    nf = orc.nf_reference(mc, step)[0] if nf_mode == "reference" else orc.nf_full(mc)
    nf_used = np.array([float("%f" % v) for v in nf])
    p, _, _, fc = orc.ttest_rows(mc, nf_used, idxA, idxB)
    adj = orc.bh(p)
    keep = np.nonzero(adj <= pmax)[0]
    assert len(keep) > 0, "test needs survivors"
    world = 2
    # rank 0 must own the rows computeNF consumes in reference mode: split at 60% of the key range
    edges = [0, int(np.quantile(np.concatenate(keys), 0.6)), 1 << 62]
    case = dict(keys=keys, cnts=cnts, edges=edges, r=r, a=a, idxA=idxA, idxB=idxB, pmax=pmax, nf_mode=nf_mode, step=step)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(rk, world, port, case, q)) for rk in range(world)]
    [pr.start() for pr in procs]
    res = sorted(q.get(timeout=180) for _ in range(world))
    [pr.join(timeout=60) for pr in procs]
    assert all(pr.exitcode == 0 for pr in procs)
    for rk, gU, gR, gD, gnf, _, _ in res:
        assert (gU, gR, gD) == (U, len(mk), len(keep))
        assert np.allclose(gnf, nf, rtol=1e-12)
    got_keys = res[0][5] + res[1][5]                       # rank order == key order
    got_adj = res[0][6] + res[1][6]
    assert got_keys == mk[keep].tolist()
    assert np.allclose(got_adj, adj[keep], rtol=1e-12)
