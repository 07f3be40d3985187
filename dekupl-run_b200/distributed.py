"""Prefix-sharded hot path over N GPUs (one process per GPU, torch.distributed for the plumbing).

Rows are independent for the merge, the filter and the per-row test, so every rank runs those on
its own k-mer prefix shard with no data-path collective (rank order == key order, so the
concatenation of the ranks' outputs is the reference's sorted output). Global pieces only:
  * NF, reference mode : computeNF consumes a PREFIX of the sorted matrix -> the ranks owning the
                         first rows compute it; here rank 0 (its shard must cover the rows consumed),
                         then one broadcast of S doubles.
  * NF, full mode      : exact global median per sample by radix select; per round one all-reduce
                         of S*256 uint64 histogram bins (8 rounds).
  * BH                 : all-reduce of R; the candidate p-values (p <= P) are routed by value range to the
                         rank that ranks that range (all-to-all), which answers with the adjusted values
                         (all-to-all back); one all-gather of a scalar per rank carries the running minimum
                         across range borders.
The per-rank compute is behind a small backend interface so the collective logic can be tested on
CPU with gloo (tests/test_distributed_cpu.py drives it with an oracle-backed backend)."""
import numpy as np
import torch
import torch.distributed as dist


def nf_text_round(nf):
    """computeNF prints "%f" (computeNF.c:174); TtestFilter reads it back with atof (TtestFilter.c:261)"""
    return np.array([float("%f" % v) for v in nf])


class GpuBackend:
    """per-rank compute through the C ABI (libdekupl_b200.so)"""

    def __init__(self, ctx, device):
        self.ctx, self.device = ctx, device
        self.m = self.t = self.nfs = None
        # The library's kernels and torch's ops / NCCL collectives hand device buffers to each other (histograms,
        # candidate lists, adjusted values): both must run on ONE stream, or nothing orders them (a context's own
        # stream is non-blocking). See dk_set_stream in include/dekupl_b200.h.
        ctx.set_stream(torch.cuda.current_stream(device).cuda_stream)

    def join(self, key_ptrs, cnt_ptrs, n, r, a):
        self.m = self.ctx.join_counts_dev(key_ptrs, cnt_ptrs, n, r, a)
        return self.m.rows, self.m.union

    def nf_reference(self, step_rows=0):
        from .binding import NF_REFERENCE
        return self.ctx.compute_nf(self.m, NF_REFERENCE, step_rows)

    # full mode building blocks
    def nf_begin(self):
        import ctypes as C
        st, n = C.c_void_p(0), C.c_uint64(0)
        self.ctx._chk(self.ctx.L.dk_nf_begin(self.ctx.h, self.m.h, C.byref(st), C.byref(n)))
        self.nfs = st
        self.hist = torch.zeros(self.m.samples * 256, dtype=torch.int64, device=self.device)
        return int(n.value)

    def nf_set_rank(self, k):
        self.ctx._chk(self.ctx.L.dk_nf_set_rank(self.ctx.h, self.nfs, k))

    def nf_hist(self):
        self.ctx._chk(self.ctx.L.dk_nf_hist_dev(self.ctx.h, self.nfs, self.hist.data_ptr()))
        self.ctx.sync()
        return self.hist

    def nf_advance(self, hist):
        self.ctx._chk(self.ctx.L.dk_nf_advance_dev(self.ctx.h, self.nfs, hist.data_ptr()))

    def nf_result(self):
        import ctypes as C
        S = self.m.samples
        out = (C.c_double * S)()
        self.ctx._chk(self.ctx.L.dk_nf_result(self.ctx.h, self.nfs, out))
        self.ctx.L.dk_nf_end(self.ctx.h, self.nfs)
        self.nfs = None
        return np.array(out[:])

    def ttest_rows(self, nf, idxA, idxB):
        self.t = self.ctx.ttest_rows(self.m, nf, idxA, idxB)

    def count_le(self, cut):
        return self.t.count_le(cut) if self.t is not None else 0

    def next_cut(self, p_cut, n_global, R_global):
        return float(self.ctx.L.dk_bh_next_cut(float(p_cut), int(n_global), int(R_global)))

    def candidates(self, p_cut, cut=None):
        """-> float64 tensor (device) of this rank's candidate p-values (p <= cut, the narrowed p_cut)"""
        n, ptr = self.t.candidates(p_cut) if cut is None or cut >= p_cut else self.t.candidates_narrowed(cut, p_cut)
        self.ctx.sync()
        if n == 0:
            return torch.empty(0, dtype=torch.float64, device=self.device)
        from .torch_interop import tensor_from_ptr
        return tensor_from_ptr(ptr, n, torch.float64, self.device)

    def adjust_global(self, all_p, R_global):
        self.t.adjust_global(all_p.data_ptr() if all_p.numel() else 0, all_p.numel(), R_global)

    def bh_rank_slice(self, p, n_below, R_global):
        """-> (sorted p, F) device tensors for one p-range of the global candidate list"""
        sp, F = torch.empty_like(p), torch.empty_like(p)
        if p.numel():
            self.ctx._chk(self.ctx.L.dk_bh_rank_slice(self.ctx.h, p.data_ptr(), p.numel(), int(n_below), int(R_global),
                                                      sp.data_ptr(), F.data_ptr()))
        return sp, F

    def apply_adjusted(self, sp, F):
        self.t.apply_adjusted(sp.data_ptr() if sp.numel() else 0, F.data_ptr() if F.numel() else 0, sp.numel())

    def filter(self, pmax, fmin):
        return self.t.filter(pmax, fmin), self.t.ties

    def free(self):
        if self.t is not None:
            self.t.free()
        if self.m is not None:
            self.m.free()
        self.t = self.m = None


def _exchange(x, send_n, recv_n):
    """personalised exchange of 1-D tensors: x holds the pieces for rank 0, 1, ... back to back (send_n elements each);
    returns the pieces received from rank 0, 1, ... back to back. NCCL: one all-to-all. gloo has none, so the CPU tests
    gather everything and keep the slices addressed to this rank (same result, test-sized data only)."""
    rank, world = dist.get_rank(), dist.get_world_size()
    if dist.get_backend() == "nccl":
        out = x.new_empty(int(sum(recv_n)))
        dist.all_to_all_single(out, x.contiguous(), output_split_sizes=recv_n, input_split_sizes=send_n)
        return out
    sizes = torch.tensor(send_n, dtype=torch.int64, device=x.device)
    all_sizes = [torch.zeros_like(sizes) for _ in range(world)]
    dist.all_gather(all_sizes, sizes)
    mx = max(int(s.sum().item()) for s in all_sizes)
    pad = torch.zeros(max(mx, 1), dtype=x.dtype, device=x.device)
    pad[: x.numel()] = x
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad)
    parts = []
    for src in range(world):
        off = int(all_sizes[src][:rank].sum().item())
        parts.append(bufs[src][off: off + int(all_sizes[src][rank].item())])
    return torch.cat(parts) if parts else x.new_empty(0)


def _all_gather_var(x, group=None):
    """all-gather of 1-D tensors of different lengths (sizes first, then padded payload)"""
    world = dist.get_world_size(group)
    n = torch.tensor([x.numel()], dtype=torch.int64, device=x.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    mx = max(sizes)
    if mx == 0:
        return x
    pad = torch.zeros(mx, dtype=x.dtype, device=x.device)
    pad[: x.numel()] = x
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    return torch.cat([o[:s] for o, s in zip(out, sizes)])


def _bh_global(backend, R_glob, pmax):
    """Benjamini-Hochberg over the candidates of all ranks (TtestFilter.c:456-476): leaves the adjusted values on
    backend.t; returns the global candidate count"""
    import os
    import time
    rank, world = dist.get_rank(), dist.get_world_size()
    prof = [] if os.environ.get("DK_BH_PROF") else None

    def mark(name):
        if prof is not None:
            torch.cuda.synchronize()
            prof.append((name, time.perf_counter()))
    mark("start")
    # candidate narrowing over the GLOBAL counts (dk_bh_next_cut): a few counting passes + scalar all-reduces shrink the
    # candidate lists of all ranks from |{p <= P}| to about the rows that can matter, and everything below with them
    cut = pmax
    if pmax < 1.0 and R_glob > 0 and hasattr(backend, "count_le"):
        for _ in range(4):
            nle = torch.tensor([backend.count_le(cut)], dtype=torch.int64, device=backend.device)
            if world > 1:
                dist.all_reduce(nle)
            nxt = backend.next_cut(pmax, int(nle.item()), R_glob)
            if int(nle.item()) <= (1 << 20) * world or not (nxt < cut * 0.75):   # a list this short sorts for free
                break
            cut = nxt
    mark("narrow")
    cand = backend.candidates(pmax, cut)
    mark("candidates")
    if world == 1:
        backend.adjust_global(cand, R_glob)
        n_cand = int(cand.numel())
    else:
        # BH is split by p-value range: rank g owns the candidates in (g*cut/world, (g+1)*cut/world] (equal values share a
        # range; under the null p is uniform, so the ranges are balanced; `cut` is the narrowed P). Every rank ROUTES its candidates to their
        # owners (all-to-all), the owner ranks its range and answers with the adjusted values (all-to-all back): the
        # traffic and the sort per rank stay those of one shard however many ranks there are.
        edges = torch.tensor([cut * g / world for g in range(1, world)], dtype=torch.float64, device=cand.device)
        send, _ = torch.sort(cand)                                    # ranges become contiguous: edges[g-1] < p <= edges[g] -> g
        cuts = torch.searchsorted(send, edges, right=True)            # elements <= edges[g]
        zero = torch.zeros(1, dtype=cuts.dtype, device=cand.device)
        full = torch.full((1,), send.numel(), dtype=cuts.dtype, device=cand.device)
        counts = torch.diff(torch.cat([zero, cuts, full])).to(torch.int64)
        mark("sort+cuts")
        table = [torch.zeros_like(counts) for _ in range(world)]
        dist.all_gather(table, counts)
        table = torch.stack(table).cpu()                              # table[src][dst]
        send_n = [int(x) for x in table[rank].tolist()]
        recv_n = [int(x) for x in table[:, rank].tolist()]
        n_below = int(table[:, :rank].sum().item())
        n_cand = int(table.sum().item())
        mark("count table")
        mine = _exchange(send, send_n, recv_n)
        mark("route")
        sp, F = backend.bh_rank_slice(mine, n_below, R_glob)
        mark("rank slice")
        # the running minimum crosses range borders: lower F by the smallest F of all higher ranges
        first = F[:1] if F.numel() else torch.full((1,), float("inf"), dtype=torch.float64, device=cand.device)
        mins = [torch.empty_like(first) for _ in range(world)]
        dist.all_gather(mins, first.contiguous())
        above = torch.cat(mins[rank + 1:]).min() if rank + 1 < world else None
        if above is not None and F.numel():
            F = torch.minimum(F, above)
        # adjusted value of every routed candidate: adj depends on the value only, so any member of a tie run will do
        adj_mine = F[torch.searchsorted(sp, mine)] if mine.numel() else mine
        mark("mins+lookup")
        adj_send = _exchange(adj_mine.contiguous(), recv_n, send_n)   # back to the senders, in `send` order
        mark("route back")
        backend.apply_adjusted(send, adj_send)                        # `send` is this rank's candidate list, sorted
        mark("apply")
    if prof and rank == 0:
        print("[bh prof] " + " ".join(f"{b[0]}={1e3 * (b[1] - a[1]):.2f}ms" for a, b in zip(prof[:-1], prof[1:])) + f" cand={int(cand.numel())}",
              file=__import__("sys").stderr)
    return n_cand


def run_sharded(backend, key_ptrs, cnt_ptrs, n, r, a, idxA, idxB, pmax, fmin, nf_mode="reference", nf_step=0,
                nf_text_roundtrip=True, device=None):
    """one pass of the whole path on this rank's shard; returns the global stats (identical on every rank)"""
    rank, world = dist.get_rank(), dist.get_world_size()
    R_loc, U_loc = backend.join(key_ptrs, cnt_ptrs, n, r, a)
    tot = torch.tensor([R_loc, U_loc], dtype=torch.int64, device=device)
    dist.all_reduce(tot)
    R_glob, U_glob = int(tot[0].item()), int(tot[1].item())
    S = len(key_ptrs)
    nf = torch.zeros(S, dtype=torch.float64, device=device)
    if nf_mode == "reference":
        ok = torch.ones(1, dtype=torch.int64, device=device)
        if rank == 0:
            v, info = backend.nf_reference(nf_step)
            nf = torch.tensor(v, dtype=torch.float64, device=device)
            # computeNF.c:138-168 consumes a PREFIX of the sorted matrix; rank 0 can answer alone only if the
            # early stop fired inside its shard (rows consumed < its rows). Otherwise the reference would have
            # read rows owned by later ranks: refuse loudly instead of returning a different NF.
            if world > 1 and not (int(info["consumed"]) < R_loc):
                ok[0] = 0
        dist.broadcast(ok, src=0)
        if int(ok.item()) == 0:
            raise RuntimeError("reference-mode NF did not stop inside rank 0's shard; use nf_mode='full' or fewer shards")
        dist.broadcast(nf, src=0)
        nf = nf.cpu().numpy()
    else:
        n_pos = backend.nf_begin()
        t = torch.tensor([n_pos], dtype=torch.int64, device=device)
        dist.all_reduce(t)
        backend.nf_set_rank(int(t.item()) // 2)               # element [size/2], computeNF.c:43
        for _ in range(8):
            h = backend.nf_hist()
            dist.all_reduce(h)
            backend.nf_advance(h)
        nf = backend.nf_result()
    nf_used = nf_text_round(nf) if nf_text_roundtrip else nf
    backend.ttest_rows(nf_used, idxA, idxB)
    n_cand = _bh_global(backend, R_glob, pmax)
    D_loc, ties = backend.filter(pmax, fmin)
    d = torch.tensor([D_loc, ties[0], ties[1], ties[2]], dtype=torch.int64, device=device)
    dist.all_reduce(d)
    return dict(U=U_glob, R=R_glob, D=int(d[0].item()), ties=tuple(int(x) for x in d[1:].tolist()), nf=nf,
                R_local=R_loc, U_local=U_loc, D_local=D_loc, n_cand=n_cand)



class StreamBackend(GpuBackend):
    """per-rank compute of the chunk-streamed path: the dumps of this rank's k-mer range stay in pinned host memory"""

    def __init__(self, ctx, device, chunk_records):
        super().__init__(ctx, device)
        self.chunk_records = int(chunk_records)
        self.stream_stats = None

    # a rank whose chunks held no candidate row has no test object; it still takes part in every collective
    def candidates(self, p_cut, cut=None):
        if self.t is None:
            return torch.empty(0, dtype=torch.float64, device=self.device)
        return super().candidates(p_cut, cut)

    def adjust_global(self, all_p, R_global):
        if self.t is not None:
            super().adjust_global(all_p, R_global)

    def apply_adjusted(self, sp, F):
        if self.t is not None:
            super().apply_adjusted(sp, F)

    def free(self):
        if self.t is not None:
            self.t.free()
        if self.m is not None:
            self.m.free()
        self.t = self.m = None


def run_streamed(backend, key_ptrs, cnt_ptrs, n, r, a, idxA, idxB, pmax, fmin, nf_step=0, nf_text_roundtrip=True, device=None,
                 nf_from="auto"):
    """One pass of the chunk-streamed path (BASELINE.json configs[3]) on this rank's shard of HOST dumps.
    computeNF's factors (computeNF.c:138-168): nf_from="auto" lets rank 0 try the head of the table (its first chunk);
    if computeNF's early stop fires there, that is the answer (one broadcast). Otherwise — and always with
    nf_from="sweep" — every rank makes a first sweep that keeps only the all-positive rows (the only rows computeNF
    uses), they are gathered on rank 0 with global row numbers and the step / early-stop logic is replayed on them:
    exact for any number of ranks, and the same two passes over the table the reference pipeline makes.
    Then every rank walks its chunks again (copy in / merge / filter / test / keep the candidate rows) with no
    collective; R and U are all-reduced and BH runs over the candidates of all ranks exactly as in run_sharded."""
    import time
    rank, world = dist.get_rank(), dist.get_world_size()
    ctx = backend.ctx
    S = len(key_ptrs)
    nf = torch.zeros(S, dtype=torch.float64, device=device)
    flag = torch.zeros(2, dtype=torch.int64, device=device)      # [0] = head answered, [1] = error on rank 0
    err = None
    t_0 = time.perf_counter()
    if nf_from == "auto":
        if rank == 0:
            try:
                v, info = ctx.stream_nf_reference(key_ptrs, cnt_ptrs, n, r, a, backend.chunk_records, nf_step, widen=(world == 1))
                if info["stop_fired"] or world == 1:
                    nf = torch.tensor(v, dtype=torch.float64, device=device)
                    flag[0] = 1
            except Exception as e:                                # every rank must leave together
                flag[1], err = 1, e
        dist.broadcast(flag, src=0)
        if int(flag[1].item()):
            raise RuntimeError(f"reference-mode NF failed on rank 0: {err}")
    nf_how = "head"
    if not int(flag[0].item()):
        nf_how = "sweep"
        st1, prow, pcnt = ctx.stream_positive_rows(key_ptrs, cnt_ptrs, n, r, a, backend.chunk_records)
        rl = torch.tensor([st1["R"]], dtype=torch.int64, device=device)
        rs = [torch.zeros_like(rl) for _ in range(world)]
        dist.all_gather(rs, rl)
        rs = [int(x.item()) for x in rs]
        grow = torch.from_numpy((prow.astype(np.int64) + sum(rs[:rank]))).to(device)
        gcnt = torch.from_numpy(pcnt.reshape(-1).astype(np.int32)).to(device)
        all_rows, all_cnts = _all_gather_var(grow), _all_gather_var(gcnt)
        if rank == 0:
            try:
                v, _ = ctx.nf_reference_sparse(all_rows.cpu().numpy().astype(np.uint64), all_cnts.cpu().numpy().reshape(-1, S),
                                               sum(rs), nf_step)
                nf = torch.tensor(v, dtype=torch.float64, device=device)
            except Exception as e:
                flag[1], err = 1, e
        dist.broadcast(flag, src=0)
        if int(flag[1].item()):
            raise RuntimeError(f"reference-mode NF failed on rank 0: {err}")
    dist.broadcast(nf, src=0)
    nf = nf.cpu().numpy()
    nf_used = nf_text_round(nf) if nf_text_roundtrip else nf
    t_1 = time.perf_counter()
    st, backend.m, backend.t = ctx.stream_path(key_ptrs, cnt_ptrs, n, r, a, backend.chunk_records, nf_used, idxA, idxB, pmax)
    t_2 = time.perf_counter()
    backend.stream_stats = st
    tot = torch.tensor([st["R"], st["U"]], dtype=torch.int64, device=device)
    dist.all_reduce(tot)
    R_glob, U_glob = int(tot[0].item()), int(tot[1].item())
    n_cand = _bh_global(backend, R_glob, pmax)
    D_loc, ties = backend.filter(pmax, fmin) if backend.t is not None else (0, (0, 0, 0))
    d = torch.tensor([D_loc, ties[0], ties[1], ties[2]], dtype=torch.int64, device=device)
    dist.all_reduce(d)
    st["nf_sweep_h2d_bytes"] = st1["h2d_bytes"] if nf_how == "sweep" else 0
    st["nf_sweep_h2d_seconds"] = st1["h2d_seconds"] if nf_how == "sweep" else 0.0
    st["wall_ms_nf_head"], st["wall_ms_sweep"], st["wall_ms_bh"] = 1e3 * (t_1 - t_0), 1e3 * (t_2 - t_1), 1e3 * (time.perf_counter() - t_2)
    return dict(U=U_glob, R=R_glob, D=int(d[0].item()), ties=tuple(int(x) for x in d[1:].tolist()), nf=nf,
                R_local=st["R"], U_local=st["U"], D_local=D_loc, n_cand=n_cand, stream=st, nf_from=nf_how)
