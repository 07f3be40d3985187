"""Synthetic per-sample k-mer dumps of the BASELINE.json shapes, generated on the
device with torch (benchmark/test input plumbing, not part of the product path).

The union is implicit: row i in [row0,row1) has key  i*stride + mix(i) % stride  with
stride = 2^62 // U_total, i.e. strictly increasing stratified-uniform 62-bit keys, so any prefix
shard [row0,row1) can be generated independently (one shard per GPU) and every dump is sorted and
duplicate-free by construction. Presence and counts are counter-based hashes of (row, sample):
  two-class model : 20% "core" rows present w.p. 0.95 per sample, 80% "rare" rows w.p. q
  genome model    : 90% of rows present in exactly one sample, 10% in >= S/2 samples (config 5)
  counts          : c = 1 + floor(Z * d_s * e), Z ~ Zipf-like with P(Z>=z) = 1/z^(alpha-1),
                    depth d_s in [0.5,2], effect e = 4 on group B for 1% of rows; clipped to [2, 2^31-1]
"""
import torch

SEED = 20261104
_M1, _M2 = -7046029254386353131, -4658895280553007687      # 0x9E3779B97F4A7C15, 0xBF58476D1CE4E5B9 as int64
_M3 = -7723592293110705685                                  # 0x94D049BB133111EB


def _wrap(v):
    """python int -> int64 two's complement"""
    v &= (1 << 64) - 1
    return v - (1 << 64) if v >= (1 << 63) else v


def _lsr(x, k):
    return (x >> k) & ((1 << (64 - k)) - 1)


def mix64(x):
    """splitmix64 finaliser on int64 tensors (wrapping arithmetic)"""
    x = x * _M1
    x = (x ^ _lsr(x, 30)) * _M2
    x = (x ^ _lsr(x, 27)) * _M3
    return x ^ _lsr(x, 31)


def sample_depths(S, seed=SEED):
    g = torch.Generator().manual_seed(seed)
    return (0.5 + 1.5 * torch.rand(S, generator=g, dtype=torch.float64)).tolist()


BURST = 1      # > 1: k-mers come in runs of BURST consecutive integers separated by gaps (what low-complexity sequence
               # and shared prefixes do to a real sorted dump) instead of one stratified-uniform key per stride


def _rows(S, U_total, a, b, *, device, model, q, alpha, group_b, seed, eff=4.0, eff_frac=0.01, index=None):
    """rows [a,b) of the implicit union (or the rows listed in `index`): (key, [(present, count) per sample]) — the one
    place the generator is defined"""
    stride = (1 << 62) // U_total
    depths = sample_depths(S, seed)
    i = torch.arange(a, b, device=device, dtype=torch.int64) if index is None else index
    h = mix64(i + seed)
    if BURST > 1:          # row i = member (i % BURST) of run (i // BURST): consecutive k-mers, then a gap of BURST strides
        key = (i // BURST) * (stride * BURST) + (i % BURST)
    else:
        key = i * stride + (_lsr(h, 2) % stride)
    cls = _lsr(mix64(h ^ 0x5851F42D), 40).to(torch.float64) / float(1 << 24)     # row class uniform [0,1)
    effr = _lsr(mix64(h ^ 0x2545F491), 40).to(torch.float64) / float(1 << 24) < eff_frac
    if model == "genome":
        single = cls < 0.9
        owner = _lsr(mix64(h ^ 0x1234567), 8) % S

    def sample(s):
        hs = mix64(h + _wrap((s + 1) * 0x632BE59BD9B4E019))
        u = _lsr(hs, 40).to(torch.float64) / float(1 << 24)                       # presence
        if model == "two_class":
            present = u < torch.where(cls < 0.2, 0.95, q)
        else:
            present = torch.where(single, owner == s, u < 0.75)
        v = ((_lsr(hs, 8) & 0xFFFFFFFF).to(torch.float64) + 1.0) / 4294967296.0   # (0,1]
        z = torch.floor(v ** (-1.0 / (alpha - 1.0)))
        e = torch.where(effr, float(eff), 1.0) if s in group_b else 1.0
        c = torch.clamp(1.0 + torch.floor(z * depths[s] * e), 2.0, 2147483647.0).to(torch.int32)
        return present, c
    return key, sample


def gen_shard(S, U_total, row0, row1, *, device, model="two_class", q=0.075, alpha=2.0, group_b=None, chunk=1 << 24,
              seed=SEED, pinned_out=False, eff=4.0, eff_frac=0.01):
    """returns (keys[s] int64 tensors holding uint64 bit patterns, cnts[s] int32 tensors) for rows [row0,row1)"""
    group_b = set(range(S // 2, S)) if group_b is None else set(group_b)
    keys_parts = [[] for _ in range(S)]
    cnts_parts = [[] for _ in range(S)]
    for a in range(row0, row1, chunk):
        b = min(row1, a + chunk)
        key, sample = _rows(S, U_total, a, b, device=device, model=model, q=q, alpha=alpha, group_b=group_b, seed=seed,
                            eff=eff, eff_frac=eff_frac)
        for s in range(S):
            present, c = sample(s)
            keys_parts[s].append(key[present])
            cnts_parts[s].append(c[present])
        del key, sample
    keys, cnts = [], []
    for s in range(S):
        k = torch.cat(keys_parts[s]) if keys_parts[s] else torch.empty(0, dtype=torch.int64, device=device)
        c = torch.cat(cnts_parts[s]) if cnts_parts[s] else torch.empty(0, dtype=torch.int32, device=device)
        keys_parts[s], cnts_parts[s] = None, None
        keys.append(k)
        cnts.append(c)
    return keys, cnts


def expected_invariants(S, U_total, row0, row1, r, a, *, device, model="two_class", q=0.075, alpha=2.0, group_b=None,
                        chunk=1 << 24, seed=SEED, eff=4.0, eff_frac=0.01):
    """What joinCounts must produce from gen_shard's dumps of rows [row0,row1), computed from the generator's own
    presence/count model without merging anything (joinCounts.c:81-83,104): U = rows present in any sample,
    R = rows with #{present and count >= a} >= r, xor of the surviving keys, sum of all surviving counts (mod 2^64)."""
    group_b = set(range(S // 2, S)) if group_b is None else set(group_b)
    U = R = 0
    kx = torch.zeros((), dtype=torch.int64, device=device)
    csum = 0
    n_rec = 0
    for a0 in range(row0, row1, chunk):
        b0 = min(row1, a0 + chunk)
        key, sample = _rows(S, U_total, a0, b0, device=device, model=model, q=q, alpha=alpha, group_b=group_b, seed=seed,
                            eff=eff, eff_frac=eff_frac)
        anyp = torch.zeros(b0 - a0, dtype=torch.bool, device=device)
        rec = torch.zeros(b0 - a0, dtype=torch.int32, device=device)
        tot = torch.zeros(b0 - a0, dtype=torch.int64, device=device)
        for s in range(S):
            present, c = sample(s)
            anyp |= present
            rec += (present & (c >= a)).to(torch.int32)
            tot += torch.where(present, c.to(torch.int64), 0)
            n_rec += int(present.sum().item())
        keep = rec >= r
        if r <= 0:
            keep = keep & anyp
        U += int(anyp.sum().item())
        R += int(keep.sum().item())
        kk = key[keep]
        if kk.numel():
            # xor-reduce: torch has no bitwise_xor reduction; fold halves
            v = kk
            while v.numel() > 1:
                if v.numel() % 2:
                    v = torch.cat([v, torch.zeros(1, dtype=v.dtype, device=device)])
                v = v[: v.numel() // 2] ^ v[v.numel() // 2:]
            kx ^= v[0]
        csum = (csum + int(tot[keep].sum().item())) & ((1 << 64) - 1)
    return dict(U=U, R=R, key_xor=int(kx.item()) & ((1 << 64) - 1), count_sum=csum, records=n_rec)


def key_stride(U_total):
    """row i of the union owns the keys [i*stride, (i+1)*stride): a k-mer boundary maps back to a generator row"""
    return (1 << 62) // U_total


def gen_row_sample(S, U_total, n_sample, *, device, model="two_class", q=0.075, alpha=2.0, seed=SEED):
    """every (U_total // n_sample)-th row of the whole union, as per-sample sorted key tensors: the coarse global view
    from which every rank derives the same count-balanced shard boundaries (dk_key_splitters)"""
    step = max(1, U_total // n_sample)
    idx = torch.arange(0, U_total, step, device=device, dtype=torch.int64)
    key, sample = _rows(S, U_total, 0, 0, device=device, model=model, q=q, alpha=alpha, group_b=set(range(S // 2, S)),
                        seed=seed, index=idx)
    return [key[sample(s)[0]].contiguous() for s in range(S)]
