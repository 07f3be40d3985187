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


def gen_shard(S, U_total, row0, row1, *, device, model="two_class", q=0.075, alpha=2.0, group_b=None, chunk=1 << 24,
              seed=SEED, pinned_out=False):
    """returns (keys[s] int64 tensors holding uint64 bit patterns, cnts[s] int32 tensors) for rows [row0,row1)"""
    stride = (1 << 62) // U_total
    depths = sample_depths(S, seed)
    group_b = set(range(S // 2, S)) if group_b is None else set(group_b)
    keys_parts = [[] for _ in range(S)]
    cnts_parts = [[] for _ in range(S)]
    for a in range(row0, row1, chunk):
        b = min(row1, a + chunk)
        i = torch.arange(a, b, device=device, dtype=torch.int64)
        h = mix64(i + seed)
        key = i * stride + (_lsr(h, 2) % stride)
        cls = _lsr(mix64(h ^ 0x5851F42D), 40).to(torch.float64) / float(1 << 24)     # row class uniform [0,1)
        eff = _lsr(mix64(h ^ 0x2545F491), 40).to(torch.float64) / float(1 << 24) < 0.01
        if model == "genome":
            single = cls < 0.9
            owner = _lsr(mix64(h ^ 0x1234567), 8) % S
        for s in range(S):
            hs = mix64(h + _wrap((s + 1) * 0x632BE59BD9B4E019))
            u = _lsr(hs, 40).to(torch.float64) / float(1 << 24)                       # presence
            if model == "two_class":
                present = u < torch.where(cls < 0.2, 0.95, q)
            else:
                present = torch.where(single, owner == s, u < 0.75)
            v = ((_lsr(hs, 8) & 0xFFFFFFFF).to(torch.float64) + 1.0) / 4294967296.0   # (0,1]
            z = torch.floor(v ** (-1.0 / (alpha - 1.0)))
            e = torch.where(eff, 4.0, 1.0) if s in group_b else 1.0
            c = torch.clamp(1.0 + torch.floor(z * depths[s] * e), 2.0, 2147483647.0).to(torch.int32)
            keys_parts[s].append(key[present])
            cnts_parts[s].append(c[present])
        del i, h, key, cls, eff
    keys, cnts = [], []
    for s in range(S):
        k = torch.cat(keys_parts[s]) if keys_parts[s] else torch.empty(0, dtype=torch.int64, device=device)
        c = torch.cat(cnts_parts[s]) if cnts_parts[s] else torch.empty(0, dtype=torch.int32, device=device)
        keys_parts[s], cnts_parts[s] = None, None
        keys.append(k)
        cnts.append(c)
    return keys, cnts
