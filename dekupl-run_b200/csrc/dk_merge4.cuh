// dk_merge4.cuh — K2 v4: the k-way join of a round WITHOUT sorting: monotone buckets + owner election
// (included by dk_join.cu; S <= 64; reference semantics: share/joinCounts/joinCounts.c:72-115).
//
// ncu of v3 (profiles/r01m_ncu_merge_v3.csv): 3.0 shared-memory wavefronts per record, 67% of the LSU
// wavefront peak — log2(S) merge levels move every record through shared memory 2x per level plus the
// merge-path searches. v4 never orders records against each other. A round still stages one window per dump
// by 1-D TMA bulk copies and takes every key below the pivot (smallest look-ahead key), but then:
//   B1  bucket(key) = floor((key - base) * M / span), M = 4096 buckets over the round's key span: MONOTONE in
//       the key, so bucket order is k-mer order; expected distinct k-mers per bucket ~0.2;
//   B2  owner election by plain racing stores (no atomics: ATOMS costs 2 cycles per lane on this part):
//       every record stores its slot into own[0][bucket]; after a barrier whichever slot survived is the
//       bucket's level-0 owner; records whose k-mer equals the owner's k-mer are resolved at level 0, the
//       others repeat at level 1.. (J = 4 tables). A bucket therefore lists its distinct k-mers, one per level,
//       in arbitrary order; a k-mer unresolved after J levels (>4 distinct k-mers in one bucket) shrinks the
//       round (pivot := that k-mer) and the round is redone with finer buckets — exact for any key distribution;
//   B3  one thread per 8 buckets counts the k-mers of its buckets, ranks the (<= 4) k-mers of a multi-k-mer
//       bucket by direct comparison, and a block scan turns that into the row id of (bucket, level);
//   B4  recurrence (joinCounts.c:81-83): records with count >= a set one BYTE of their row in a byte matrix
//       (plain stores), a thread per row sums it; verdict rec >= r (:104) -> scan -> in-order output slots;
//   B5  records of surviving rows write their 4-byte count into the row image (odd row pitch: conflict-free),
//       the owner record writes the k-mer; coalesced 128-bit copy-out into the chunked matrix.
// A k-mer that occurs twice inside one dump (never produced by jellyfish) raises the overflow flag and the host
// reruns the join with v3 (which keeps such copies as separate rows exactly like the reference).
#pragma once

#ifndef M4_NTL
#define M4_NTL 8
#define M4_CTAS 4
#endif
#define M4_NT (1 << M4_NTL)         // 256 threads; 4 CTAs per SM
#ifndef M4_VT
#define M4_VT 11
#endif
#define M4_W (M4_NT * M4_VT)       // 2816 record slots per round (A/B at cfg 2: VT 8/9/10/11/12 = 51.7/48.0/46.4/45.6/45.7 ms)
#define M4_M (8 * M4_NT)            // 2048 buckets per round (8 per thread in the bucket scan)
#ifndef M4_J
#define M4_J 4                     // owner levels (distinct k-mers a bucket can hold)
#endif
#define M4_EMPTY 0xFFFFu
#define M4_ALIAS_BYTES ((M4_J - 1) * M4_M * 2)   // own[1..J-1] is reused for the byte matrix and the row images

struct M4Cfg {
    int S, SP, levels, caps, ks, S4, BR, abw, RMAX, RS;
    uint32_t divS4;
    uint32_t off_own, off_info, off_orow, off_small, keybuf_off, smem_bytes;
};

struct M4Small {
    unsigned long long red[3][M4_NT / 32];
    unsigned long long bar;
    unsigned long long kcut;
    uint32_t scan[M4_NT / 32];
    uint32_t part, total_real;
    uint32_t chunk_cur, rowpos, seq, newbase, nnew, have_cur;
    uint32_t prow, puni;
};
template <int SP>
struct M4Samp {
    unsigned long long cur[SP], endg[SP];
    const int32_t *cptr[SP];
    uint16_t n[SP], L[SP];
};

__device__ __forceinline__ uint32_t m4_block_excl_scan(uint32_t v, M4Small &sm, uint32_t &total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    __syncthreads();                       // protects sm.scan from the previous use
    if (lane == 31) sm.scan[w] = x;
    __syncthreads();
    uint32_t t = lane < M4_NT / 32 ? sm.scan[lane] : 0u;
#pragma unroll
    for (int o = 1; o < M4_NT / 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, t, o); if (lane >= o) t += y; }
    total = __shfl_sync(0xffffffffu, t, M4_NT / 32 - 1);
    const uint32_t off = __shfl_sync(0xffffffffu, t, w > 0 ? w - 1 : 0);
    return (w > 0 ? off : 0u) + x - v;
}
__device__ __forceinline__ unsigned long long m4_warp_min(unsigned long long v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) { unsigned long long y = __shfl_xor_sync(0xffffffffu, v, o); v = y < v ? y : v; }
    return v;
}
__device__ __forceinline__ unsigned long long m4_warp_max(unsigned long long v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) { unsigned long long y = __shfl_xor_sync(0xffffffffu, v, o); v = y > v ? y : v; }
    return v;
}
__device__ __forceinline__ uint32_t m4_half(const uint4 &v, int q)
{
    const uint32_t w = (q >> 1) == 0 ? v.x : (q >> 1) == 1 ? v.y : (q >> 1) == 2 ? v.z : v.w;
    return (q & 1) ? (w >> 16) : (w & 0xffffu);
}

// -DDK_M4_PROF: thread 0 of every CTA accumulates the cycles between phase marks into prof[1..11] (units of 16 cycles)
#ifdef DK_M4_PROF
#define M4_MARK(i) do { if (tid == 0) { long long t__ = clock64(); atomicAdd(&prof[i], (uint32_t)((t__ - tprev) >> 4)); tprev = t__; } } while (0)
#else
#define M4_MARK(i) do { } while (0)
#endif

template <int LV>
__global__ void __launch_bounds__(M4_NT, M4_CTAS) k_merge4(MergeParams P, M4Cfg G, uint32_t *overflow)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned long long *bufK = reinterpret_cast<unsigned long long *>(smem_raw);
    uint16_t *own = reinterpret_cast<uint16_t *>(smem_raw + G.off_own);          // [J][M] owner slots
    uint32_t *info = reinterpret_cast<uint32_t *>(smem_raw + G.off_info);        // [M] row base | level ranks << 16
    uint16_t *orow = reinterpret_cast<uint16_t *>(smem_raw + G.off_orow);        // [W] output slot of a row
    uint32_t *abw = reinterpret_cast<uint32_t *>(own + M4_M);                    // byte matrix rows (aliases own[1..])
    int32_t *rowbuf = reinterpret_cast<int32_t *>(own + M4_M);                   // row images (after the verdicts)
    unsigned long long *keybuf = reinterpret_cast<unsigned long long *>(reinterpret_cast<unsigned char *>(own + M4_M) + G.keybuf_off);
    M4Small &sm = *reinterpret_cast<M4Small *>(smem_raw + G.off_small);

    constexpr int SP = 1 << LV, caps = M4_W >> LV, ks = (caps + 4 + 3) & ~3, TPSL = M4_NTL - LV, TPS = 1 << TPSL;
    M4Samp<SP> &ss = *reinterpret_cast<M4Samp<SP> *>(smem_raw + G.off_small + sizeof(M4Small));
    const int S = G.S, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t npart = *P.npart;
    const uint32_t bar = smem_u32(&sm.bar);
    const bool owner = tid < S;
    if (tid == 0) {
        mbar_init(bar, (uint32_t)S);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (tid < SP) ss.cptr[tid] = owner ? P.T.cnts[tid] : nullptr;
    __syncthreads();
    uint32_t phase = 0;
    const int min_ab = P.min_abund, min_rec = P.min_rec;
    const int ps = tid >> TPSL, psub = tid & (TPS - 1);
    const uint32_t ABW = (uint32_t)G.abw, RS = (uint32_t)G.RS;
#ifdef DK_M4_PROF
    uint32_t *prof = overflow + 1;     // scal[4..15]
    long long tprev = 0;
#endif

    for (;;) {
        if (tid == 0) sm.part = atomicAdd(P.work_counter, 1u);
        __syncthreads();
        const uint32_t part = sm.part;
        if (part >= npart) break;
        if (tid < SP) {
            ss.cur[tid] = owner ? P.starts[(size_t)part * S + tid] : 0ull;
            ss.endg[tid] = owner ? P.starts[(size_t)(part + 1) * S + tid] : 0ull;
        }
        __syncthreads();
        if (tid == 0) { sm.chunk_cur = CHUNK_NONE; sm.rowpos = 0; sm.seq = 0; sm.prow = 0; sm.puni = 0; }

        for (;;) {   // ---- rounds
            // R0: stage the key windows (TMA), read the (count >= a) bits while the copies fly
#ifdef DK_M4_PROF
            if (tid == 0) tprev = clock64();
#endif
            uint32_t lo = 0, L = 0;
            if (owner) {
                // per-dump state lives in shared / constant-cached memory, not in registers (64-register budget)
                const unsigned long long cur = ss.cur[tid], endg = ss.endg[tid], nsrc = P.T.n[tid];
                const uint64_t *ksrc = P.T.keys[tid];
                const unsigned long long a0 = cur & ~3ull;
                lo = (uint32_t)(cur - a0);
                const unsigned long long left = endg - cur;
                L = (uint32_t)(left < (unsigned long long)(caps + 1) ? left : (unsigned long long)(caps + 1));
                uint32_t tq = 0;
                unsigned long long mend = a0;
                if (L) {
                    const uint32_t nq = (lo + L + 3u) >> 2;
                    const unsigned long long avail = (nsrc - a0) >> 2;
                    tq = (uint32_t)(avail < nq ? avail : nq);
                    mend = a0 + 4ull * nq;
                    if (mend > nsrc) mend = nsrc;
                }
                unsigned long long *kd = bufK + (size_t)tid * ks;
                if (tq) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    mbar_expect_tx(bar, tq * 32u);
                    tma_bulk_g2s(smem_u32(kd), ksrc + a0, tq * 32u, bar);
                } else mbar_arrive(bar);
                for (unsigned long long g = a0 + 4ull * tq; g < mend; g++) kd[g - a0] = ksrc[g];   // array tail (< 4 records)
            }
            const unsigned long long pcur = ss.cur[ps];
            const int32_t *cp = ss.cptr[ps] + pcur;
            uint32_t abm = 0;
            {
                const unsigned long long pleft = ss.endg[ps] - pcur;
                const uint32_t pL = (uint32_t)(pleft < (unsigned long long)caps ? pleft : (unsigned long long)caps);
#pragma unroll
                for (int k = 0; k < M4_VT; k++) {
                    const uint32_t i = (uint32_t)(psub + k * TPS);
                    if (i < pL && __ldg(cp + i) >= min_ab) abm |= 1u << k;                                  // joinCounts.c:81-83
                }
            }
            mbar_wait(bar, phase);
            phase ^= 1u;
            __syncthreads();
            M4_MARK(1);
            // R1: pivot (smallest look-ahead key), base (smallest first key), largest staged key
            unsigned long long candk = DK_SENTINEL, firstk = DK_SENTINEL, lastk = 0;
            uint32_t ncm = 0;
            if (owner && L) {
                const unsigned long long *kw = bufK + (size_t)tid * ks + lo;
                firstk = kw[0];
                if (L == (uint32_t)(caps + 1)) { candk = kw[caps]; ncm = (uint32_t)caps; } else ncm = L;
                lastk = kw[ncm - 1];
            }
            constexpr int OW = (SP + 31) / 32;                     // warps that hold dump owners (tid < S <= SP)
            if (warp < OW) {
                candk = m4_warp_min(candk);
                firstk = m4_warp_min(firstk);
                lastk = m4_warp_max(lastk);
                if (lane == 0) { sm.red[0][warp] = candk; sm.red[1][warp] = firstk; sm.red[2][warp] = lastk; }
            }
            if (tid < SP) ss.L[tid] = (uint16_t)L;
            __syncthreads();
            unsigned long long pivot = DK_SENTINEL, base = DK_SENTINEL, maxk = 0;
#pragma unroll
            for (int i = 0; i < OW; i++) {
                unsigned long long a = sm.red[0][i], b = sm.red[1][i], c = sm.red[2][i];
                pivot = a < pivot ? a : pivot;
                base = b < base ? b : base;
                maxk = c > maxk ? c : maxk;
            }
            if (base == DK_SENTINEL) break;                        // every window empty: partition done
            if (maxk != DK_SENTINEL && maxk + 1ull < pivot) pivot = maxk + 1ull;   // span of the keys actually staged
            M4_MARK(2);

            const uint32_t slo = (uint32_t)pcur & 3u;
            const unsigned long long *kw = bufK + (size_t)ps * ks + slo;
            const uint32_t slot0 = (uint32_t)ps * ks + slo;
            uint32_t n = 0, T = 0, nrows = 0, consm = 0, lvl = 0, headm = 0;
            uint32_t bkp[(M4_VT + 1) / 2];            // bucket of record k: 16 bits each
#define M4_BK(k) ((bkp[(k) >> 1] >> (16 * ((k) & 1))) & 0xffffu)
            bool stop = false;
            for (;;) {   // ---- bucket pass, redone with a smaller pivot when a bucket cannot be resolved
                if (tid == 0) sm.total_real = 0;
                {
                    uint4 *o4 = reinterpret_cast<uint4 *>(own);
                    const uint4 e4 = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
#pragma unroll
                    for (int i = 0; i < (M4_J * M4_M * 2 / 16) / M4_NT; i++) o4[tid + i * M4_NT] = e4;
                }
                __syncthreads();                                   // (A)
                if (tid == 0) sm.kcut = DK_SENTINEL;
                n = 0;
                if (owner && ncm) {
                    const unsigned long long *kwo = bufK + (size_t)tid * ks + lo;
                    uint32_t a = 0, b = ncm;
                    while (a < b) { uint32_t mid = (a + b) >> 1; if (kwo[mid] < pivot) a = mid + 1; else b = mid; }
                    n = a;
                    if (n) atomicAdd(&sm.total_real, n);
                }
                if (tid < SP) ss.n[tid] = (uint16_t)n;
                __syncthreads();                                   // (B)
                M4_MARK(3);
                T = sm.total_real;
                if (T == 0) {                                      // a window full of one k-mer: not this kernel's case
                    if (tid == 0) atomicExch(overflow, 1u);
                    stop = true;
                    break;
                }
                // B1: buckets
                const unsigned long long span = pivot - base;     // >= 1
                const int wsh = (64 - __clzll((long long)span)) > 16 ? (64 - __clzll((long long)span)) - 16 : 0;
                const uint32_t fmax = (uint32_t)((span - 1ull) >> wsh);
                const uint32_t scale = (fmax + 1u <= (uint32_t)M4_M) ? 65536u : (uint32_t)(((uint32_t)M4_M << 16) / (fmax + 1u));
                const uint32_t sn = ss.n[ps], sL = ss.L[ps];
                consm = 0;
#pragma unroll
                for (int k = 0; k < (M4_VT + 1) / 2; k++) bkp[k] = 0;
                bool dupf = false;
#pragma unroll
                for (int k = 0; k < M4_VT; k++) {
                    // straight-line code (no branch per record): slots past the consumable prefix hold stale keys that
                    // are inside the window buffer; their bucket is masked into range and never used
                    const uint32_t i = (uint32_t)(psub + k * TPS);
                    const unsigned long long key = kw[i], nxt = kw[i + 1];
                    const bool c = i < sn;
                    consm |= (c ? 1u : 0u) << k;
                    const uint32_t f = (uint32_t)((key - base) >> wsh);
                    bkp[k >> 1] |= (((f * scale) >> 16) & (uint32_t)(M4_M - 1)) << (16 * (k & 1));
                    dupf |= c && i + 1u < sL && nxt == key;
                }
                if (dupf) atomicExch(overflow, 1u);                 // copies of a k-mer inside a dump: v3's case
                // B2: owner election. Level 0 takes every record (unrolled, buckets in registers); the few records
                // still unresolved after it walk levels 1.. in data-dependent loops (bucket recomputed from the key)
                uint32_t unres = consm;
                lvl = 0; headm = 0;
                bool all_done = false;
#pragma unroll
                for (int k = 0; k < M4_VT; k++)
                    if ((unres >> k) & 1u) own[M4_BK(k)] = (uint16_t)(slot0 + (uint32_t)(psub + k * TPS));
                __syncthreads();
#pragma unroll
                for (int k = 0; k < M4_VT; k++)
                    if ((unres >> k) & 1u) {
                        const uint32_t i = (uint32_t)(psub + k * TPS);
                        const uint32_t h = own[M4_BK(k)];
                        const bool ishead = h == slot0 + i;
                        if (ishead || bufK[h] == kw[i]) {
                            unres &= ~(1u << k);
                            if (ishead) headm |= 1u << k;
                        }
                    }
#pragma unroll 1
                for (int l = 1; l < M4_J; l++) {
                    uint16_t *ow = own + l * M4_M;
                    for (uint32_t m = unres; m; m &= m - 1) {
                        const uint32_t i = (uint32_t)psub + (uint32_t)(__ffs(m) - 1) * TPS;
                        const uint32_t f = (uint32_t)((kw[i] - base) >> wsh);
                        ow[(f * scale) >> 16] = (uint16_t)(slot0 + i);
                    }
                    if (!__syncthreads_or(unres != 0u)) { all_done = true; break; }
                    for (uint32_t m = unres; m; m &= m - 1) {
                        const uint32_t k = (uint32_t)(__ffs(m) - 1);
                        const uint32_t i = (uint32_t)psub + k * TPS;
                        const unsigned long long key = kw[i];
                        const uint32_t f = (uint32_t)((key - base) >> wsh);
                        const uint32_t h = ow[(f * scale) >> 16];
                        const bool ishead = h == slot0 + i;
                        if (ishead || bufK[h] == key) {
                            unres &= ~(1u << k);
                            lvl |= (uint32_t)l << (2u * k);
                            if (ishead) headm |= 1u << k;
                        }
                    }
                }
                if (!all_done) {
                    if (unres) {
                        unsigned long long mk = DK_SENTINEL;
#pragma unroll
                        for (int k = 0; k < M4_VT; k++)
                            if ((unres >> k) & 1u) { const unsigned long long key = kw[psub + k * TPS]; mk = key < mk ? key : mk; }
                        atomicMin(&sm.kcut, mk);
                    }
                    __syncthreads();                               // (C)
                    const unsigned long long kc = sm.kcut;
                    if (kc != DK_SENTINEL) {                       // > J distinct k-mers in one bucket: finer buckets
                        pivot = kc > base + 1ull ? kc : base + 1ull;
                        continue;
                    }
                }
                M4_MARK(4);
                // B3: k-mers per bucket, ranks inside multi-k-mer buckets, row ids
                uint32_t nkm = 0, mine = 0;                        // nkm: 3 bits per bucket
                {
                    const uint4 o0 = reinterpret_cast<const uint4 *>(own)[tid];
                    if ((o0.x & o0.y & o0.z & o0.w) != 0xffffffffu) {
#pragma unroll
                        for (int q = 0; q < 8; q++) nkm |= (m4_half(o0, q) != M4_EMPTY ? 1u : 0u) << (3 * q);
#pragma unroll
                        for (int l = 1; l < M4_J; l++) {
                            const uint4 ol = reinterpret_cast<const uint4 *>(own + l * M4_M)[tid];
                            if ((ol.x & ol.y & ol.z & ol.w) != 0xffffffffu) {
#pragma unroll
                                for (int q = 0; q < 8; q++) nkm += (m4_half(ol, q) != M4_EMPTY ? 1u : 0u) << (3 * q);
                            }
                        }
#pragma unroll
                        for (int q = 0; q < 8; q++) mine += (nkm >> (3 * q)) & 7u;
                    }
                }
                uint32_t tot;
                uint32_t rb = m4_block_excl_scan(mine, sm, tot);
                nrows = tot;
                if (nrows > (uint32_t)G.RMAX) {                    // more rows than the byte matrix holds: halve the span
                    pivot = base + (span >> 1);
                    continue;
                }
                {
                    uint32_t inf[8];
#pragma unroll
                    for (int q = 0; q < 8; q++) {
                        const uint32_t nk = (nkm >> (3 * q)) & 7u;
                        uint32_t v = rb;
                        if (nk >= 2) {
                            unsigned long long kk[M4_J];
#pragma unroll
                            for (int l = 0; l < M4_J; l++)
                                kk[l] = (uint32_t)l < nk ? bufK[own[l * M4_M + 8 * tid + q]] : DK_SENTINEL;
#pragma unroll
                            for (int l = 0; l < M4_J; l++) {
                                uint32_t r = 0;
#pragma unroll
                                for (int l2 = 0; l2 < M4_J; l2++) r += kk[l2] < kk[l] ? 1u : 0u;
                                if ((uint32_t)l < nk) v |= r << (16 + 2 * l);
                            }
                        }
                        inf[q] = v;
                        rb += nk;
                    }
                    uint4 *i4 = reinterpret_cast<uint4 *>(info) + 2 * tid;
                    i4[0] = make_uint4(inf[0], inf[1], inf[2], inf[3]);
                    i4[1] = make_uint4(inf[4], inf[5], inf[6], inf[7]);
                }
                __syncthreads();                                   // own[1..] is dead from here on
                break;
            }
            if (stop) break;
            if (owner) ss.cur[tid] += n;                           // read again only at the top of the next round
#ifdef DK_M4_PREFETCH   // A/B at cfg 2: 48.8 ms without, 51.5 ms with (the owner threads stall on the issue) — off
            if (owner) {   // next round's windows (keys and counts) on their way into L2 while this round is finished
                const unsigned long long nc = ss.cur[tid] + n, e = ss.endg[tid];
                if (nc < e) {
                    const unsigned long long a0 = nc & ~3ull;
                    unsigned long long a1 = nc + (unsigned long long)(caps + 1);
                    a1 = a1 < e ? a1 : e;
                    a1 &= ~3ull;                                   // whole 4-record groups only: never past the arrays
                    if (a1 > a0) {
                        const uint32_t nrec = (uint32_t)(a1 - a0);
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(P.T.keys[tid] + a0), "r"(nrec * 8u) : "memory");
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ss.cptr[tid] + a0), "r"(nrec * 4u) : "memory");
                    }
                }
            }
#endif
            M4_MARK(5);
            // B4: recurrence. Zero the byte matrix (it aliases own[1..])
            {
                const uint32_t nw4 = (nrows * ABW + 3u) >> 2;
                uint4 *z = reinterpret_cast<uint4 *>(abw);
                for (uint32_t i = tid; i < nw4; i += M4_NT) z[i] = make_uint4(0, 0, 0, 0);
            }
            __syncthreads();
            {
                unsigned char *abb = reinterpret_cast<unsigned char *>(abw);
                const uint32_t m = consm & abm;
                uint32_t rwp[(M4_VT + 1) / 2];                     // from here on bkp holds the ROW of record k
#pragma unroll
                for (int k = 0; k < (M4_VT + 1) / 2; k++) rwp[k] = 0;
#pragma unroll
                for (int k = 0; k < M4_VT; k++) {                  // all loads first (a store in between would order them)
                    const uint32_t inf = info[M4_BK(k)];           // unconditional: in range for every slot, used under consm only
                    const uint32_t row = (inf & 0x1fffu) + ((inf >> (16 + 2 * ((lvl >> (2 * k)) & 3u))) & 3u);
                    rwp[k >> 1] |= row << (16 * (k & 1));
                }
#pragma unroll
                for (int k = 0; k < (M4_VT + 1) / 2; k++) bkp[k] = rwp[k];
#pragma unroll
                for (int k = 0; k < M4_VT; k++)
                    if ((m >> k) & 1u) abb[M4_BK(k) * (ABW * 4u) + (uint32_t)ps] = 1;
            }
            __syncthreads();
            M4_MARK(6);
            const uint32_t RPT = (nrows + M4_NT - 1) / M4_NT, rb0 = (uint32_t)tid * RPT;
            uint32_t passmask = 0;
            for (uint32_t i = 0; i < RPT && rb0 + i < nrows; i++) {
                const uint32_t *rw = abw + (rb0 + i) * ABW;
                uint32_t rec = 0;
                for (uint32_t j = 0; j < ABW; j++) rec += __popc(rw[j]);
                if ((int)rec >= min_rec) passmask |= 1u << i;                                              // joinCounts.c:104
            }
            uint32_t nout;
            const uint32_t obase = m4_block_excl_scan(__popc(passmask), sm, nout);
            for (uint32_t i = 0; i < RPT && rb0 + i < nrows; i++)
                orow[rb0 + i] = ((passmask >> i) & 1u) ? (uint16_t)(obase + __popc(passmask & ((1u << i) - 1u))) : (uint16_t)0xFFFFu;
            // chunk allocation for nout rows appended to the partition's chunk sequence: thread 0 ISSUES the global atomic
            // here and consumes its result only after the row images are built (the round trip hides behind that pass)
            uint32_t a_have = 0, a_nnew = 0, a_nb = 0;
            if (tid == 0) {
                a_have = (sm.chunk_cur != CHUNK_NONE) ? 1u : 0u;
                const uint32_t slots = a_have ? (DK_CH - sm.rowpos) : 0u;
                a_nnew = nout > slots ? (nout - slots + DK_CH - 1) / DK_CH : 0u;
                if (a_nnew) a_nb = atomicAdd(P.chunk_counter, a_nnew);
            }
            {   // the first batch of row images is zeroed while thread 0's chunk allocation is in flight
                // (the byte matrix under it is dead: every verdict was computed before the scan's barriers)
                const uint32_t nbr0 = nout < (uint32_t)G.BR ? nout : (uint32_t)G.BR;
                const uint32_t n4 = (nbr0 * RS + 3u) >> 2;
                int4 *z = reinterpret_cast<int4 *>(rowbuf);
                for (uint32_t i = tid; i < n4; i += M4_NT) z[i] = make_int4(0, 0, 0, 0);
            }
            __syncthreads();
            M4_MARK(7);
            // B5: row images of the surviving rows, coalesced copy-out
            {
                for (uint32_t b0 = 0; b0 < nout; b0 += (uint32_t)G.BR) {
                    const uint32_t nbr = (nout - b0) < (uint32_t)G.BR ? (nout - b0) : (uint32_t)G.BR;
                    if (b0) {
                        const uint32_t n4 = (nbr * RS + 3u) >> 2;
                        int4 *z = reinterpret_cast<int4 *>(rowbuf);
                        for (uint32_t i = tid; i < n4; i += M4_NT) z[i] = make_int4(0, 0, 0, 0);
                        __syncthreads();
                    }
                    M4_MARK(8);
                    {
                        // three passes so that the loads of all 9 records are in flight together
                        uint32_t op[(M4_VT + 1) / 2], live = 0;
                        int32_t cv[M4_VT];
#pragma unroll
                        for (int k = 0; k < (M4_VT + 1) / 2; k++) op[k] = 0;
#pragma unroll
                        for (int k = 0; k < M4_VT; k++) {
                            if (!((consm >> k) & 1u)) continue;
                            const uint32_t o = (uint32_t)orow[M4_BK(k)] - b0;
                            if (o < nbr) { live |= 1u << k; op[k >> 1] |= o << (16 * (k & 1)); }
                        }
#pragma unroll
                        for (int k = 0; k < M4_VT; k++) {
                            cv[k] = 0;
                            if ((live >> k) & 1u) cv[k] = __ldg(cp + (uint32_t)(psub + k * TPS));
                        }
#pragma unroll
                        for (int k = 0; k < M4_VT; k++) {
                            if (!((live >> k) & 1u)) continue;
                            const uint32_t o = (op[k >> 1] >> (16 * (k & 1))) & 0xffffu;
                            rowbuf[o * RS + (uint32_t)ps] = cv[k];
                            if ((headm >> k) & 1u) keybuf[o] = kw[(uint32_t)(psub + k * TPS)];
                        }
                    }
                    if (tid == 0 && b0 == 0) {                      // the allocation is needed from here on
                        for (uint32_t k = 0; k < a_nnew; k++) {
                            const uint32_t id = a_nb + k;
                            if (id < P.max_chunks) { P.chunk_part[id] = part; P.chunk_seq[id] = sm.seq + k; P.chunk_rows[id] = DK_CH; }
                        }
                        sm.have_cur = a_have; sm.newbase = a_nb;
                    }
                    __syncthreads();
                    M4_MARK(9);
                    const uint32_t have = sm.have_cur, nb = sm.newbase, ccur = sm.chunk_cur;
                    const uint32_t rp = have ? sm.rowpos : 0u;
                    if (G.S4) {
                        const uint32_t S4 = (uint32_t)G.S4, n4 = nbr * S4;
                        for (uint32_t i = tid; i < n4; i += M4_NT) {
                            const uint32_t rr = S4 == 1 ? i : __umulhi(i, G.divS4), c4 = i - rr * S4;
                            const uint32_t v = rp + b0 + rr, k = v / DK_CH, off = v % DK_CH;
                            const uint32_t chunk = (have && k == 0) ? ccur : nb + (k - have);
                            const int32_t *src = rowbuf + rr * RS + 4u * c4;
                            if (chunk < P.max_chunks)
                                reinterpret_cast<int4 *>(P.out_cnts + ((size_t)chunk * DK_CH + off) * S)[c4] = make_int4(src[0], src[1], src[2], src[3]);
                        }
                    } else {
                        for (uint32_t rr = tid / 32u; rr < nbr; rr += M4_NT / 32u) {
                            const uint32_t v = rp + b0 + rr, k = v / DK_CH, off = v % DK_CH;
                            const uint32_t chunk = (have && k == 0) ? ccur : nb + (k - have);
                            if (chunk < P.max_chunks)
                                for (uint32_t c = lane; c < (uint32_t)S; c += 32u) P.out_cnts[((size_t)chunk * DK_CH + off) * S + c] = rowbuf[rr * RS + c];
                        }
                    }
                    for (uint32_t rr = tid; rr < nbr; rr += M4_NT) {
                        const uint32_t v = rp + b0 + rr, k = v / DK_CH, off = v % DK_CH;
                        const uint32_t chunk = (have && k == 0) ? ccur : nb + (k - have);
                        if (chunk < P.max_chunks) P.out_keys[(size_t)chunk * DK_CH + off] = keybuf[rr];
                    }
                    __syncthreads();
                    M4_MARK(10);
                }
            }
            // R7: advance the partition state
            if (tid == 0) {
                const uint32_t v = (a_have ? sm.rowpos : 0u) + nout;
                if (a_nnew) {
                    sm.chunk_cur = a_nb + a_nnew - 1;
                    sm.seq += a_nnew;
                    sm.rowpos = v - (a_have + a_nnew - 1) * DK_CH;
                } else sm.rowpos = v;
                if (sm.chunk_cur != CHUNK_NONE && sm.rowpos == DK_CH) { sm.chunk_cur = CHUNK_NONE; sm.rowpos = 0; }
                sm.prow += nout; sm.puni += nrows;
            }
            // no barrier here: the next round touches shared state only in ways the barriers above already order
            M4_MARK(11);
        }
        __syncthreads();
        if (tid == 0) {
            if (sm.chunk_cur != CHUNK_NONE && sm.chunk_cur < P.max_chunks) P.chunk_rows[sm.chunk_cur] = sm.rowpos;
            P.part_nchunks[part] = sm.seq;
            P.part_rows[part] = sm.prow;
            P.part_union[part] = sm.puni;
        }
        __syncthreads();
    }
}

// host side: geometry for S samples
static bool m4_make_cfg(int S, size_t smem_budget, M4Cfg &G)
{
    if (S < 1 || S > 64) return false;
    int sp = 1, lv = 0;
    while (sp < S) { sp <<= 1; lv++; }
    G.S = S; G.SP = sp; G.levels = lv;
    G.caps = M4_W / sp;
    G.ks = (G.caps + 4 + 3) & ~3;
    if ((size_t)sp * G.ks >= 8192) return false;                  // slots are 13-bit (info) / 16-bit (own) values
    G.S4 = (S % 4 == 0) ? S / 4 : 0;
    G.divS4 = G.S4 > 1 ? (uint32_t)(((1ull << 32) + G.S4 - 1) / G.S4) : 0;
    G.abw = ((S + 3) / 4) | 1;                                    // odd word pitch of the byte matrix
    G.RS = S | 1;                                                 // odd word pitch of the row images
    int rmax = (int)(M4_ALIAS_BYTES / (4 * (size_t)G.abw));
    if (rmax > 8191) rmax = 8191;
    G.RMAX = rmax;
    int BR = (int)((M4_ALIAS_BYTES - 32) / ((size_t)4 * G.RS + 8));
    if (BR > 512) BR = 512;
    if (BR < 4) return false;
    G.BR = BR;
    G.keybuf_off = (uint32_t)dk_round_up((size_t)BR * G.RS * 4, 16);
    if (G.keybuf_off + (size_t)BR * 8 > M4_ALIAS_BYTES) return false;
    size_t off = dk_round_up((size_t)sp * G.ks * 8, 128);
    G.off_own = (uint32_t)off; off += (size_t)M4_J * M4_M * 2;
    G.off_info = (uint32_t)off; off += (size_t)M4_M * 4;
    G.off_orow = (uint32_t)off; off += dk_round_up((size_t)(G.RMAX < M4_W ? G.RMAX : M4_W) * 2 + 16, 128);   // one slot per row
    G.off_small = (uint32_t)off; off += dk_round_up(sizeof(M4Small) + (size_t)sp * 28 + 16, 128);
    G.smem_bytes = (uint32_t)off;
    return off <= smem_budget;
}
