// dk_merge3.cuh — K2 v3: tile merge by a balanced pairwise merge-path tree in shared memory
// (included by dk_join.cu; any S <= 256; reference semantics: share/joinCounts/joinCounts.c:72-115).
//
// r01b ncu of v2 (profiles/r01b_ncu_merge_v2.csv): a thread-sequential 16-way merge keeps only
// 13 of 32 lanes busy (key ranges hold very different numbers of rows) and pays O(S) per row.
// v3 has no data-dependent trip counts at all. A CTA walks its partition in ROUNDS:
//   R0  every sample contributes a window of <= caps keys (+1 look-ahead), staged by 1-D TMA bulk
//       copies (cp.async.bulk + mbarrier complete_tx) into bufK; counts never occupy shared memory
//       (which is what lets three CTAs share an SM): while the key copies are in flight every thread
//       reads the counts of the records it will pack and keeps only the (count >= a) bits;
//   R1  pivot = smallest look-ahead key, so every key < pivot is complete in the staged windows;
//   R2  each staged record becomes ONE 64-bit word  (key - base) << 16 | dup << 14 | ab << 13 | slot:
//       order by word == order by (key, copy index inside the dump, ...); ab = (count >= a) is read
//       from global memory with consecutive lanes on consecutive records; slot addresses the record;
//       the round's key span is clipped to 2^48 so the word is exact (never lossy);
//   R3  log2(S_pad) merge levels, ping-pong between two buffers: every thread finds its diagonal
//       by merge-path binary search and emits exactly 17 words (odd => conflict-free blocked stores);
//       runs are padded with PAD and fenced with GUARD words so the serial merge needs no bounds checks;
//   R4  rows = runs of equal (key,dup): head flags + block scan give row ids, the scan of the ab bits
//       gives each row's recurrence (joinCounts.c:81-83), a scan of the verdicts (rec >= r, :104) gives
//       in-order output slots;
//   R5  surviving rows are assembled in the merge buffer that is free after the last level (zero-filled,
//       one 4-byte count per record re-read from L2) and copied out with coalesced 128-bit stores.
// Duplicate k-mers inside one dump become consecutive rows exactly like the reference; more than
// 4 copies of one k-mer in one dump (never produced by jellyfish) raise the overflow flag and the
// host reruns the join with the v1 kernel.
// r01d ncu: the kernel is bound by the shared-memory data pipe (68% of peak wavefronts, half of them
// bank conflicts of the per-lane LDS.64 in the merge-path search and the serial merge), not by HBM.
#pragma once

#define M3_NT 256
#define M3_VT 17
#define M3_TP (M3_NT * M3_VT)      // 4352 word slots per round
#define M3_SB 13                   // slot bits (slot < S_pad*ks < 8192)
#define M3_ABB M3_SB               // bit 13: count >= min_abund (the verdict needs no second look at the count)
#define M3_DSH (M3_SB + 1)         // bits 14..15: copy index inside the dump
#define M3_DB 2
#define M3_KSH (M3_DSH + M3_DB)    // key - base above bit 16
#define M3_PAD 0xFFFFFFFFFFFFFFFEull
#define M3_GUARD 0xFFFFFFFFFFFFFFFFull
#define M3_SPAN ((1ull << (64 - M3_KSH)) - 1ull)   // consumable keys: key - base < M3_SPAN

struct M3Cfg {
    int S, S_pad, caps, ks, levels, tps_log2, BR, S4;
    uint32_t divKS, divS, divS4;                  // ceil(2^32 / x): exact floor(q/x) for q < 2^16
    uint32_t off_A, off_small, yoff_rowbuf, yoff_keybuf, smem_bytes;
};

struct M3Small {
    unsigned long long red[2][M3_NT / 32];
    unsigned long long bar;
    uint32_t scan[M3_NT / 32];
    uint32_t part, total_real, any;
    uint32_t chunk_cur, rowpos, seq, newbase, nnew, have_cur;
    uint32_t prow, puni;
};
// per-sample round state, sized by the padded sample count (follows M3Small in shared memory)
template <int SP>
struct M3Samp {
    unsigned long long cur[SP], endg[SP];   // cursor / partition end of every dump (record indices)
    const int32_t *cptr[SP];          // count arrays stay in global memory / L2
    uint16_t n[SP];
};

__device__ __forceinline__ uint32_t m3_block_excl_scan(uint32_t v, M3Small &sm, uint32_t &total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    __syncthreads();                       // protects sm.scan from the previous use
    if (lane == 31) sm.scan[w] = x;
    __syncthreads();
    uint32_t off = 0, tot = 0;
#pragma unroll
    for (int i = 0; i < M3_NT / 32; i++) { uint32_t s = sm.scan[i]; if (i < w) off += s; tot += s; }
    total = tot;
    return off + x - v;
}

__device__ __forceinline__ unsigned long long m3_warp_min(unsigned long long v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) { unsigned long long y = __shfl_xor_sync(0xffffffffu, v, o); v = y < v ? y : v; }
    return v;
}

// shared-memory loads by 32-bit shared address: the merge path below selects ONE address per step and issues ONE
// load for the whole warp (two predicated loads would each replay their own bank conflicts), and the binary search
// probes the high 32-bit halves first (a 32-bit access conflicts on 32 banks, a 64-bit one on 16 bank pairs)
__device__ __forceinline__ unsigned long long m3_lds64(uint32_t addr)
{
    unsigned long long v;
    asm volatile("ld.shared.u64 %0, [%1];" : "=l"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t m3_lds32(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

// -DDK_M3_PROF: thread 0 of every CTA accumulates the cycles between phase marks into prof[1..7] (units of 16 cycles)
#ifdef DK_M3_PROF
#define M3_MARK(i) do { if (tid == 0) { long long t__ = clock64(); atomicAdd(&prof[i], (uint32_t)((t__ - tprev) >> 4)); tprev = t__; } } while (0)
#else
#define M3_MARK(i) do { } while (0)
#endif

template <int LV>
__global__ void __launch_bounds__(M3_NT, 3) k_merge3(MergeParams P, M3Cfg G, uint32_t *overflow)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned long long *bufK = reinterpret_cast<unsigned long long *>(smem_raw);                 // staging keys / merge buffer B
    unsigned long long *bufA = reinterpret_cast<unsigned long long *>(smem_raw + G.off_A);
    M3Small &sm = *reinterpret_cast<M3Small *>(smem_raw + G.off_small);

    const int S = G.S, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // geometry is compile-time per level count: S_pad = 2^LV runs of caps words, TPS threads per run in the pack pass
    constexpr int SP = 1 << LV, caps = M3_TP >> LV, ks = (caps + 4 + 3) & ~3, TPSL = 8 - LV, TPS = 1 << TPSL;
    M3Samp<SP> &ss = *reinterpret_cast<M3Samp<SP> *>(smem_raw + G.off_small + sizeof(M3Small));
    const uint32_t npart = *P.npart;
    const uint32_t bar = smem_u32(&sm.bar);
    const bool owner = tid < S;
    if (tid == 0) {
        mbar_init(bar, (uint32_t)S);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    const uint64_t *ksrc = owner ? P.T.keys[tid] : nullptr;
    const int32_t *csrc = owner ? P.T.cnts[tid] : nullptr;
    const unsigned long long nsrc = owner ? P.T.n[tid] : 0ull;
    if (tid < SP) ss.cptr[tid] = csrc;
    __syncthreads();
    uint32_t phase = 0;
    const int min_ab = P.min_abund, min_rec = P.min_rec;
#ifdef DK_M3_PROF
    uint32_t *prof = overflow + 1;     // scal[4..11]
    long long tprev = 0;
#endif

    for (;;) {
        if (tid == 0) sm.part = atomicAdd(P.work_counter, 1u);
        __syncthreads();
        const uint32_t part = sm.part;
        if (part >= npart) break;
        unsigned long long cur = 0, endg = 0;
        if (owner) { cur = P.starts[(size_t)part * S + tid]; endg = P.starts[(size_t)(part + 1) * S + tid]; }
        if (tid < SP) { ss.cur[tid] = cur; ss.endg[tid] = endg; }
        __syncthreads();
        if (tid == 0) { sm.chunk_cur = CHUNK_NONE; sm.rowpos = 0; sm.seq = 0; sm.prow = 0; sm.puni = 0; }

        for (;;) {   // ---- rounds
#ifdef DK_M3_PROF
            if (tid == 0) tprev = clock64();
#endif
            // R0: stage the windows
            uint32_t lo = 0, L = 0;
            if (tid == 0) { sm.total_real = 0; }
            if (owner) {
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
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // earlier generic accesses precede the async writes
                    mbar_expect_tx(bar, tq * 32u);
                    tma_bulk_g2s(smem_u32(kd), ksrc + a0, tq * 32u, bar);
                } else mbar_arrive(bar);
                for (unsigned long long g = a0 + 4ull * tq; g < mend; g++) kd[g - a0] = ksrc[g];   // array tail (< 4 records)
            }
            // counts are never staged: while the key windows are in flight every thread reads the counts of the
            // records it will pack (consecutive lanes -> consecutive records) and keeps only the (count >= a) bits
            const int ps = tid >> TPSL, psub = tid & (TPS - 1);
            const unsigned long long pcur = ss.cur[ps];
            uint32_t abm = 0;
            {
                const unsigned long long pleft = ss.endg[ps] - pcur;
                const uint32_t pL = (uint32_t)(pleft < (unsigned long long)caps ? pleft : (unsigned long long)caps);
                const int32_t *cp = ss.cptr[ps] + pcur;
#pragma unroll
                for (int k = 0; k < M3_VT; k++) {
                    const uint32_t i = (uint32_t)(psub + k * TPS);
                    if (i < pL && __ldg(cp + i) >= min_ab) abm |= 1u << k;                                  // joinCounts.c:81-83
                }
            }
            mbar_wait(bar, phase);
            phase ^= 1u;
            __syncthreads();
            M3_MARK(1);
            // R1: pivot (smallest look-ahead key) and base (smallest first key)
            unsigned long long candk = DK_SENTINEL, firstk = DK_SENTINEL;
            uint32_t ncm = 0;
            if (owner && L) {
                const unsigned long long *kw = bufK + (size_t)tid * ks + lo;
                firstk = kw[0];
                if (L == (uint32_t)(caps + 1)) { candk = kw[caps]; ncm = (uint32_t)caps; } else ncm = L;
            }
            candk = m3_warp_min(candk);
            firstk = m3_warp_min(firstk);
            if (lane == 0) { sm.red[0][warp] = candk; sm.red[1][warp] = firstk; }
            __syncthreads();
            unsigned long long pivot = DK_SENTINEL, base = DK_SENTINEL;
#pragma unroll
            for (int i = 0; i < M3_NT / 32; i++) {
                unsigned long long a = sm.red[0][i], b = sm.red[1][i];
                pivot = a < pivot ? a : pivot;
                base = b < base ? b : base;
            }
            if (base == DK_SENTINEL) break;                        // every window empty: partition done
            {
                unsigned long long lim = base + M3_SPAN;
                if (lim < base) lim = DK_SENTINEL;                 // saturate
                pivot = lim < pivot ? lim : pivot;
            }
            // R2: consumable prefix of every window
            uint32_t n = 0;
            if (owner && ncm) {
                const unsigned long long *kw = bufK + (size_t)tid * ks + lo;
                uint32_t a = 0, b = ncm;
                while (a < b) { uint32_t mid = (a + b) >> 1; if (kw[mid] < pivot) a = mid + 1; else b = mid; }
                n = a;
                if (n) atomicAdd(&sm.total_real, n);
            }
            if (tid < SP) ss.n[tid] = (uint16_t)n;
            __syncthreads();
            M3_MARK(2);
            const uint32_t T = sm.total_real;
            if (T == 0) {                                          // >= caps copies of one k-mer inside a dump
                if (tid == 0) atomicExch(overflow, 1u);
                break;
            }
            // R3: pack (key,dup,slot) words into the level-0 runs
            {
                const int s = ps, sub = psub;
                const uint32_t slo = (uint32_t)pcur & 3u, sn = ss.n[s];
                const unsigned long long *kw = bufK + (size_t)s * ks + slo;
                unsigned long long *run = bufA + (size_t)s * (caps + 1);
                const uint32_t slot0 = (uint32_t)s * ks + slo;
#pragma unroll
                for (int k = 0; k < M3_VT; k++) {
                    const uint32_t i = (uint32_t)(sub + k * TPS);
                    unsigned long long w = M3_PAD;
                    if (i < sn) {
                        const unsigned long long key = kw[i];
                        uint32_t d = 0;
                        if (i > 0 && kw[i - 1] == key) {           // copies of one k-mer inside a dump (rare)
                            d = 1;
                            while (d < i && kw[i - 1 - d] == key) d++;
                            if (d > 3) { atomicExch(overflow, 1u); d = 3; }
                        }
                        w = ((key - base) << M3_KSH) | ((unsigned long long)d << M3_DSH) | (unsigned long long)(((abm >> k) & 1u) << M3_ABB) |
                            (unsigned long long)(slot0 + i);
                    }
                    run[i] = w;
                }
                if (sub == 0) run[caps] = M3_GUARD;
            }
            __syncthreads();
            M3_MARK(3);
            // R4: merge tree
            unsigned long long *X = bufA, *Y = bufK;
            {
#pragma unroll
                for (int l = 0; l < LV; l++) {
                    const int len = caps << l;
                    const int sh = TPSL + 1 + l;
                    const int pair = tid >> sh;
                    const int d = (tid & ((1 << sh) - 1)) * M3_VT;
                    const unsigned long long *a_base = X + (size_t)(2 * pair) * (len + 1);
                    const unsigned long long *b_base = a_base + (len + 1);
                    const uint32_t xa = smem_u32(a_base), xb = smem_u32(b_base);
                    int alo = d > len ? d - len : 0, ahi = d < len ? d : len;
                    while (alo < ahi) {
                        const int mid = (alo + ahi) >> 1;
                        const uint32_t qa = xa + 8u * (uint32_t)mid, qb = xb + 8u * (uint32_t)(d - 1 - mid);
                        const uint32_t ah = m3_lds32(qa + 4u), bh = m3_lds32(qb + 4u);
                        bool le = ah < bh;
                        if (ah == bh) le = m3_lds32(qa) <= m3_lds32(qb);
                        if (le) alo = mid + 1; else ahi = mid;
                    }
                    uint32_t pa = xa + 8u * (uint32_t)alo, pb = xb + 8u * (uint32_t)(d - alo);
                    unsigned long long a = m3_lds64(pa), b = m3_lds64(pb);
                    unsigned long long *out = Y + (size_t)pair * (2 * len + 1) + d;
#pragma unroll
                    for (int i = 0; i < M3_VT; i++) {
                        const bool ta = a <= b;
                        out[i] = ta ? a : b;
                        pa += ta ? 8u : 0u;
                        pb += ta ? 0u : 8u;
                        const unsigned long long v = m3_lds64(ta ? pa : pb);
                        a = ta ? v : a;
                        b = ta ? b : v;
                    }
                    if (d + M3_VT == 2 * len) out[M3_VT] = M3_GUARD;
                    __syncthreads();
                    unsigned long long *t = X; X = Y; Y = t;
                }
            }
            M3_MARK(4);
            // R5: rows. X = merged words (T real, then PAD); Y is free: rowab u16[TP+1], orow u16[TP]
            uint16_t *rowab = reinterpret_cast<uint16_t *>(Y);
            constexpr int ROWAB_N = (M3_TP + 8) > (SP * ks + 8) ? (M3_TP + 8) : (SP * ks + 8);
            uint16_t *orow = rowab + ROWAB_N;
            uint16_t *slot2row = rowab;                      // rowab is dead once the verdicts are out
            int32_t *rowbuf = reinterpret_cast<int32_t *>(reinterpret_cast<unsigned char *>(Y) + G.yoff_rowbuf);
            unsigned long long *keybuf = reinterpret_cast<unsigned long long *>(reinterpret_cast<unsigned char *>(Y) + G.yoff_keybuf);
            const uint32_t j0 = (uint32_t)tid * M3_VT;
            uint32_t headmask = 0, abmask = 0;
            {
                unsigned long long prevkp = (tid == 0) ? ~0ull : (X[j0 - 1] >> M3_DSH);
#pragma unroll
                for (int i = 0; i < M3_VT; i++) {
                    const unsigned long long w = X[j0 + i];
                    const bool real = w < M3_PAD;
                    const unsigned long long kp = w >> M3_DSH;
                    if (real && kp != prevkp) headmask |= 1u << i;
                    prevkp = kp;
                    if (real && (((uint32_t)w >> M3_ABB) & 1u)) abmask |= 1u << i;
                }
            }
            uint32_t tot1;
            const uint32_t ex1 = m3_block_excl_scan((__popc(headmask) << 16) | __popc(abmask), sm, tot1);
            const uint32_t rowbase = ex1 >> 16, abbase = ex1 & 0xffffu, nrows = tot1 >> 16;
            {
                uint32_t hm = headmask;
                while (hm) {
                    const int i = __ffs(hm) - 1;
                    hm &= hm - 1;
                    const uint32_t below = (1u << i) - 1u;
                    rowab[rowbase + __popc(headmask & below)] = (uint16_t)(abbase + __popc(abmask & below));
                }
                if (tid == M3_NT - 1) rowab[nrows] = (uint16_t)(tot1 & 0xffffu);
            }
            __syncthreads();
            M3_MARK(5);
            // verdicts: rows are dealt out in blocks of RPT = ceil(nrows/NT) per thread (nrows is usually << TP)
            const uint32_t RPT = (nrows + M3_NT - 1) / M3_NT, rb = (uint32_t)tid * RPT;
            uint32_t passmask = 0;
            if (rb < nrows) {
                uint32_t prev = rowab[rb];
                for (uint32_t i = 0; i < RPT && rb + i < nrows; i++) {
                    const uint32_t cur_ab = rowab[rb + i + 1];
                    if ((int)(cur_ab - prev) >= min_rec) passmask |= 1u << i;                              // joinCounts.c:104
                    prev = cur_ab;
                }
            }
            uint32_t nout;
            const uint32_t obase = m3_block_excl_scan(__popc(passmask), sm, nout);
            for (uint32_t i = 0; i < RPT && rb + i < nrows; i++)
                orow[rb + i] = ((passmask >> i) & 1u) ? (uint16_t)(obase + __popc(passmask & ((1u << i) - 1u))) : (uint16_t)0xFFFFu;
            // chunk allocation for nout rows appended to the partition's chunk sequence
            if (tid == 0) {
                const uint32_t have = (sm.chunk_cur != CHUNK_NONE) ? 1u : 0u;
                const uint32_t slots = have ? (DK_CH - sm.rowpos) : 0u;
                const uint32_t nnew = nout > slots ? (nout - slots + DK_CH - 1) / DK_CH : 0u;
                uint32_t nb = 0;
                if (nnew) {
                    nb = atomicAdd(P.chunk_counter, nnew);
                    for (uint32_t k = 0; k < nnew; k++) {
                        const uint32_t id = nb + k;
                        if (id < P.max_chunks) { P.chunk_part[id] = part; P.chunk_seq[id] = sm.seq + k; P.chunk_rows[id] = DK_CH; }
                    }
                }
                sm.have_cur = have; sm.newbase = nb; sm.nnew = nnew;
            }
            __syncthreads();
            M3_MARK(6);
            // R6: assemble surviving rows in shared memory, copy out coalesced
            {
                const uint32_t have = sm.have_cur, nb = sm.newbase, ccur = sm.chunk_cur;
                const uint32_t rp = have ? sm.rowpos : 0u;
                // entry sweep (shared memory only): every record learns the output row of its k-mer, by staged slot
                if (nout) {
                    const uint32_t nreal = T > j0 ? (T - j0 < (uint32_t)M3_VT ? T - j0 : (uint32_t)M3_VT) : 0u;
                    const uint32_t *Xlo = reinterpret_cast<const uint32_t *>(X);
                    uint32_t r = rowbase - 1u;
#pragma unroll
                    for (int i = 0; i < M3_VT; i++) {
                        r += (headmask >> i) & 1u;
                        if ((uint32_t)i < nreal) slot2row[Xlo[2 * (j0 + i)] & ((1u << M3_SB) - 1u)] = orow[r];
                    }
                }
                for (uint32_t b0 = 0; b0 < nout; b0 += (uint32_t)G.BR) {
                    const uint32_t nbr = (nout - b0) < (uint32_t)G.BR ? (nout - b0) : (uint32_t)G.BR;
                    const uint32_t words = nbr * (uint32_t)S;
                    {
                        int4 *z = reinterpret_cast<int4 *>(rowbuf);
                        const uint32_t n4 = (words + 3u) >> 2;
                        for (uint32_t i = tid; i < n4; i += M3_NT) z[i] = make_int4(0, 0, 0, 0);
                    }
                    __syncthreads();
                    M3_MARK(8);
                    {   // sample-major sweep: the same (dump, record) mapping as the pack pass, so the count reads are
                        // coalesced (consecutive lanes -> consecutive records) and hit L1/L2
                        const uint32_t sn = ss.n[ps], slot0 = (uint32_t)ps * ks + ((uint32_t)pcur & 3u);
                        const int32_t *cp = ss.cptr[ps] + pcur;
#pragma unroll
                        for (int k = 0; k < M3_VT; k++) {
                            const uint32_t i = (uint32_t)(psub + k * TPS);
                            if (i < sn) {
                                const uint32_t o = (uint32_t)slot2row[slot0 + i] - b0;
                                if (o < nbr) rowbuf[o * (uint32_t)S + (uint32_t)ps] = __ldg(cp + i);
                            }
                        }
                        uint32_t hm = headmask;                  // keys of the surviving rows, from their head records
                        while (hm) {
                            const int i = __ffs(hm) - 1;
                            hm &= hm - 1;
                            const uint32_t o = (uint32_t)orow[rowbase + __popc(headmask & ((1u << i) - 1u))] - b0;
                            if (o < nbr) keybuf[o] = base + (X[j0 + i] >> M3_KSH);
                        }
                    }
                    __syncthreads();
                    M3_MARK(9);
                    if (G.S4) {
                        const uint32_t S4 = (uint32_t)G.S4, n4 = nbr * S4;
                        const int4 *src4 = reinterpret_cast<const int4 *>(rowbuf);
                        for (uint32_t i = tid; i < n4; i += M3_NT) {
                            const uint32_t rr = S4 == 1 ? i : __umulhi(i, G.divS4), c4 = i - rr * S4;   // ceil(2^32/1) does not fit 32 bits
                            const uint32_t v = rp + b0 + rr, k = v / DK_CH, off = v % DK_CH;
                            const uint32_t chunk = (have && k == 0) ? ccur : nb + (k - have);
                            if (chunk < P.max_chunks)
                                reinterpret_cast<int4 *>(P.out_cnts + ((size_t)chunk * DK_CH + off) * S)[c4] = src4[i];
                        }
                    } else {
                        for (uint32_t i = tid; i < words; i += M3_NT) {
                            const uint32_t rr = S == 1 ? i : __umulhi(i, G.divS), c = i - rr * (uint32_t)S;
                            const uint32_t v = rp + b0 + rr, k = v / DK_CH, off = v % DK_CH;
                            const uint32_t chunk = (have && k == 0) ? ccur : nb + (k - have);
                            if (chunk < P.max_chunks) P.out_cnts[((size_t)chunk * DK_CH + off) * S + c] = rowbuf[i];
                        }
                    }
                    for (uint32_t rr = tid; rr < nbr; rr += M3_NT) {
                        const uint32_t v = rp + b0 + rr, k = v / DK_CH, off = v % DK_CH;
                        const uint32_t chunk = (have && k == 0) ? ccur : nb + (k - have);
                        if (chunk < P.max_chunks) P.out_keys[(size_t)chunk * DK_CH + off] = keybuf[rr];
                    }
                    __syncthreads();
                    M3_MARK(10);
                }
            }
            // R7: advance the partition state
            cur += n;
            if (tid < SP) ss.cur[tid] = cur;                     // read again only after the barrier below
            if (tid == 0) {
                const uint32_t v = (sm.have_cur ? sm.rowpos : 0u) + nout;
                if (sm.nnew) {
                    sm.chunk_cur = sm.newbase + sm.nnew - 1;
                    sm.seq += sm.nnew;
                    sm.rowpos = v - (sm.have_cur + sm.nnew - 1) * DK_CH;
                } else sm.rowpos = v;
                if (sm.chunk_cur != CHUNK_NONE && sm.rowpos == DK_CH) { sm.chunk_cur = CHUNK_NONE; sm.rowpos = 0; }
                sm.prow += nout; sm.puni += nrows;
            }
            __syncthreads();
            M3_MARK(7);
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
static bool m3_make_cfg(int S, size_t smem_budget, M3Cfg &G)
{
    if (S < 1 || S > 256) return false;
    int sp = 1, lv = 0;
    while (sp < S) { sp <<= 1; lv++; }
    G.S = S; G.S_pad = sp; G.levels = lv; G.tps_log2 = 8 - lv;
    G.caps = M3_TP / sp;
    G.ks = (G.caps + 4 + 3) & ~3;
    if ((size_t)sp * G.ks >= (1u << M3_SB)) return false;
    G.S4 = (S % 4 == 0) ? S / 4 : 0;
    auto cdiv32 = [](uint32_t x) { return (uint32_t)(((1ull << 32) + x - 1) / x); };
    G.divKS = cdiv32((uint32_t)G.ks); G.divS = cdiv32((uint32_t)S); G.divS4 = G.S4 ? cdiv32((uint32_t)G.S4) : 0;
    size_t bK = (size_t)sp * G.ks * 8, bM = (size_t)(M3_TP + sp) * 8;
    if (bK < bM) bK = bM;
    // the merge buffer that is free after the last level also holds rowab/orow (u16 x TP each), the row
    // assembly buffer (BR rows x S counts) and BR keys
    const size_t rowab_n = (size_t)(M3_TP + 8) > (size_t)sp * G.ks + 8 ? (size_t)(M3_TP + 8) : (size_t)sp * G.ks + 8;
    const size_t yrow = dk_round_up((rowab_n + M3_TP + 8) * 2, 16);
    if (bM < yrow + 64) return false;
    int BR = (int)((bM - yrow - 32) / ((size_t)4 * S + 8));
    if (BR > 256) BR = 256;
    if (BR < 4) return false;
    G.BR = BR;
    G.yoff_rowbuf = (uint32_t)yrow;
    G.yoff_keybuf = (uint32_t)(yrow + dk_round_up((size_t)BR * S * 4, 16));
    size_t off = dk_round_up(bK, 128);
    G.off_A = (uint32_t)off; off += dk_round_up(bM, 128);
    G.off_small = (uint32_t)off; off += dk_round_up(sizeof(M3Small) + (size_t)sp * 26 + 16, 128);
    G.smem_bytes = (uint32_t)off;
    return off <= smem_budget;
}
