// dk_merge5.cuh — K2 v5: the k-way join of a round through a direct-address table, no election and no sorting
// (included by dk_join.cu; S <= 32; reference semantics: share/joinCounts/joinCounts.c:72-115).
//
// ncu of v4 (profiles/r01u_ncu_merge_v4.csv): 6.6 warp instructions per record, ~17 barriers per round, issue slots 60 %
// busy. Its cost sits in finding out which records share a k-mer (owner election over 4 table levels, ranks inside
// multi-k-mer buckets, two block scans, a byte matrix for the recurrence). v5 keeps v4's round (one window per dump staged
// by 1-D TMA bulk copies, every key below the smallest look-ahead key is complete, MONOTONE buckets over the round's key
// span) and replaces everything after it:
//   T   a table T[bucket][dump] of one byte: the window index of the FIRST record of that dump in that bucket. A dump is
//       sorted and buckets are monotone, so the records of a dump in a bucket are a run; only the run's first record
//       stores (its predecessor lies in another bucket), hence no two stores ever race and nothing is elected. A per-thread
//       bit mask says where runs start, so a reader knows where a run ends;
//   G   a bucket with >= min(r,3) dumps present is gathered by a GROUP of SP lanes, lane s = dump s: one byte load gives
//       the lane its record, the group's smallest k-mer comes from two redux.sync, the lanes holding it are the row
//       (joinCounts.c:78-83), a ballot of count >= a is the recurrence (:81-83,104) and the lanes store the row image
//       directly; lanes that took part step to their next record of the run. Rows of a bucket come out in k-mer order, a
//       k-mer occurring twice in one dump becomes two rows exactly as in the reference (no fallback kernel needed);
//   L   a bucket with fewer dumps present can never pass the filter (rec <= dumps present): one thread only counts its rows;
//   O   surviving rows take staging slots from a shared-memory counter in arrival order and remember (bucket, index in
//       bucket); one block scan over the per-bucket survivor counts turns that into the output position, so the copy-out
//       writes rows in k-mer order into the chunked matrix.
// Counts are staged by the same TMA as the keys (no global load after the copy), the next round's copies are issued as
// soon as this round's buckets are done (they fly during the scan and the copy-out), and a round has 7 block barriers.
#pragma once

#define M5_NTL 8
#define M5_NT (1 << M5_NTL)        // 256 threads
#ifndef M5_VT
#define M5_VT 11                   // consecutive records of one dump per thread (odd: conflict-free 64-bit key loads)
#endif
#ifndef M5_MB
#define M5_MB 4                    // buckets per thread: M = 1024 buckets per round
#endif
#ifndef M5_CAP
#define M5_CAP 160                 // staged (surviving) rows per round; more shrink the round
#endif
#ifndef M5_CTAS
#define M5_CTAS 3
#endif
#define M5_W (M5_NT * M5_VT)
#define M5_M (M5_MB * M5_NT)
#define M5_LB (32 * M5_MB)         // buckets a warp gathers
#define M5_EMPTY 0xFFu
#define M5_BK 32                   // the thread that keeps the partition's chunk bookkeeping (first lane of warp 1)

struct M5Cfg {
    int S;
    uint32_t RS, S4, divS4;
    uint32_t off_cnt, off_fb, off_T, off_binfo, off_bpos, off_stage, off_skey, off_smeta, off_list, off_mflag, off_clist, off_small, smem_bytes;
};
struct M5Small {
    unsigned long long bar;
    unsigned long long pivot, base;
    uint32_t scan[M5_NT / 32];
    uint32_t part, nstage, uround, ncomplex;
    uint32_t chunk_cur, rowpos, seq, prow, puni;
    uint32_t cp_nout, cp_have, cp_nb, cp_rp, cp_ccur;      // the previous round's rows: copied out behind this round's first barrier
};
template <int SP>
struct M5Samp {
    unsigned long long cur[SP], endg[SP];
    const uint64_t *ksrc[SP];      // the dump tables, read from global memory once per CTA
    const int32_t *csrc[SP];
    unsigned long long nsrc[SP];
    uint32_t koff[SP];
    uint32_t n[SP], ncm[SP];
};

__device__ __forceinline__ uint32_t m5_block_excl_scan(uint32_t v, M5Small &sm, uint32_t &total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) sm.scan[w] = x;        // the previous use of sm.scan lies several block barriers back
    __syncthreads();
    uint32_t t = lane < M5_NT / 32 ? sm.scan[lane] : 0u;
#pragma unroll
    for (int o = 1; o < M5_NT / 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, t, o); if (lane >= o) t += y; }
    total = __shfl_sync(0xffffffffu, t, M5_NT / 32 - 1);
    const uint32_t off = __shfl_sync(0xffffffffu, t, w > 0 ? w - 1 : 0);
    return (w > 0 ? off : 0u) + x - v;
}
// bounded wait: a lost transaction count becomes a trap (an error the host sees), never a hung GPU
__device__ __forceinline__ void m5_mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done = 0;
    for (uint32_t it = 0; !done; it++) {
        asm volatile(
            "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (it > (1u << 24)) __trap();
    }
}
__device__ __forceinline__ unsigned long long m5_warp_min64(unsigned long long v)
{
    const uint32_t hi = (uint32_t)(v >> 32), lo = (uint32_t)v;
    const uint32_t mh = __reduce_min_sync(0xffffffffu, hi);
    const uint32_t ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
    return ((unsigned long long)mh << 32) | ml;
}
__device__ __forceinline__ unsigned long long m5_warp_max64(unsigned long long v)
{
    const uint32_t hi = (uint32_t)(v >> 32), lo = (uint32_t)v;
    const uint32_t mh = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
    return ((unsigned long long)mh << 32) | ml;
}
// most significant bit of every byte of x that is not 0xFF
__device__ __forceinline__ uint32_t m5_present_msb(uint32_t x)
{
    const uint32_t y = ~x;
    return (((y & 0x7f7f7f7fu) + 0x7f7f7f7fu) | y) & 0x80808080u;
}

template <int TW>
__device__ __forceinline__ void m5_load_row(const unsigned char *row, uint32_t (&xw)[TW])
{
    if (TW == 4) { const uint4 v = *reinterpret_cast<const uint4 *>(row); xw[0] = v.x; xw[1] = v.y; xw[2] = v.z; xw[3] = v.w; }
    else {
        const uint4 v0 = reinterpret_cast<const uint4 *>(row)[0], v1 = reinterpret_cast<const uint4 *>(row)[1];
        xw[0] = v0.x; xw[1] = v0.y; xw[2] = v0.z; xw[3] = v0.w;
        xw[4 % TW] = v1.x; xw[5 % TW] = v1.y; xw[6 % TW] = v1.z; xw[7 % TW] = v1.w;
    }
}
template <int TW>
__device__ __forceinline__ void m5_clear_row(unsigned char *row)
{
    reinterpret_cast<uint4 *>(row)[0] = make_uint4(~0u, ~0u, ~0u, ~0u);
    if (TW == 8) reinterpret_cast<uint4 *>(row)[1] = make_uint4(~0u, ~0u, ~0u, ~0u);
}

#ifdef DK_M5_PROF
#define M5_MARK(i) do { if (tid == 0) { long long t__ = clock64(); atomicAdd(&prof[i], (uint32_t)((t__ - tprev) >> 4)); tprev = t__; } } while (0)
#else
#define M5_MARK(i) do { } while (0)
#endif

template <int LV>
__global__ void __launch_bounds__(M5_NT, M5_CTAS) k_merge5(MergeParams P, M5Cfg G, uint32_t *overflow)
{
    constexpr int SP = 1 << LV, TPSL = M5_NTL - LV, TPS = 1 << TPSL, caps = M5_W >> LV, ks = (caps + 4 + 3) & ~3;
    constexpr int TW = SP / 4;                       // 32-bit words of a table row
    constexpr int GPW = 32 / SP;                     // lane groups per warp
    constexpr int FBS = caps + 4;                    // run-start flags per dump (the look-ahead slot always starts a run)
    static_assert(caps + 4 < 255, "window index must fit one byte");
    static_assert(M5_VT <= 16 && M5_LB <= 256 && M5_MB % 2 == 0, "geometry");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    unsigned long long *bufK = reinterpret_cast<unsigned long long *>(smem_raw);              // [SP][ks] keys
    int32_t *bufC = reinterpret_cast<int32_t *>(smem_raw + G.off_cnt);                        // [SP][ks] counts
    unsigned char *fb = smem_raw + G.off_fb;                                                  // [SP][FBS] 1 = this record starts a run
    unsigned char *Tb = smem_raw + G.off_T;                                                   // [M][SP] first record
    uint16_t *binfo = reinterpret_cast<uint16_t *>(smem_raw + G.off_binfo);                   // [M] surviving rows of a bucket
    uint16_t *bpos = reinterpret_cast<uint16_t *>(smem_raw + G.off_bpos);                     // [M] output slot of its first one
    int32_t *stage = reinterpret_cast<int32_t *>(smem_raw + G.off_stage);                     // [CAP][RS] row images
    unsigned long long *skey = reinterpret_cast<unsigned long long *>(smem_raw + G.off_skey); // [CAP]
    uint32_t *smeta = reinterpret_cast<uint32_t *>(smem_raw + G.off_smeta);                   // [CAP] bucket | index << 16
    unsigned char *lists = smem_raw + G.off_list;                                             // [warps][LB]
    unsigned char *mflag = smem_raw + G.off_mflag;                                            // [M] 1 = some dump has a run here
    uint16_t *clist = reinterpret_cast<uint16_t *>(smem_raw + G.off_clist);                   // [M] complex buckets of the round
    M5Small &sm = *reinterpret_cast<M5Small *>(smem_raw + G.off_small);
    M5Samp<SP> &ss = *reinterpret_cast<M5Samp<SP> *>(smem_raw + G.off_small + sizeof(M5Small));

    const int S = G.S, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t npart = *P.npart;
    const uint32_t bar = smem_u32(&sm.bar);
    const bool owner = tid < S;                      // S <= 32: all dump owners sit in warp 0
    const int ps = tid >> TPSL, psub = tid & (TPS - 1);
    const int gl = lane & (SP - 1);
    const int gidx = lane >> LV;
    const int min_ab = P.min_abund, min_rec = P.min_rec;
    const uint32_t hth = (uint32_t)(min_rec < 3 ? (min_rec < 1 ? 1 : min_rec) : 3);   // dumps present from which a group gathers
    const uint32_t RS = G.RS;
#ifdef DK_M5_PROF
    uint32_t *prof = overflow + 1;
    long long tprev = 0;
#endif
    if (tid == 0) {
        mbar_init(bar, (uint32_t)S);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        sm.nstage = 0; sm.uround = 0; sm.ncomplex = 0;
    }
    if (owner) { ss.ksrc[tid] = P.T.keys[tid]; ss.csrc[tid] = P.T.cnts[tid]; ss.nsrc[tid] = P.T.n[tid]; }
    {   // table empty, no survivors, look-ahead slots start a run
        uint4 *t4 = reinterpret_cast<uint4 *>(Tb);
        for (int i = tid; i < M5_M * SP / 16; i += M5_NT) t4[i] = make_uint4(~0u, ~0u, ~0u, ~0u);
        for (int i = tid; i < M5_M; i += M5_NT) { binfo[i] = 0; mflag[i] = 0; }
        for (int i = tid; i < SP * FBS; i += M5_NT) fb[i] = 1;
    }
    __syncthreads();
    uint32_t phase = 0, shrink = 0;
    uint32_t o_lo = 0, o_L = 0, o_cap = 0;           // owner threads: window geometry of the copies in flight

    // owner threads: stage the next window of their dump (keys and counts, one transaction count)
    auto issue = [&]() {
        const unsigned long long cur = ss.cur[tid], endg = ss.endg[tid], nsrc = ss.nsrc[tid];
        const uint64_t *ksrc = ss.ksrc[tid];
        const int32_t *csrc = ss.csrc[tid];
        const unsigned long long a0 = cur & ~3ull;
        o_lo = (uint32_t)(cur - a0);
        o_cap = (uint32_t)caps >> shrink;
        const unsigned long long left = endg - cur;
        o_L = (uint32_t)(left < (unsigned long long)(o_cap + 1) ? left : (unsigned long long)(o_cap + 1));
        uint32_t tq = 0;
        unsigned long long mend = a0;
        if (o_L) {
            const uint32_t nq = (o_lo + o_L + 3u) >> 2;
            const unsigned long long avail = (nsrc - a0) >> 2;
            tq = (uint32_t)(avail < nq ? avail : nq);
            mend = a0 + 4ull * nq;
            if (mend > nsrc) mend = nsrc;
        }
        unsigned long long *kd = bufK + (size_t)tid * ks;
        int32_t *cd = bufC + (size_t)tid * ks;
        if (tq) {
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_expect_tx(bar, tq * 48u);
            tma_bulk_g2s(smem_u32(kd), ksrc + a0, tq * 32u, bar);
            tma_bulk_g2s(smem_u32(cd), csrc + a0, tq * 16u, bar);
        } else mbar_arrive(bar);
        for (unsigned long long g = a0 + 4ull * tq; g < mend; g++) { kd[g - a0] = ksrc[g]; cd[g - a0] = csrc[g]; }   // array tail (< 4 records)
        ss.koff[tid] = (uint32_t)tid * ks + o_lo;
    };

    for (;;) {
        if (tid == 0) sm.part = atomicAdd(P.work_counter, 1u);
        __syncthreads();
        const uint32_t part = sm.part;
        if (part >= npart) break;
        if (tid < SP) {
            ss.cur[tid] = owner ? P.starts[(size_t)part * S + tid] : 0ull;
            ss.endg[tid] = owner ? P.starts[(size_t)(part + 1) * S + tid] : 0ull;
            ss.koff[tid] = (uint32_t)tid * ks;
            ss.n[tid] = 0;
        }
        if (tid == 0) { sm.chunk_cur = CHUNK_NONE; sm.rowpos = 0; sm.seq = 0; sm.prow = 0; sm.puni = 0; }
        if (owner) issue();
        bool stop = false;
        // thread M5_BK: the chunk allocation of a round's rows is ISSUED when the round ends and CONSUMED at the top of the next
        // round (its latency hides behind the scan, the copies' wait and the pivot); the rows are copied out there too
        bool pend = false;
        uint32_t a_have = 0, a_nnew = 0, a_nb = 0, p_nout = 0;

        for (;;) {   // ---- rounds
#ifdef DK_M5_PROF
            if (tid == 0) tprev = clock64();
#endif
#ifdef M5_OVERLAP
            // the chunk bookkeeping is published one barrier early, so that warps 1..7 copy the previous round's rows out
            // WHILE warp 0 waits for the windows and computes the pivot (a single warp's dependent chain: 10 % of a round)
            if (tid == M5_BK) {   // chunk bookkeeping on warp 1: warp 0 is the round's critical path (pivot, copy issue)
                uint32_t cpn = 0;
                if (pend) {
                    for (uint32_t k = 0; k < a_nnew; k++) {
                        const uint32_t id = a_nb + k;
                        if (id < P.max_chunks) { P.chunk_part[id] = part; P.chunk_seq[id] = sm.seq + k; P.chunk_rows[id] = DK_CH; }
                    }
                    sm.cp_have = a_have; sm.cp_nb = a_nb; sm.cp_rp = a_have ? sm.rowpos : 0u; sm.cp_ccur = sm.chunk_cur;
                    cpn = p_nout;
                    const uint32_t v = (a_have ? sm.rowpos : 0u) + p_nout;
                    if (a_nnew) {
                        sm.chunk_cur = a_nb + a_nnew - 1;
                        sm.seq += a_nnew;
                        sm.rowpos = v - (a_have + a_nnew - 1) * DK_CH;
                    } else sm.rowpos = v;
                    if (sm.chunk_cur != CHUNK_NONE && sm.rowpos == DK_CH) { sm.chunk_cur = CHUNK_NONE; sm.rowpos = 0; }
                    pend = false;
                }
                sm.cp_nout = cpn;
            }
            __syncthreads();                                            // B0
            if (warp == 0) m5_mbar_wait(bar, phase);
#else
            m5_mbar_wait(bar, phase);
#endif
            M5_MARK(1);
            // R1 (warp 0 alone): pivot = smallest look-ahead key, base = smallest first key
            if (warp == 0) {
                unsigned long long candk = DK_SENTINEL, firstk = DK_SENTINEL, lastk = 0;
                uint32_t ncm = 0;
                if (owner && o_L) {
                    const unsigned long long *kw = bufK + (size_t)tid * ks + o_lo;
                    firstk = kw[0];
                    if (o_L == o_cap + 1u) { candk = kw[o_cap]; ncm = o_cap; } else ncm = o_L;
                    lastk = kw[ncm - 1];
                }
                candk = m5_warp_min64(candk);
                firstk = m5_warp_min64(firstk);
                lastk = m5_warp_max64(lastk);
                if (firstk != DK_SENTINEL && lastk + 1ull < candk) candk = lastk + 1ull;   // span of the keys staged
                if (lane < SP) ss.ncm[lane] = ncm;
                if (lane == 0) { sm.pivot = candk; sm.base = firstk; }
            }
#ifndef M5_OVERLAP
            if (tid == M5_BK) {   // chunk bookkeeping on warp 1: warp 0 is the round's critical path (pivot, copy issue)
                uint32_t cpn = 0;
                if (pend) {
                    for (uint32_t k = 0; k < a_nnew; k++) {
                        const uint32_t id = a_nb + k;
                        if (id < P.max_chunks) { P.chunk_part[id] = part; P.chunk_seq[id] = sm.seq + k; P.chunk_rows[id] = DK_CH; }
                    }
                    sm.cp_have = a_have; sm.cp_nb = a_nb; sm.cp_rp = a_have ? sm.rowpos : 0u; sm.cp_ccur = sm.chunk_cur;
                    cpn = p_nout;
                    const uint32_t v = (a_have ? sm.rowpos : 0u) + p_nout;
                    if (a_nnew) {
                        sm.chunk_cur = a_nb + a_nnew - 1;
                        sm.seq += a_nnew;
                        sm.rowpos = v - (a_have + a_nnew - 1) * DK_CH;
                    } else sm.rowpos = v;
                    if (sm.chunk_cur != CHUNK_NONE && sm.rowpos == DK_CH) { sm.chunk_cur = CHUNK_NONE; sm.rowpos = 0; }
                    pend = false;
                }
                sm.cp_nout = cpn;
            }
            M5_MARK(9);
            __syncthreads();                                            // B1
            M5_MARK(2);
            const uint32_t cw = (uint32_t)tid, cn = M5_NT;              // every thread copies out
            // the previous round's rows go out now: staging, bpos and the copy descriptor are rewritten behind B3 / B4 only
            {
#else
            const uint32_t cw = (uint32_t)tid - 32u, cn = M5_NT - 32u;  // warps 1..7 copy out, beside warp 0's pivot
            if (warp != 0) {
#endif
                const uint32_t nprev = sm.cp_nout;
                if (nprev) {
                    const uint32_t have = sm.cp_have, nb = sm.cp_nb, rp = sm.cp_rp, ccur = sm.cp_ccur;
                    if (G.S4) {
                        const uint32_t S4 = G.S4, n4 = nprev * S4;
                        for (uint32_t i = cw; i < n4; i += cn) {
                            const uint32_t rr = S4 == 1 ? i : __umulhi(i, G.divS4), c4 = i - rr * S4;
                            const uint32_t meta = smeta[rr];
                            const uint32_t v = rp + (uint32_t)bpos[meta & 0xffffu] + (meta >> 16), k = v / DK_CH, off = v % DK_CH;
                            const uint32_t chunk = (have && k == 0) ? ccur : nb + (k - have);
                            const int32_t *src = stage + rr * RS + 4u * c4;
                            if (chunk < P.max_chunks)
                                reinterpret_cast<int4 *>(P.out_cnts + ((size_t)chunk * DK_CH + off) * S)[c4] = make_int4(src[0], src[1], src[2], src[3]);
                        }
                    } else {
                        for (uint32_t rr = cw / 32u; rr < nprev; rr += cn / 32u) {
                            const uint32_t meta = smeta[rr];
                            const uint32_t v = rp + (uint32_t)bpos[meta & 0xffffu] + (meta >> 16), k = v / DK_CH, off = v % DK_CH;
                            const uint32_t chunk = (have && k == 0) ? ccur : nb + (k - have);
                            if (chunk < P.max_chunks)
                                for (uint32_t c = lane; c < (uint32_t)S; c += 32u) P.out_cnts[((size_t)chunk * DK_CH + off) * S + c] = stage[rr * RS + c];
                        }
                    }
                    for (uint32_t rr = cw; rr < nprev; rr += cn) {
                        const uint32_t meta = smeta[rr];
                        const uint32_t v = rp + (uint32_t)bpos[meta & 0xffffu] + (meta >> 16), k = v / DK_CH, off = v % DK_CH;
                        const uint32_t chunk = (have && k == 0) ? ccur : nb + (k - have);
                        if (chunk < P.max_chunks) P.out_keys[(size_t)chunk * DK_CH + off] = skey[rr];
                    }
                }
            }
#ifdef M5_OVERLAP
            if (warp != 0) m5_mbar_wait(bar, phase);                    // every thread observes the copies itself
            M5_MARK(9);
            __syncthreads();                                            // B1
            M5_MARK(2);
#endif
            phase ^= 1u;
            M5_MARK(3);
            unsigned long long pivot = sm.pivot;
            const unsigned long long base = sm.base;
            if (base == DK_SENTINEL) break;                             // every window empty: partition done
            uint32_t nout = 0;
            for (;;) {   // ---- one attempt at the round; redone with a smaller pivot when the staging area overflows
                // T: bucket of every record below the pivot; the first record of a run stores its buffer slot
                const unsigned long long span = pivot - base;          // >= 1 whenever a record lies below the pivot
                const int wsh = (64 - __clzll((long long)span)) > 16 ? (64 - __clzll((long long)span)) - 16 : 0;
                const uint32_t fmax = (uint32_t)((span - 1ull) >> wsh);
                const uint32_t scale = (fmax + 1u <= (uint32_t)M5_M) ? 65536u : (uint32_t)(((uint32_t)M5_M << 16) / (fmax + 1u));
                uint32_t cntv = 0;
                {
                    const uint32_t ncm = ss.ncm[ps];
                    const uint32_t lo_ps = ss.koff[ps] - (uint32_t)ps * ks;
                    const unsigned long long *kw = bufK + ss.koff[ps];
                    const uint32_t i0 = (uint32_t)psub * M5_VT;
                    uint32_t bk[M5_VT], vm = 0;
#pragma unroll
                    for (int j = 0; j < M5_VT; j++) {
                        const unsigned long long key = kw[i0 + j];                       // slots past the window: stale, never used
                        vm |= ((i0 + j < ncm && key < pivot) ? 1u : 0u) << j;            // keys are sorted: the valid records are a prefix
                        const uint32_t f = (uint32_t)((key - base) >> wsh);
                        bk[j] = (f * scale) >> 16;
                    }
                    cntv = __popc(vm);
                    {   // consumable prefix of the dump: exactly one thread of the dump sees its end
                        const bool v0 = vm & 1u, vl = (vm >> (M5_VT - 1)) & 1u;
                        const bool pvl = __shfl_up_sync(0xffffffffu, vl ? 1u : 0u, 1) != 0u;
                        if (v0 && !vl) ss.n[ps] = i0 + cntv;
                        else if (vl && psub == TPS - 1) ss.n[ps] = i0 + M5_VT;
                        else if (!v0 && (psub == 0 || pvl)) ss.n[ps] = i0;
                    }
                    const uint32_t prev = __shfl_up_sync(0xffffffffu, bk[M5_VT - 1], 1);
                    const uint32_t sl0 = lo_ps + i0;                    // buffer slot of my first record
                    unsigned char *fbp = fb + ps * FBS + sl0;
#pragma unroll
                    for (int j = 0; j < M5_VT; j++) {
                        const bool valid = (vm >> j) & 1u;
                        const bool first = j ? (bk[j] != bk[j - 1]) : (psub == 0 || bk[0] != prev);
                        if (valid && first) Tb[bk[j] * SP + (uint32_t)ps] = (unsigned char)(sl0 + j);
                        if (valid && !first) mflag[bk[j]] = 1;          // this bucket holds a run: gathered by a lane group
                        fbp[j] = (first || !valid) ? 1 : 0;
                    }
                    if (psub == TPS - 1) fbp[M5_VT] = 1;                // the look-ahead slot always starts a run
                }
                if (!__syncthreads_or(cntv != 0u)) {                    // B3. No record below the pivot: a window full of ONE
                    if (tid == 0) atomicExch(overflow, 1u);            // k-mer (its copies fill a dump's window): v3's case
                    stop = true;
                    break;
                }
                M5_MARK(4);
                // ---- C1: buckets of this warp: rows counted (L) or listed (G)
                uint32_t urows = 0;
                {
                    unsigned char *lst = lists + warp * M5_LB;
                    const uint32_t wb0 = (uint32_t)warp * M5_LB;
                    uint32_t hcnt = 0, lcnt = 0;                        // heavy list from the front of lst, light list from its back
#pragma unroll
                    for (int q = 0; q < M5_MB; q++) {                   // classification only: straight-line code
                        const uint32_t bq = wb0 + q * 32 + lane;
                        uint32_t xw[TW], np = 0;
                        m5_load_row<TW>(Tb + (size_t)bq * SP, xw);
#pragma unroll
                        for (int w = 0; w < TW; w++) np += __popc(m5_present_msb(xw[w]));
                        const bool runs = mflag[bq] != 0;
                        const bool heavy = np >= hth;
                        const bool easy = !heavy && np == 1 && !runs;   // one record: one row that cannot pass (rec <= 1 < r)
                        const bool light = np != 0 && !heavy && !easy;
                        urows += easy ? 1u : 0u;
                        const uint32_t hb = __ballot_sync(0xffffffffu, heavy), lb = __ballot_sync(0xffffffffu, light);
                        const uint32_t below = (1u << lane) - 1u;
                        if (heavy) lst[hcnt + __popc(hb & below)] = (unsigned char)(q * 32 + lane);
                        if (light) lst[M5_LB - 1 - (lcnt + __popc(lb & below))] = (unsigned char)(q * 32 + lane);
                        hcnt += __popc(hb); lcnt += __popc(lb);
                    }
                    __syncwarp();
                    // L: buckets with fewer dumps present than a passing row needs (rec <= np < r, joinCounts.c:104) only have
                    // their rows COUNTED; one lane per listed bucket
                    for (uint32_t e0 = 0; e0 < lcnt; e0 += 32) {
                        if (e0 + lane < lcnt) {
                            const uint32_t bq = wb0 + lst[M5_LB - 1 - (e0 + lane)];
                            unsigned char *trow = Tb + (size_t)bq * SP;
                            uint32_t xw[TW], tw[TW], np = 0, pm = 0;
                            m5_load_row<TW>(trow, xw);
#pragma unroll
                            for (int w = 0; w < TW; w++) {
                                tw[w] = m5_present_msb(xw[w]); np += __popc(tw[w]);
                                pm |= (((tw[w] >> 7) * 0x10204080u) >> 28) << (4 * w);
                            }
                            const bool runs = mflag[bq] != 0;
                            if (np == 2 && !runs) {                     // two records of two dumps: one row if they hold one k-mer
                                const uint32_t s1 = __ffs(pm) - 1, s2 = 31 - __clz(pm);
                                const uint32_t i1 = trow[s1], i2 = trow[s2];
                                urows += bufK[s1 * ks + i1] == bufK[s2 * ks + i2] ? 1u : 2u;
                            } else if (np == 1) {                       // a run of one dump: as many rows as records
                                const uint32_t s1 = __ffs(pm) - 1, i1 = trow[s1];
                                const unsigned char *f = fb + s1 * FBS + i1;
                                uint32_t rows = 1;
                                while (!f[rows]) { if (++rows > (uint32_t)FBS) __trap(); }   // (a run ends at the look-ahead slot at the latest)
                                urows += rows;
                            } else {                                    // walk the rows in k-mer order (cursors in the table row)
                                for (uint32_t guard = 0; pm; guard++) {
                                    if (guard > (uint32_t)(SP * FBS)) __trap();
                                    unsigned long long kmin = DK_SENTINEL;
                                    for (uint32_t m = pm; m; m &= m - 1) {
                                        const uint32_t s2 = __ffs(m) - 1;
                                        const unsigned long long k = bufK[s2 * ks + trow[s2]];
                                        kmin = k < kmin ? k : kmin;
                                    }
                                    for (uint32_t m = pm; m; m &= m - 1) {
                                        const uint32_t s2 = __ffs(m) - 1, i = trow[s2];
                                        if (bufK[s2 * ks + i] == kmin) {
                                            if (!fb[s2 * FBS + i + 1]) trow[s2] = (unsigned char)(i + 1); else pm &= ~(1u << s2);
                                        }
                                    }
                                    urows++;
                                }
                            }
                        }
                    }
                    __syncwarp();
                    M5_MARK(5);
                    // ---- C2: one thread per listed bucket: the usual bucket holds ONE k-mer and no run. Every dump's record
                    // is compared with the first present one, the counts are gathered, the row is judged and staged — straight
                    // line code over the dumps. Anything else goes to the CTA's list for the lane groups.
                    for (uint32_t e0 = 0; e0 < hcnt; e0 += 32) {
                        if (e0 + lane < hcnt) {
                            const uint32_t b = wb0 + lst[e0 + lane];
                            unsigned char *trow = Tb + (size_t)b * SP;
                            uint32_t xw[TW];
                            m5_load_row<TW>(trow, xw);
                            bool cplx = mflag[b] != 0;
                            uint32_t sref = 0, iref = 0;
#pragma unroll
                            for (int w = TW - 1; w >= 0; w--) {
                                const uint32_t t = m5_present_msb(xw[w]);
                                if (t) { const uint32_t by = (uint32_t)(__ffs(t) - 1) >> 3; sref = 4u * w + by; iref = (xw[w] >> (8u * by)) & 0xffu; }
                            }
                            const unsigned long long kref = bufK[sref * ks + iref];
                            int32_t cv[SP];
                            int rec = 0;
#pragma unroll
                            for (int s2 = 0; s2 < SP; s2++) {
                                const uint32_t sl = (xw[s2 >> 2] >> (8 * (s2 & 3))) & 0xffu;
                                const bool p = sl != M5_EMPTY;
                                const uint32_t sc = p ? sl : 0u;
                                cplx |= p && bufK[s2 * ks + sc] != kref;
                                const int32_t c = bufC[s2 * ks + sc];
                                rec += (p && c >= min_ab) ? 1 : 0;                                   // joinCounts.c:81-83
                                cv[s2] = p ? c : 0;                                                  // :96
                            }
                            if (cplx) {
                                const uint32_t pos = atomicAdd(&sm.ncomplex, 1u);
                                clist[pos] = (uint16_t)b;
                            } else {
                                const bool pass = rec >= min_rec;                                    // :104
                                if (pass) {
                                    const uint32_t slot = atomicAdd(&sm.nstage, 1u);
                                    if (slot < (uint32_t)M5_CAP) {
                                        int32_t *dst = stage + slot * RS;
#pragma unroll
                                        for (int s2 = 0; s2 < SP; s2++) if (s2 < S) dst[s2] = cv[s2];
                                        skey[slot] = kref;
                                        smeta[slot] = b;
                                    }
                                }
                                binfo[b] = pass ? 1 : 0;
                                urows++;
                            }
                        }
                    }
                }
                __syncthreads();                                        // B3b: the CTA's list of complex buckets is complete
                M5_MARK(6);
                {
                    // ---- C3: lane groups, lane gl = dump gl, walk the rows of a complex bucket in k-mer order
                    const unsigned long long *kg = bufK + gl * ks;
                    const int32_t *cg = bufC + gl * ks;
                    const unsigned char *fg = fb + gl * FBS + 1;
                    const uint32_t gbits = SP == 32 ? 0xffffffffu : ((1u << SP) - 1u);
                    const int gsh = lane & ~(SP - 1);
                    const uint32_t ncx = sm.ncomplex;
                    for (uint32_t h = (uint32_t)warp * GPW; h < ncx; h += (M5_NT / 32) * GPW) {
                        const bool act = h + gidx < ncx;
                        const uint32_t b = clist[act ? h + gidx : h];
                        uint32_t i = act ? (uint32_t)Tb[(size_t)b * SP + gl] : M5_EMPTY;
                        uint32_t rows = 0, npass = 0;
                        bool present = i != M5_EMPTY;
                        uint32_t pw = __ballot_sync(0xffffffffu, present);
                        for (uint32_t guard = 0; pw; guard++) {
                            if (guard > (uint32_t)(SP * FBS)) __trap();                              // (a bucket holds < SP * FBS records)
                            unsigned long long key = DK_SENTINEL;
                            if (present) key = kg[i];
                            const uint32_t hi = (uint32_t)(key >> 32), lo32 = (uint32_t)key;
                            const uint32_t gp = (pw >> gsh) & gbits;                                 // lanes of my group holding a record
                            // smallest k-mer of the group (:97-100). One full-mask redux per group: a redux over half a warp
                            // compiles to a divergent slow path
                            uint32_t mh, ml;
                            if (GPW == 1) {
                                mh = __reduce_min_sync(0xffffffffu, hi);
                                ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo32 : 0xffffffffu);
                            } else {
                                const uint32_t h0 = __reduce_min_sync(0xffffffffu, gidx == 0 ? hi : 0xffffffffu);
                                const uint32_t h1 = __reduce_min_sync(0xffffffffu, gidx == 1 ? hi : 0xffffffffu);
                                mh = gidx ? h1 : h0;
                                const uint32_t l0 = __reduce_min_sync(0xffffffffu, (gidx == 0 && hi == mh) ? lo32 : 0xffffffffu);
                                const uint32_t l1 = __reduce_min_sync(0xffffffffu, (gidx == 1 && hi == mh) ? lo32 : 0xffffffffu);
                                ml = gidx ? l1 : l0;
                            }
                            const bool member = present && hi == mh && lo32 == ml;                   // joinCounts.c:78
                            int32_t c = 0;
                            if (member) c = cg[i];
                            const uint32_t aw = __ballot_sync(0xffffffffu, member && c >= min_ab);   // :81-83
                            const bool pass = gp && (int)__popc((aw >> gsh) & gbits) >= min_rec;     // :104
                            uint32_t slot = 0;
                            if (pass && gl == 0) slot = atomicAdd(&sm.nstage, 1u);
                            slot = __shfl_sync(0xffffffffu, slot, gsh);
                            if (pass && slot < (uint32_t)M5_CAP) {
                                if (gl < S) stage[slot * RS + (uint32_t)gl] = c;
                                if (gl == 0) { skey[slot] = ((unsigned long long)mh << 32) | ml; smeta[slot] = b | (npass << 16); }
                            }
                            npass += pass ? 1u : 0u;
                            rows += gp ? 1u : 0u;
                            if (member) { present = !fg[i]; i++; }                                   // next record of the run, if any
                            pw = __ballot_sync(0xffffffffu, present);
                        }
                        if (act && gl == 0) { binfo[b] = (uint16_t)npass; urows += rows; }
                    }
                }
                {
                    const uint32_t ut = __reduce_add_sync(0xffffffffu, urows);
                    if (lane == 0 && ut) atomicAdd(&sm.uround, ut);
                }
                __syncthreads();                                        // B4
                {   // every bucket has been read: table and run flags are emptied for the next pass in one sweep
                    uint4 *t4 = reinterpret_cast<uint4 *>(Tb);
#pragma unroll
                    for (int i = 0; i < M5_M * SP / 16 / M5_NT; i++) t4[tid + i * M5_NT] = make_uint4(~0u, ~0u, ~0u, ~0u);
                    if (tid < M5_M / 16) reinterpret_cast<uint4 *>(mflag)[tid] = make_uint4(0u, 0u, 0u, 0u);
                }
                M5_MARK(7);
                const uint32_t nst = sm.nstage;
                if (nst > (uint32_t)M5_CAP) {                           // more survivors than staging slots: smaller round
                    for (int i = tid; i < M5_M; i += M5_NT) binfo[i] = 0;
                    if (span <= 1ull) {                                 // one k-mer, > CAP copies of it in a dump: v3's case
                        if (tid == 0) atomicExch(overflow, 1u);
                        stop = true;
                        break;
                    }
                    unsigned long long ns = (span / nst) * (unsigned long long)(M5_CAP * 3 / 4);
                    if (ns < 1ull) ns = 1ull;
                    if (ns >= span) ns = span - 1ull;
                    pivot = base + ns;
                    if (shrink < 4u) shrink++;
                    __syncthreads();                                    // every thread has read the counters
                    if (tid == 0) { sm.nstage = 0; sm.uround = 0; sm.ncomplex = 0; }
                    continue;                                           // (they are next touched behind the redone pass's barrier)
                }
                nout = nst;
                break;
            }
            if (stop) break;
            if (shrink && nout < (uint32_t)M5_CAP / 4u) shrink--;
            // the next round's windows start flying now; the chunk allocation of this round's rows too
            if (owner) { ss.cur[tid] += ss.n[tid]; issue(); }
            if (tid == M5_BK) {
                a_have = (sm.chunk_cur != CHUNK_NONE) ? 1u : 0u;
                const uint32_t slots = a_have ? (DK_CH - sm.rowpos) : 0u;
                a_nnew = nout > slots ? (nout - slots + DK_CH - 1) / DK_CH : 0u;
                a_nb = 0;
                if (a_nnew) a_nb = atomicAdd(P.chunk_counter, a_nnew);  // consumed at the top of the next round
                p_nout = nout;
                pend = nout != 0u;
            }
            // O: output slot of every bucket's first surviving row
            {
                uint32_t *bi = reinterpret_cast<uint32_t *>(binfo) + (size_t)tid * (M5_MB / 2);
                uint32_t *bp = reinterpret_cast<uint32_t *>(bpos) + (size_t)tid * (M5_MB / 2);
                uint32_t v[M5_MB / 2], mine = 0;
#pragma unroll
                for (int q = 0; q < M5_MB / 2; q++) { v[q] = bi[q]; mine += (v[q] & 0xffffu) + (v[q] >> 16); bi[q] = 0u; }
                uint32_t tot;
                uint32_t run = m5_block_excl_scan(mine, sm, tot);       // (one block barrier inside)
#pragma unroll
                for (int q = 0; q < M5_MB / 2; q++) {
                    const uint32_t p0 = run, p1 = run + (v[q] & 0xffffu);
                    bp[q] = p0 | (p1 << 16);
                    run = p1 + (v[q] >> 16);
                }
            }
            M5_MARK(8);
            if (tid == 0) {     // behind the scan's barrier every thread has read this round's counters
                sm.prow += nout; sm.puni += sm.uround;
                sm.nstage = 0; sm.uround = 0; sm.ncomplex = 0;
            }
        }
        __syncthreads();                                                // (no copy is in flight here: a failing attempt issues none;
                                                                        //  the last round's rows went out behind B1 above)
        if (tid == 0) {
            if (sm.chunk_cur != CHUNK_NONE && sm.chunk_cur < P.max_chunks) P.chunk_rows[sm.chunk_cur] = sm.rowpos;
            P.part_nchunks[part] = sm.seq;
            P.part_rows[part] = sm.prow;
            P.part_union[part] = sm.puni;
            sm.nstage = 0; sm.uround = 0; sm.ncomplex = 0;
        }
        __syncthreads();
    }
}

// host side: geometry for S samples
static bool m5_make_cfg(int S, size_t smem_budget, M5Cfg &G, int &lv)
{
    if (S < 1 || S > 32) return false;
    lv = S <= 16 ? 4 : 5;
    const int sp = 1 << lv, caps = M5_W >> lv, ks = (caps + 4 + 3) & ~3;
    G.S = S;
    G.RS = (uint32_t)(S | 1);
    G.S4 = (S % 4 == 0) ? S / 4 : 0;
    G.divS4 = G.S4 > 1 ? (uint32_t)(((1ull << 32) + G.S4 - 1) / G.S4) : 0;
    size_t off = dk_round_up((size_t)sp * ks * 8, 128);
    G.off_cnt = (uint32_t)off; off += dk_round_up((size_t)sp * ks * 4, 128);
    G.off_fb = (uint32_t)off; off += dk_round_up((size_t)sp * (caps + 4), 128);
    G.off_T = (uint32_t)off; off += (size_t)M5_M * sp;
    G.off_binfo = (uint32_t)off; off += (size_t)M5_M * 2;
    G.off_bpos = (uint32_t)off; off += (size_t)M5_M * 2;
    G.off_stage = (uint32_t)off; off += dk_round_up((size_t)M5_CAP * G.RS * 4, 128);
    G.off_skey = (uint32_t)off; off += (size_t)M5_CAP * 8;
    G.off_smeta = (uint32_t)off; off += dk_round_up((size_t)M5_CAP * 4, 128);
    G.off_list = (uint32_t)off; off += (size_t)(M5_NT / 32) * M5_LB;
    G.off_mflag = (uint32_t)off; off += (size_t)M5_M;
    G.off_clist = (uint32_t)off; off += (size_t)M5_M * 2;
    G.off_small = (uint32_t)off; off += dk_round_up(sizeof(M5Small) + (size_t)sp * 56 + 16, 128);
    G.smem_bytes = (uint32_t)off;
    return off <= smem_budget;
}
