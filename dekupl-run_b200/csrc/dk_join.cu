// dk_join.cu — joinCounts on the GPU: k-way merge of S sorted (key,count) dumps
// with the recurrence filter fused in (reference: share/joinCounts/joinCounts.c:72-115).
//
// Design:
//  K1  key-space partitioning: NB uniform buckets over [kmin,kmax]; per bucket the
//      total record count below its lower edge (S binary searches); partitions are
//      runs of buckets holding ~PART_TARGET records; per partition S start cursors.
//  K2  persistent merge kernel. A "stream" is a group of GW lanes (16 or 32) that
//      walks one partition; lane l owns samples l, l+GW, ... Each sample has a small
//      ring buffer in shared memory refilled by 1-D TMA bulk copies
//      (cp.async.bulk + mbarrier complete_tx) issued by the owning lane. Per output
//      row: group-min of the head keys (2x redux.sync on the 32-bit halves), ballot
//      of (cnt >= a) for the recurrence filter, coalesced row-major store.
//      Output rows go to DK_CH-row chunks taken from a bump allocator; a tiny
//      post-pass orders the chunks (reference order = partition order).
#include "dk_common.cuh"

// ---------------------------------------------------------------- PTX helpers ---
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity)
        : "memory");
}
// 1-D TMA bulk copy global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void tma_bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

// ------------------------------------------------------------- K1 partitioning ---
struct JoinTables {
    const uint64_t *const *keys;   // [S] device pointers
    const int32_t *const *cnts;    // [S]
    const uint64_t *n;             // [S] record counts (device)
};

// range[0]=kmin, range[1]=kmax over all samples, range[2]=shift, range[3]=total records
__global__ void k_key_range(JoinTables T, int S, int nb_log2, uint64_t *range)
{
    uint64_t mn = ~0ull, mx = 0, tot = 0;
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
        uint64_t ns = T.n[s];
        tot += ns;
        if (ns) {
            uint64_t a = T.keys[s][0], b = T.keys[s][ns - 1];
            mn = a < mn ? a : mn;
            mx = b > mx ? b : mx;
        }
    }
    __shared__ uint64_t smn[32], smx[32], stot[32];
    for (int o = 16; o; o >>= 1) {
        uint64_t a = __shfl_down_sync(0xffffffffu, mn, o), b = __shfl_down_sync(0xffffffffu, mx, o);
        mn = a < mn ? a : mn;
        mx = b > mx ? b : mx;
        tot += __shfl_down_sync(0xffffffffu, tot, o);
    }
    if ((threadIdx.x & 31) == 0) { smn[threadIdx.x >> 5] = mn; smx[threadIdx.x >> 5] = mx; stot[threadIdx.x >> 5] = tot; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (blockDim.x + 31) / 32; w++) {
            mn = smn[w] < mn ? smn[w] : mn;
            mx = smx[w] > mx ? smx[w] : mx;
            tot += stot[w];
        }
        if (tot == 0) { mn = 0; mx = 0; }
        uint64_t span = mx - mn;
        int shift = 0;
        while (shift < 64 && (span >> shift) >= (1ull << nb_log2)) shift++;
        range[0] = mn; range[1] = mx; range[2] = (uint64_t)shift; range[3] = tot;
    }
}

__device__ __forceinline__ uint64_t lower_bound_u64(const uint64_t *a, uint64_t n, uint64_t key)
{
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        uint64_t mid = (lo + hi) >> 1;
        if (__ldg(a + mid) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// below[b] = #records (all samples) with key < lower edge of bucket b, b in [0, NB]
__global__ void k_bucket_below(JoinTables T, int S, uint32_t NB, const uint64_t *range, uint64_t *below)
{
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > NB) return;
    uint64_t kmin = range[0], shift = range[2];
    uint64_t tot = 0;
    if (b == NB) tot = range[3];
    else if (b > 0) {
        // edge may overflow past kmax for unused high buckets: saturate
        uint64_t off = (shift >= 64) ? 0 : ((uint64_t)b << shift);
        bool ovf = shift < 64 && shift > 0 && (off >> shift) != b;
        uint64_t edge = kmin + off;
        if (ovf || edge < kmin) tot = range[3];
        else for (int s = 0; s < S; s++) tot += lower_bound_u64(T.keys[s], T.n[s], edge);
    }
    below[b] = tot;
}

// flag[b]=1 iff bucket b starts a new partition: partition id = below[b] / target
__global__ void k_part_flags(const uint64_t *below, uint32_t NB, uint64_t target, uint32_t *flag)
{
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= NB) return;
    uint32_t f;
    if (b == 0) f = 1;
    else {
        // bucket b starts a partition iff it is non-empty-from-here and crosses a target multiple
        uint64_t p0 = below[b - 1] / target, p1 = below[b] / target;
        f = (p1 != p0) ? 1u : 0u;
    }
    flag[b] = f;
}

// starts[(k)*S+s] for partition k (k = rank of flagged bucket); last row = n[s]
__global__ void k_part_starts(JoinTables T, int S, uint32_t NB, const uint64_t *range, const uint32_t *flag,
                              const uint32_t *rank /*exclusive scan of flag, rank[NB]=NPART*/, uint64_t *starts,
                              uint32_t *npart_out)
{
    uint32_t b = blockIdx.x;
    uint32_t NP = rank[NB];
    if (b == NB) {
        for (int s = threadIdx.x; s < S; s += blockDim.x) starts[(size_t)NP * S + s] = T.n[s];
        if (threadIdx.x == 0) *npart_out = NP;
        return;
    }
    if (!flag[b]) return;
    uint32_t k = rank[b];
    uint64_t kmin = range[0], shift = range[2];
    uint64_t off = (shift >= 64) ? 0 : ((uint64_t)b << shift);
    uint64_t edge = kmin + off;
    bool ovf = (shift < 64 && shift > 0 && (off >> shift) != b) || edge < kmin;
    for (int s = threadIdx.x; s < S; s += blockDim.x)
        starts[(size_t)k * S + s] = (b == 0) ? 0 : (ovf ? T.n[s] : lower_bound_u64(T.keys[s], T.n[s], edge));
}

// ------------------------------------------------------------------ K2 merge ---
struct MergeParams {
    JoinTables T;
    int S, min_rec, min_abund;
    const uint64_t *starts;      // [(NP+1)*S]
    const uint32_t *npart;       // device scalar
    uint32_t *work_counter;      // dynamic partition scheduler
    uint64_t *out_keys;          // [max_chunks*DK_CH]
    int32_t *out_cnts;           // [max_chunks*DK_CH*S]
    uint32_t *chunk_rows, *chunk_part, *chunk_seq;   // [max_chunks]
    uint32_t *chunk_counter;     // bump allocator (keeps counting past max_chunks => exact size for a retry)
    uint32_t max_chunks;
    uint32_t *part_nchunks, *part_rows, *part_union; // [NPmax]
};

#define CHUNK_NONE 0xFFFFFFFFu
#define CHUNK_OVER 0xFFFFFFFEu

template <int GW, int SPL, int CHK, int NSLOT>
__global__ void __launch_bounds__(128) k_merge(MergeParams P)
{
    constexpr int RING = CHK * NSLOT;
    constexpr int GPW = 32 / GW;
    extern __shared__ __align__(128) unsigned char smem[];
    const int S = P.S;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lig = lane % GW, grp = lane / GW;
    const unsigned gmask = (GW == 32) ? 0xffffffffu : (((1u << GW) - 1u) << (grp * GW));
    const size_t stream_bytes = ((size_t)S * (RING * 12 + NSLOT * 8) + 15) & ~(size_t)15;
    unsigned char *base = smem + (size_t)(warp * GPW + grp) * stream_bytes;
    uint64_t *rkeys = (uint64_t *)base;
    int32_t *rcnts = (int32_t *)(base + (size_t)S * RING * 8);
    uint64_t *bars = (uint64_t *)(base + (size_t)S * RING * 12);

#pragma unroll
    for (int j = 0; j < SPL; j++) {
        int s = lig + GW * j;
        if (s < S)
            for (int k = 0; k < NSLOT; k++) mbar_init(smem_u32(&bars[s * NSLOT + k]), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();

    uint64_t kcur[SPL], pos[SPL], end[SPL], nrec[SPL];
    int32_t ccur[SPL];
    uint32_t phase[SPL];
#pragma unroll
    for (int j = 0; j < SPL; j++) { kcur[j] = DK_SENTINEL; pos[j] = end[j] = nrec[j] = 0; ccur[j] = 0; phase[j] = 0; }

    const uint32_t npart = *P.npart;
    bool done = false, active = false;
    uint32_t part = 0, rowpos = 0, chunk = CHUNK_NONE, seq = 0, prow = 0, puni = 0;

    // issue the load of chunk c of sample s into its ring slot
    auto issue_chunk = [&](int s, uint64_t c, uint64_t Ns) {
        uint32_t slot = (uint32_t)(c % NSLOT);
        uint64_t g0 = c * CHK;
        uint32_t bar = smem_u32(&bars[s * NSLOT + slot]);
        uint64_t *kd = &rkeys[s * RING + slot * CHK];
        int32_t *cd = &rcnts[s * RING + slot * CHK];
        const uint64_t *ksrc = P.T.keys[s];
        const int32_t *csrc = P.T.cnts[s];
        if (g0 + CHK <= Ns) {
            mbar_expect_tx(bar, CHK * 12);
            tma_bulk_g2s(smem_u32(kd), ksrc + g0, CHK * 8, bar);
            tma_bulk_g2s(smem_u32(cd), csrc + g0, CHK * 4, bar);
        } else {   // array tail: a bulk copy would read past the end; copy by hand
            for (uint64_t i = g0; i < Ns; i++) { kd[i - g0] = ksrc[i]; cd[i - g0] = csrc[i]; }
            mbar_arrive(bar);
        }
    };

    while (__any_sync(0xffffffffu, !done)) {
        if (!done && !active) {
            uint32_t p = 0;
            if (lig == 0) p = atomicAdd(P.work_counter, 1u);
            p = __shfl_sync(gmask, p, 0, GW);
            if (p >= npart) done = true;
            else {
                part = p; active = true; seq = 0; prow = 0; puni = 0; rowpos = 0; chunk = CHUNK_NONE;
#pragma unroll
                for (int j = 0; j < SPL; j++) {
                    int s = lig + GW * j;
                    kcur[j] = DK_SENTINEL;
                    if (s < S) {
                        pos[j] = P.starts[(size_t)p * S + s];
                        end[j] = P.starts[(size_t)(p + 1) * S + s];
                        nrec[j] = P.T.n[s];
                        if (pos[j] < end[j]) {
                            uint64_t c0 = pos[j] / CHK;
                            for (int k = 0; k < NSLOT; k++)
                                if ((c0 + k) * CHK < end[j]) issue_chunk(s, c0 + k, nrec[j]);
                        }
                    } else pos[j] = end[j] = 0;
                }
#pragma unroll
                for (int j = 0; j < SPL; j++) {
                    int s = lig + GW * j;
                    if (s < S && pos[j] < end[j]) {
                        uint32_t slot = (uint32_t)((pos[j] / CHK) % NSLOT);
                        mbar_wait(smem_u32(&bars[s * NSLOT + slot]), (phase[j] >> slot) & 1u);
                        phase[j] ^= 1u << slot;
                        uint32_t r = (uint32_t)(pos[j] % RING);
                        kcur[j] = rkeys[s * RING + r];
                        ccur[j] = rcnts[s * RING + r];
                    }
                }
            }
        }
        if (active) {
            uint64_t lk = kcur[0];
#pragma unroll
            for (int j = 1; j < SPL; j++) lk = kcur[j] < lk ? kcur[j] : lk;
            uint32_t hi = (uint32_t)(lk >> 32), lo = (uint32_t)lk;
            uint32_t mhi = __reduce_min_sync(gmask, hi);
            uint32_t cand = (hi == mhi) ? lo : 0xffffffffu;
            uint32_t mlo = __reduce_min_sync(gmask, cand);
            uint64_t m = ((uint64_t)mhi << 32) | mlo;
            if (m == DK_SENTINEL) {      // partition exhausted
                if (lig == 0) {
                    if (chunk < CHUNK_OVER && rowpos > 0) P.chunk_rows[chunk] = rowpos;
                    P.part_nchunks[part] = seq;
                    P.part_rows[part] = prow;
                    P.part_union[part] = puni;
                }
                active = false;
            } else {
                int rec = 0;
                int32_t c[SPL];
#pragma unroll
                for (int j = 0; j < SPL; j++) {
                    bool hit = (kcur[j] == m);
                    c[j] = hit ? ccur[j] : 0;
                    rec += __popc(__ballot_sync(gmask, hit && ccur[j] >= P.min_abund));   // joinCounts.c:81-83
                }
                puni++;
                if (rec >= P.min_rec) {                                                    // joinCounts.c:104
                    if (chunk == CHUNK_NONE) {
                        uint32_t id = 0;
                        if (lig == 0) {
                            id = atomicAdd(P.chunk_counter, 1u);
                            if (id < P.max_chunks) { P.chunk_part[id] = part; P.chunk_seq[id] = seq; }
                        }
                        id = __shfl_sync(gmask, id, 0, GW);
                        chunk = id < P.max_chunks ? id : CHUNK_OVER;
                        seq++;
                    }
                    if (chunk != CHUNK_OVER) {
                        size_t phys = (size_t)chunk * DK_CH + rowpos;
#pragma unroll
                        for (int j = 0; j < SPL; j++) {
                            int s = lig + GW * j;
                            if (s < S) P.out_cnts[phys * S + s] = c[j];
                        }
                        if (lig == 0) P.out_keys[phys] = m;
                    }
                    rowpos++; prow++;
                    if (rowpos == DK_CH) {
                        if (lig == 0 && chunk != CHUNK_OVER) P.chunk_rows[chunk] = DK_CH;
                        chunk = CHUNK_NONE; rowpos = 0;
                    }
                }
                // advance every cursor that sat on the minimum (joinCounts.c:84-92)
#pragma unroll
                for (int j = 0; j < SPL; j++) {
                    if (kcur[j] == m) {
                        int s = lig + GW * j;
                        pos[j]++;
                        if (pos[j] >= end[j]) kcur[j] = DK_SENTINEL;
                        else {
                            if ((pos[j] % CHK) == 0) {
                                uint64_t cc = pos[j] / CHK;
                                uint32_t slot = (uint32_t)(cc % NSLOT);
                                mbar_wait(smem_u32(&bars[s * NSLOT + slot]), (phase[j] >> slot) & 1u);
                                phase[j] ^= 1u << slot;
                                uint64_t cn = cc - 1 + NSLOT;
                                if (cn * CHK < end[j]) issue_chunk(s, cn, nrec[j]);
                            }
                            uint32_t r = (uint32_t)(pos[j] % RING);
                            kcur[j] = rkeys[s * RING + r];
                            ccur[j] = rcnts[s * RING + r];
                        }
                    }
                }
            }
        }
    }
}

#include "dk_merge3.cuh"
#include "dk_merge4.cuh"
#include "dk_merge5.cuh"

// -------------------------------------------------------- chunk bookkeeping ---
// totals[0]=R, totals[1]=U, totals[2]=chunks requested, totals[3]=npart
__global__ void k_join_totals(const uint32_t *part_rows, const uint32_t *part_union, const uint32_t *npart,
                              const uint32_t *chunk_counter, const uint32_t *overflow, uint64_t *totals)
{
    __shared__ uint64_t sr[32], su[32];
    uint32_t NP = *npart;
    uint64_t r = 0, u = 0;
    for (uint32_t p = threadIdx.x; p < NP; p += blockDim.x) { r += part_rows[p]; u += part_union[p]; }
    for (int o = 16; o; o >>= 1) { r += __shfl_down_sync(0xffffffffu, r, o); u += __shfl_down_sync(0xffffffffu, u, o); }
    if ((threadIdx.x & 31) == 0) { sr[threadIdx.x >> 5] = r; su[threadIdx.x >> 5] = u; }
    __syncthreads();
    if (threadIdx.x == 0) {
        r = 0; u = 0;
        for (int w = 0; w < (blockDim.x + 31) / 32; w++) { r += sr[w]; u += su[w]; }
        totals[0] = r; totals[1] = u; totals[2] = *chunk_counter; totals[3] = NP; totals[4] = *overflow;
    }
}
__global__ void k_chunk_order(const uint32_t *chunk_part, const uint32_t *chunk_seq, const uint32_t *part_choff,
                              uint32_t n_chunks, uint32_t *order)
{
    uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id < n_chunks) order[part_choff[chunk_part[id]] + chunk_seq[id]] = id;
}
__global__ void k_chunk_rows_in_order(const uint32_t *order, const uint32_t *chunk_rows, uint32_t n_chunks, uint64_t *out)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_chunks) out[i] = chunk_rows[order[i]];
}
// the reference's extra all-zero row for min_rec <= 0 (joinCounts.c:104-114)
__global__ void k_stale_tail(uint64_t *keys, int32_t *cnts, uint32_t *chunk_rows, uint32_t *order, uint64_t *lbase,
                             uint32_t chunk, uint32_t opos, int S, const uint64_t *range)
{
    for (int s = threadIdx.x; s < S; s += blockDim.x) cnts[(size_t)chunk * DK_CH * S + s] = 0;
    if (threadIdx.x == 0) {
        keys[(size_t)chunk * DK_CH] = range[1];
        chunk_rows[chunk] = 1;
        order[opos] = chunk;
        lbase[opos + 1] = lbase[opos] + 1;
    }
}

// records and distinct k-mers of partitions [0, n): the probe that picks the merge kernel (see dk_join_impl)
__global__ void k_probe_stats(const uint32_t *part_union, const uint64_t *starts, int S, uint32_t n, uint64_t *out)
{
    __shared__ uint64_t su[32];
    uint64_t u = 0;
    for (uint32_t p = threadIdx.x; p < n; p += blockDim.x) u += part_union[p];
    for (int o = 16; o; o >>= 1) u += __shfl_down_sync(0xffffffffu, u, o);
    if ((threadIdx.x & 31) == 0) su[threadIdx.x >> 5] = u;
    __syncthreads();
    if (threadIdx.x == 0) {
        u = 0;
        for (int w = 0; w < (blockDim.x + 31) / 32; w++) u += su[w];
        uint64_t rec = 0;
        for (int s = 0; s < S; s++) rec += starts[(size_t)n * S + s] - starts[s];
        out[0] = u; out[1] = rec;
    }
}

// ------------------------------------------------------------- host driver ---
struct MergeCfg { int gw, spl, chk, nslot; };
static MergeCfg pick_cfg(int S)
{
    if (S <= 16) return {16, 1, 8, 4};
    if (S <= 32) return {32, 1, 8, 4};
    if (S <= 64) return {32, 2, 8, 4};
    if (S <= 128) return {32, 4, 8, 2};
    return {32, 8, 8, 2};
}
template <int GW, int SPL, int CHK, int NSLOT>
static cudaError_t launch_merge(dk_ctx *ctx, const MergeParams &P, int warps, int grid, size_t smem)
{
    auto kfn = k_merge<GW, SPL, CHK, NSLOT>;
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    kfn<<<grid, warps * 32, smem, ctx->stream>>>(P);
    return cudaGetLastError();
}

void dk_matrix_release(dk_ctx *ctx, dk_matrix *m)
{
    if (!m) return;
    dk_dev_release(ctx, m->keys); dk_dev_release(ctx, m->cnts); dk_dev_release(ctx, m->chunk_rows);
    dk_dev_release(ctx, m->order); dk_dev_release(ctx, m->lbase);
    delete m;
}

int dk_join_impl(dk_ctx *ctx, int S, const uint64_t *const *keys_dev, const int32_t *const *cnts_dev,
                 const uint64_t *n, int min_rec, int min_abund, int emit_stale_tail, dk_matrix **out)
{
    if (!out) return dk_fail(ctx, DK_ERR_ARG, "join: out is NULL");
    *out = nullptr;
    if (S < 1 || S > DK_MAX_SAMPLES) return dk_fail(ctx, DK_ERR_ARG, "join: S=%d outside [1,%d]", S, DK_MAX_SAMPLES);
    uint64_t total = 0;
    for (int s = 0; s < S; s++) {
        total += n[s];
        if (n[s] && (((uintptr_t)keys_dev[s] & 15) || ((uintptr_t)cnts_dev[s] & 15)))
            return dk_fail(ctx, DK_ERR_ARG, "join: sample %d device pointers must be 16-byte aligned", s);
        if (n[s] >= (1ull << 40)) return dk_fail(ctx, DK_ERR_ARG, "join: sample %d too large", s);
    }
    cudaStream_t st = ctx->stream;
    DK_CUDA(ctx, cudaEventRecord(ctx->ev[0], st));

    // ---- tables on device
    size_t tab_bytes = (size_t)S * (8 + 8 + 8);
    unsigned char *tab = (unsigned char *)dk_dev_alloc(ctx, tab_bytes);
    if (!tab) return DK_ERR_NOMEM;
    std::vector<unsigned char> htab(tab_bytes);
    memcpy(htab.data(), keys_dev, (size_t)S * 8);
    memcpy(htab.data() + (size_t)S * 8, cnts_dev, (size_t)S * 8);
    memcpy(htab.data() + (size_t)S * 16, n, (size_t)S * 8);
    DK_CUDA(ctx, cudaMemcpyAsync(tab, htab.data(), tab_bytes, cudaMemcpyHostToDevice, st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));   // htab is a stack-lifetime staging buffer
    JoinTables T;
    T.keys = (const uint64_t *const *)tab;
    T.cnts = (const int32_t *const *)(tab + (size_t)S * 8);
    T.n = (const uint64_t *)(tab + (size_t)S * 16);

    // kernel choice: v4 (bucket join, S <= 64) by default, v3 (merge-path tree, any S) otherwise and whenever a dump
    // holds copies of a k-mer; DK_MERGE=1|3|4 forces the older kernels (profiling, tests)
    size_t budget = ctx->smem_optin ? ctx->smem_optin : 227 * 1024;
    int which = 5;
    bool auto_pick = true;           // v4 hands sparse unions to v3 after a probe of the data (below) unless DK_MERGE forces a kernel
    if (const char *e = getenv("DK_MERGE")) { which = atoi(e); auto_pick = false; }
    M3Cfg G3;
    M4Cfg G4;
    M5Cfg G5;
    int lv5 = 4;
    memset(&G3, 0, sizeof G3);
    memset(&G4, 0, sizeof G4);
    memset(&G5, 0, sizeof G5);
    if (which == 5 && !m5_make_cfg(S, budget, G5, lv5)) which = 4;      // v5: S <= 32
    if (which == 4 && !m4_make_cfg(S, budget, G4)) which = 3;
    const bool have3 = m3_make_cfg(S, budget, G3);
    if (which == 3 && !have3) which = 1;
    if (which == 2) which = 1;                                          // v2 (thread per key range) was removed in round 2

    // ---- K1: partitions of ~PART_TARGET records (v3 consumes ~4K records per round: longer partitions
    //      amortise the partial last round)
    uint64_t PART_TARGET = which >= 3 ? 65536 : 16384;
    if (which == 5) auto_pick = false;
    if (which >= 3) {       // big joins: fewer, longer partitions (less partition bookkeeping, fewer partly filled chunks) while
                            // every CTA still gets >= 32 of them (A/B at cfg 2: 65536 -> 262144 = 104.3 -> 102.7 ms per step)
        const uint64_t per_cta = total / ((uint64_t)ctx->sm_count * 3 * 32);
        if (per_cta > PART_TARGET) PART_TARGET = per_cta < 262144 ? per_cta : 262144;
    }
    if (const char *e = getenv("DK_PART_TARGET")) { long v = atol(e); if (v >= 4096) PART_TARGET = (uint64_t)v; }   // tuning runs
    uint64_t np_want = total / PART_TARGET + 1;
    int nb_log2 = 4;
    while ((1ull << nb_log2) < np_want * 4 && nb_log2 < 24) nb_log2++;
    uint32_t NB = 1u << nb_log2;
    uint64_t target = total / (NB / 4) + 1;      // ~NB/4 partitions
    if (target < PART_TARGET) target = PART_TARGET;
    uint32_t NPmax = NB;

    uint64_t *range = (uint64_t *)dk_dev_alloc(ctx, 64);
    uint64_t *below = (uint64_t *)dk_dev_alloc(ctx, sizeof(uint64_t) * (NB + 1));
    uint32_t *flag = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * (NB + 1));
    uint32_t *rank = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * (NB + 2));
    uint64_t *starts = (uint64_t *)dk_dev_alloc(ctx, sizeof(uint64_t) * (size_t)(NPmax + 1) * S);
    // scalars: [0]=npart [1]=work_counter [2]=chunk_counter ; then part tables
    uint32_t *scal = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * (16 + 4 * (size_t)NPmax));
    uint64_t *totals = (uint64_t *)dk_dev_alloc(ctx, 64);
    auto drop_tables = [&]() {
        dk_dev_release(ctx, tab); dk_dev_release(ctx, range); dk_dev_release(ctx, below); dk_dev_release(ctx, flag);
        dk_dev_release(ctx, rank); dk_dev_release(ctx, starts); dk_dev_release(ctx, scal); dk_dev_release(ctx, totals);
    };
    if (!range || !below || !flag || !rank || !starts || !scal || !totals) { drop_tables(); return DK_ERR_NOMEM; }
    uint32_t *npart_d = scal, *work_d = scal + 1, *chunkctr_d = scal + 2, *overflow_d = scal + 3;
    uint32_t *part_nchunks = scal + 16, *part_rows = part_nchunks + NPmax, *part_union = part_rows + NPmax,
             *part_choff = part_union + NPmax;   // NPmax entries (+1 lives in rank reuse below)

    k_key_range<<<1, 256, 0, st>>>(T, S, nb_log2, range); DK_LAUNCHED(ctx);
    k_bucket_below<<<(NB + 1 + 127) / 128, 128, 0, st>>>(T, S, NB, range, below); DK_LAUNCHED(ctx);
    k_part_flags<<<(NB + 127) / 128, 128, 0, st>>>(below, NB, target, flag); DK_LAUNCHED(ctx);
    DK_TRY(dk_scan_excl_u32(ctx, flag, rank, NB));
    k_part_starts<<<NB + 1, 32, 0, st>>>(T, S, NB, range, flag, rank, starts, npart_d); DK_LAUNCHED(ctx);

    uint32_t NP_host = 0;
    bool picked = false;
    if (which == 4 && auto_pick) {
        DK_CUDA(ctx, cudaMemcpyAsync(&NP_host, npart_d, 4, cudaMemcpyDeviceToHost, st));
        DK_CUDA(ctx, cudaStreamSynchronize(st));
    }

    // ---- K2 launch geometry
    MergeCfg cfg = pick_cfg(S);
    size_t stream_bytes = dk_round_up((size_t)S * (cfg.chk * cfg.nslot * 12 + cfg.nslot * 8), 16);
    int gpw = 32 / cfg.gw;
    size_t warp_bytes = stream_bytes * gpw;
    int warps = 4;
    while (warps > 1 && warp_bytes * warps > budget / 2) warps >>= 1;
    if (warp_bytes * warps > budget) return dk_fail(ctx, DK_ERR_UNSUPPORTED, "join: S=%d needs %zu B smem per warp", S, warp_bytes);
    size_t smem = warp_bytes * warps;
    int ctas_per_sm = (int)((budget + 1024) / (smem + 1024));
    if (ctas_per_sm < 1) ctas_per_sm = 1;
    if (ctas_per_sm * warps > 48) ctas_per_sm = 48 / warps;
    int grid = ctx->sm_count * ctas_per_sm;

    // ---- output capacity: previous need, else a guess; retried with the exact need on overflow
    uint64_t guess = ctx->chunk_hint ? ctx->chunk_hint + ctx->chunk_hint / 16 + 64
                                  : (total / (uint64_t)(min_rec > 1 ? min_rec : 1) / 4) / DK_CH + NPmax / 4 + 64;
    if (!ctx->chunk_hint) {
        // a first guess must not eat the device: at most half of what is free now (cached blocks of this context count as
        // free). Too small a guess costs one more pass with the exact need (the kernel keeps counting past the capacity).
        size_t mfree = 0, mtot = 0;
        if (cudaMemGetInfo(&mfree, &mtot) == cudaSuccess) {
            const size_t per_chunk = (size_t)DK_CH * (8 + 4 * (size_t)S) + 12;
            const uint64_t fit = (uint64_t)((mfree + ctx->cached_bytes) / 2 / per_chunk);
            if (guess > fit && fit > 1024) guess = fit;
        }
    }
    dk_matrix *m = new dk_matrix();
    m->S = S;
    uint32_t *chunk_part = nullptr, *chunk_seq = nullptr;
    uint64_t h_tot[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    auto drop_chunks = [&]() {
        dk_dev_release(ctx, chunk_part); dk_dev_release(ctx, chunk_seq);
        chunk_part = chunk_seq = nullptr;
    };
    // capacity retries and kernel fallbacks are counted apart: a fallback (v5/v4 -> v3 -> v1) is not a failed attempt
    for (int attempt = 0, fallbacks = 0; attempt < 3 && fallbacks < 3;) {
        uint64_t cap = guess + 1;   // +1: room for the stale-tail row
        if (cap >= 0xFFFFFFF0ull) { dk_matrix_release(ctx, m); drop_tables(); return dk_fail(ctx, DK_ERR_NOMEM, "join: too many chunks"); }
        m->cap_chunks = cap;
        m->keys = (uint64_t *)dk_dev_alloc(ctx, sizeof(uint64_t) * cap * DK_CH);
        m->cnts = (int32_t *)dk_dev_alloc(ctx, sizeof(int32_t) * cap * DK_CH * S);
        m->chunk_rows = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * cap);
        chunk_part = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * cap);
        chunk_seq = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * cap);
        if (!m->keys || !m->cnts || !m->chunk_rows || !chunk_part || !chunk_seq) { dk_matrix_release(ctx, m); drop_chunks(); drop_tables(); return DK_ERR_NOMEM; }
        DK_CUDA(ctx, cudaMemsetAsync(scal + 1, 0, sizeof(uint32_t) * (15 + 4 * (size_t)NPmax), st));

        MergeParams P;
        P.T = T; P.S = S; P.min_rec = min_rec; P.min_abund = min_abund;
        P.starts = starts; P.npart = npart_d; P.work_counter = work_d;
        P.out_keys = m->keys; P.out_cnts = m->cnts;
        P.chunk_rows = m->chunk_rows; P.chunk_part = chunk_part; P.chunk_seq = chunk_seq;
        P.chunk_counter = chunkctr_d; P.max_chunks = (uint32_t)guess;
        P.part_nchunks = part_nchunks; P.part_rows = part_rows; P.part_union = part_union;

        DK_CUDA(ctx, cudaEventRecord(ctx->ev[2], st));
        cudaError_t e = cudaSuccess;
        auto launch_k = [&](int wk, int max_ctas) -> cudaError_t {
            cudaError_t e;
            auto clampg = [&](int g) { return (max_ctas > 0 && g > max_ctas) ? max_ctas : g; };
            if (wk == 5) {
                typedef void (*m5_fn)(MergeParams, M5Cfg, uint32_t *);
                const m5_fn k5 = lv5 == 4 ? (m5_fn)k_merge5<4> : (m5_fn)k_merge5<5>;
                e = cudaFuncSetAttribute(k5, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G5.smem_bytes);
                if (e == cudaSuccess) {
                    int per_sm = (int)((budget + 1024) / ((size_t)G5.smem_bytes + 1024));
                    if (per_sm < 1) per_sm = 1;
                    if (per_sm > M5_CTAS) per_sm = M5_CTAS;
                    if (const char *pe = getenv("DK_M5_CTAS")) { int v = atoi(pe); if (v >= 1 && v < per_sm) per_sm = v; }   // tuning runs
                    k5<<<clampg(ctx->sm_count * per_sm), M5_NT, G5.smem_bytes, ctx->stream>>>(P, G5, overflow_d);
                    e = cudaGetLastError();
                }
            } else if (wk == 4) {
                typedef void (*m4_fn)(MergeParams, M4Cfg, uint32_t *);
                static const m4_fn m4_tab[7] = {k_merge4<0>, k_merge4<1>, k_merge4<2>, k_merge4<3>, k_merge4<4>, k_merge4<5>, k_merge4<6>};
                const m4_fn k4 = m4_tab[G4.levels];
                e = cudaFuncSetAttribute(k4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G4.smem_bytes);
                if (e == cudaSuccess) {
                    int per_sm = (int)((budget + 1024) / ((size_t)G4.smem_bytes + 1024));
                    if (per_sm < 1) per_sm = 1;
                    if (per_sm > M4_CTAS) per_sm = M4_CTAS;
                    k4<<<clampg(ctx->sm_count * per_sm), M4_NT, G4.smem_bytes, ctx->stream>>>(P, G4, overflow_d);
                    e = cudaGetLastError();
                }
            } else if (wk == 3) {
                typedef void (*m3_fn)(MergeParams, M3Cfg, uint32_t *);
                static const m3_fn m3_tab[9] = {k_merge3<0>, k_merge3<1>, k_merge3<2>, k_merge3<3>, k_merge3<4>,
                                                k_merge3<5>, k_merge3<6>, k_merge3<7>, k_merge3<8>};
                const m3_fn k3 = m3_tab[G3.levels];
                e = cudaFuncSetAttribute(k3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G3.smem_bytes);
                if (e == cudaSuccess) {
                    int per_sm = (int)((budget + 1024) / ((size_t)G3.smem_bytes + 1024));
                    if (per_sm < 1) per_sm = 1;
                    if (per_sm > 3) per_sm = 3;
                    k3<<<clampg(ctx->sm_count * per_sm), M3_NT, G3.smem_bytes, ctx->stream>>>(P, G3, overflow_d);
                    e = cudaGetLastError();
                }
            } else if (cfg.gw == 16) e = launch_merge<16, 1, 8, 4>(ctx, P, warps, grid, smem);
            else if (cfg.spl == 1) e = launch_merge<32, 1, 8, 4>(ctx, P, warps, grid, smem);
            else if (cfg.spl == 2) e = launch_merge<32, 2, 8, 4>(ctx, P, warps, grid, smem);
            else if (cfg.spl == 4) e = launch_merge<32, 4, 8, 2>(ctx, P, warps, grid, smem);
            else e = launch_merge<32, 8, 8, 2>(ctx, P, warps, grid, smem);
            ctx->launches++;
            return e;
        };
        // Probe: v4's cost is per ROW of the union (bucket tables, byte matrix), v3's per record. On data where a k-mer
        // occurs in few dumps (config 5: 2.1 records per k-mer) v3 is the faster kernel (34 vs 74 ms), on cohort-like data
        // v4 is (cfg 2: 44 vs 71 ms, cfg 3: 20 vs 35 ms). So v4 first merges a few partitions (~1-3% of the input), the
        // records-per-k-mer ratio of those decides, and the chosen kernel takes the remaining partitions: both kernels
        // append to the same chunked matrix through the same partition tables.
        if (which == 4 && auto_pick && !picked && have3 && NP_host >= 64) {
            uint32_t probe_n = NP_host / 32;
            const uint32_t cap_n = (uint32_t)ctx->sm_count * M4_CTAS;
            if (probe_n > cap_n) probe_n = cap_n;
            uint32_t *probe_d = (uint32_t *)(totals + 6);
            uint64_t *pstat_d = totals + 4, h_ps[2] = {0, 0};   // totals[4..5] are rewritten by k_join_totals afterwards
            DK_CUDA(ctx, cudaMemcpyAsync(probe_d, &probe_n, 4, cudaMemcpyHostToDevice, st));
            P.npart = probe_d;
            e = launch_k(4, (int)probe_n);
            if (e == cudaSuccess) {
                k_probe_stats<<<1, 256, 0, st>>>(part_union, starts, S, probe_n, pstat_d); DK_LAUNCHED(ctx);
                DK_CUDA(ctx, cudaMemcpyAsync(h_ps, pstat_d, 16, cudaMemcpyDeviceToHost, st));
                DK_CUDA(ctx, cudaStreamSynchronize(st));
                if (h_ps[0] && (double)h_ps[1] < 4.0 * (double)h_ps[0]) which = 3;      // < 4 records per k-mer
                picked = true;
                DK_CUDA(ctx, cudaMemcpyAsync(work_d, &probe_n, 4, cudaMemcpyHostToDevice, st));
                DK_CUDA(ctx, cudaStreamSynchronize(st));                                // probe_n is a stack variable
                P.npart = npart_d;
            }
        }
        if (e == cudaSuccess) e = launch_k(which, 0);
        if (e != cudaSuccess) { dk_matrix_release(ctx, m); drop_chunks(); drop_tables(); return dk_fail(ctx, DK_ERR_CUDA, "merge launch: %s", cudaGetErrorString(e)); }
        DK_CUDA(ctx, cudaEventRecord(ctx->ev[3], st));
        k_join_totals<<<1, 256, 0, st>>>(part_rows, part_union, npart_d, chunkctr_d, overflow_d, totals); DK_LAUNCHED(ctx);
        DK_CUDA(ctx, cudaMemcpyAsync(h_tot, totals, sizeof h_tot, cudaMemcpyDeviceToHost, st));
        DK_CUDA(ctx, cudaStreamSynchronize(st));
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]);
        ctx->stage_ms[4] = ms;
#ifdef DK_M3_PROF
        if (which == 3) {
            uint32_t pr[12];
            cudaMemcpy(pr, scal + 4, sizeof pr, cudaMemcpyDeviceToHost);
            double tot = 0;
            for (int i = 1; i < 11; i++) tot += pr[i];
            fprintf(stderr, "[m3 prof] kernel %.3f ms; phase shares: stage+wait %.3f pivot %.3f pack %.3f merge %.3f heads %.3f verdict %.3f | zero %.3f scatter %.3f copyout %.3f advance %.3f\n",
                    ms, pr[1] / tot, pr[2] / tot, pr[3] / tot, pr[4] / tot, pr[5] / tot, pr[6] / tot, pr[8] / tot, pr[9] / tot, pr[10] / tot, pr[7] / tot);
        }
#endif
#ifdef DK_M5_PROF
        if (which == 5) {
            uint32_t pr[12];
            cudaMemcpy(pr, scal + 4, sizeof pr, cudaMemcpyDeviceToHost);
            double tot = 0;
            for (int i = 1; i < 10; i++) tot += pr[i];
            fprintf(stderr, "[m5 prof] kernel %.3f ms; phase shares: tma wait %.3f pivot(warp 0) %.3f B1 wait %.3f copy-out %.3f table %.3f classify+light %.3f "
                    "simple rows %.3f complex rows %.3f scan+alloc %.3f\n",
                    ms, pr[1] / tot, pr[9] / tot, pr[2] / tot, pr[3] / tot, pr[4] / tot, pr[5] / tot, pr[6] / tot, pr[7] / tot, pr[8] / tot);
        }
#endif
#ifdef DK_M4_PROF
        if (which == 4) {
            uint32_t pr[12];
            cudaMemcpy(pr, scal + 4, sizeof pr, cudaMemcpyDeviceToHost);
            double tot = 0;
            for (int i = 1; i < 12; i++) tot += pr[i];
            fprintf(stderr, "[m4 prof] kernel %.3f ms; phase shares: stage+wait %.3f pivot %.3f clear+prefix %.3f buckets+election %.3f "
                    "bucket scan+ranks %.3f recurrence bytes %.3f verdicts %.3f | zero %.3f scatter %.3f copyout %.3f advance %.3f\n",
                    ms, pr[1] / tot, pr[2] / tot, pr[3] / tot, pr[4] / tot, pr[5] / tot, pr[6] / tot, pr[7] / tot, pr[8] / tot,
                    pr[9] / tot, pr[10] / tot, pr[11] / tot);
        }
#endif
        if (h_tot[4] && which >= 3) {
            // v5: one k-mer fills a whole window of a dump; v4: a dump holds copies of a k-mer -> redo with v3 (copies become
            // separate rows, as in the reference); v3: a dump holds one k-mer more than 4 times (its copy-index field
            // overflowed) -> redo with the v1 kernel. The chunk need seen so far still sizes the next try.
            which = (which >= 4 && have3) ? 3 : 1;
            dk_dev_release(ctx, m->keys); dk_dev_release(ctx, m->cnts); dk_dev_release(ctx, m->chunk_rows);
            drop_chunks();
            m->keys = nullptr; m->cnts = nullptr; m->chunk_rows = nullptr;
            if (h_tot[2] > guess) guess = h_tot[2];
            fallbacks++;
            continue;
        }
        if (h_tot[2] <= guess) break;
        // overflow: the counter kept counting, so h_tot[2] is the exact need
        dk_dev_release(ctx, m->keys); dk_dev_release(ctx, m->cnts); dk_dev_release(ctx, m->chunk_rows);
        drop_chunks();
        m->keys = nullptr; m->cnts = nullptr; m->chunk_rows = nullptr;
        guess = h_tot[2];
        attempt++;
        if (attempt == 3) { dk_matrix_release(ctx, m); drop_tables(); return dk_fail(ctx, DK_ERR_CUDA, "join: output overflow twice"); }
    }
    if (!m->keys) { dk_matrix_release(ctx, m); drop_tables(); return dk_fail(ctx, DK_ERR_CUDA, "join: no merge kernel accepted the input"); }
    ctx->chunk_hint = h_tot[2];
    m->R = h_tot[0]; m->U = h_tot[1]; m->n_chunks = h_tot[2];
    uint32_t NP = (uint32_t)h_tot[3];
    uint32_t nch = (uint32_t)m->n_chunks;

    // ---- order chunks: reference order == (partition, seq)
    m->order = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * (nch + 2));
    m->lbase = (uint64_t *)dk_dev_alloc(ctx, sizeof(uint64_t) * (nch + 3));
    uint32_t *choff = (uint32_t *)dk_dev_alloc(ctx, sizeof(uint32_t) * ((size_t)NP + 2));
    if (!m->order || !m->lbase || !choff) { dk_matrix_release(ctx, m); dk_dev_release(ctx, choff); drop_chunks(); drop_tables(); return DK_ERR_NOMEM; }
    (void)part_choff;
    DK_TRY(dk_scan_excl_u32(ctx, part_nchunks, choff, NP));
    if (nch) {
        k_chunk_order<<<(nch + 255) / 256, 256, 0, st>>>(chunk_part, chunk_seq, choff, nch, m->order); DK_LAUNCHED(ctx);
        k_chunk_rows_in_order<<<(nch + 255) / 256, 256, 0, st>>>(m->order, m->chunk_rows, nch, m->lbase); DK_LAUNCHED(ctx);
    }
    DK_TRY(dk_scan_excl_u64(ctx, m->lbase, m->lbase, nch));
    if (emit_stale_tail && min_rec <= 0 && m->U > 0) {
        k_stale_tail<<<1, 64, 0, st>>>(m->keys, m->cnts, m->chunk_rows, m->order, m->lbase, nch, nch, S, range); DK_LAUNCHED(ctx);
        m->n_chunks++; m->R++;
    }
    DK_CUDA(ctx, cudaEventRecord(ctx->ev[1], st));
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
    ctx->stage_ms[0] = ms;
    drop_tables();
    drop_chunks(); dk_dev_release(ctx, choff);
    *out = m;
    return DK_OK;
}
