// dk_merge2.cuh — K2 v2: thread-per-key-range k-way merge for S <= 16 (included by dk_join.cu).
//
// r01 ncu of v1 (lane-per-sample, profiles/r01_ncu_merge_v1.csv) showed the merge issue-bound at
// ~77 warp-instructions per output row: every row costs a warp-wide min/vote, and partial-mask
// collectives of the two 16-lane streams of a warp execute serially. v2 removes all warp-wide
// work from the per-row path: a CTA walks its partition in ROUNDS; per round every sample
// contributes a window of <= CAPS records staged in shared memory by 1-D TMA bulk copies; the
// round's key range ends at the smallest first-unloaded key (so the staged prefix is complete);
// the staged prefix is split into NT key ranges by quantiles of the longest slice; each THREAD
// merges its own range sequentially with its 16 heads in registers (fully unrolled, no dynamic
// indexing). Pass 1 records per row the presence mask and the recurrence verdict; after a block
// scan of the per-thread pass counts, pass 2 replays the masks and writes the surviving rows in
// key order (joinCounts.c:72-115 semantics, including duplicate keys inside one dump).
#pragma once

#define M2_S 16          // samples handled in registers
#define M2_NT 128        // threads per CTA
#define M2_CAPS 256      // window records per sample per round
#define M2_SLICE (M2_CAPS + 8)

struct M2Smem {
    uint64_t K[M2_S][M2_SLICE];
    int32_t C[M2_S][M2_SLICE];
    uint32_t rowmask[M2_S * M2_CAPS];
    uint16_t st[M2_S][M2_NT + 1];
    unsigned long long cur[M2_S], endg[M2_S], cand[M2_S];
    uint32_t lo[M2_S], n[M2_S], w[M2_S];
    uint32_t scan[M2_NT / 32];
    uint32_t sc_total;
    unsigned long long bar;
    unsigned long long pivot;
    uint32_t part, pivot_s, total_n, use_upper;
    uint32_t chunk_cur, rowpos, seq, newbase, nnew, have_cur;
    uint32_t prow, puni;
};

__device__ __forceinline__ uint32_t m2_block_excl_scan(uint32_t v, M2Smem &sm, uint32_t &total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    __syncthreads();                       // protect sm.scan from the previous use
    if (lane == 31) sm.scan[w] = x;
    __syncthreads();
    uint32_t off = 0, tot = 0;
#pragma unroll
    for (int i = 0; i < M2_NT / 32; i++) { uint32_t s = sm.scan[i]; if (i < w) off += s; tot += s; }
    total = tot;
    return off + x - v;
}

__global__ void __launch_bounds__(M2_NT) k_merge2(MergeParams P)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    M2Smem &sm = *reinterpret_cast<M2Smem *>(smem_raw);
    const int S = P.S, tid = threadIdx.x;
    const uint32_t npart = *P.npart;
    const uint32_t bar = smem_u32(&sm.bar);
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    uint32_t phase = 0;

    for (;;) {
        if (tid == 0) sm.part = atomicAdd(P.work_counter, 1u);
        __syncthreads();
        const uint32_t part = sm.part;
        if (part >= npart) break;
        if (tid < M2_S) {
            sm.cur[tid] = tid < S ? P.starts[(size_t)part * S + tid] : 0;
            sm.endg[tid] = tid < S ? P.starts[(size_t)(part + 1) * S + tid] : 0;
        }
        if (tid == 0) { sm.chunk_cur = CHUNK_NONE; sm.rowpos = 0; sm.seq = 0; sm.prow = 0; sm.puni = 0; }
        __syncthreads();

        for (;;) {   // ---- rounds
            // R0: windows + loads. Thread 0 posts the expected byte count, threads < S issue the copies.
            if (tid < M2_S) {
                unsigned long long c = sm.cur[tid], e = sm.endg[tid];
                uint32_t w = (uint32_t)((e - c) < (unsigned long long)M2_CAPS ? (e - c) : M2_CAPS);
                sm.w[tid] = w;
                sm.lo[tid] = (uint32_t)(c & 3ull);
                sm.cand[tid] = (c + w < e) ? P.T.keys[tid][c + w] : DK_SENTINEL;   // first unloaded key
            }
            __syncthreads();
            bool any = false;
#pragma unroll
            for (int s = 0; s < M2_S; s++) any = any || (sm.w[s] != 0);
            if (!any) break;
            if (tid == 0) {
                uint32_t bytes = 0;
                for (int s = 0; s < S; s++) {
                    unsigned long long c = sm.cur[s];
                    unsigned long long a_lo = (c + 3) & ~3ull, a_hi = (c + sm.w[s]) & ~3ull;
                    if (a_hi > a_lo) bytes += (uint32_t)(a_hi - a_lo) * 12u;
                }
                mbar_expect_tx(bar, bytes);
            }
            if (tid < S && sm.w[tid]) {
                const int s = tid;
                const unsigned long long c = sm.cur[s], base_g = c & ~3ull, e = c + sm.w[s];
                unsigned long long a_lo = (c + 3) & ~3ull, a_hi = e & ~3ull;
                const uint64_t *ks = P.T.keys[s];
                const int32_t *cs = P.T.cnts[s];
                if (a_hi > a_lo) {
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // last round's generic reads precede these async writes
                    tma_bulk_g2s(smem_u32(&sm.K[s][a_lo - base_g]), ks + a_lo, (uint32_t)(a_hi - a_lo) * 8u, bar);
                    tma_bulk_g2s(smem_u32(&sm.C[s][a_lo - base_g]), cs + a_lo, (uint32_t)(a_hi - a_lo) * 4u, bar);
                } else { a_lo = e; a_hi = e; }
                for (unsigned long long g = c; g < a_lo && g < e; g++) { sm.K[s][g - base_g] = ks[g]; sm.C[s][g - base_g] = cs[g]; }
                for (unsigned long long g = (a_hi > c ? a_hi : c); g < e; g++) { sm.K[s][g - base_g] = ks[g]; sm.C[s][g - base_g] = cs[g]; }
            }
            // R1: pivot = smallest first-unloaded key
            if (tid == 0) {
                unsigned long long pv = DK_SENTINEL;
                for (int s = 0; s < S; s++) pv = sm.cand[s] < pv ? sm.cand[s] : pv;
                sm.pivot = pv;
                sm.use_upper = 0;
            }
            mbar_wait(bar, phase);
            phase ^= 1u;
            __syncthreads();
            // R2: per-sample consumable prefix n[s] = #keys < pivot in the window
            for (int attempt = 0; attempt < 2; attempt++) {
                if (tid < M2_S) {
                    uint32_t n = 0;
                    if (tid < S && sm.w[tid]) {
                        const unsigned long long pv = sm.pivot;
                        const uint32_t lo0 = sm.lo[tid], w = sm.w[tid];
                        if (pv == DK_SENTINEL) n = w;
                        else {
                            uint32_t lo = 0, hi = w;
                            const bool upper = sm.use_upper != 0;
                            while (lo < hi) {
                                uint32_t mid = (lo + hi) >> 1;
                                unsigned long long k = sm.K[tid][lo0 + mid];
                                if (upper ? (k <= pv) : (k < pv)) lo = mid + 1; else hi = mid;
                            }
                            n = lo;
                        }
                    }
                    sm.n[tid] = n;
                }
                __syncthreads();
                if (tid == 0) {
                    uint32_t tot = 0, best = 0, bs = 0;
                    for (int s = 0; s < M2_S; s++) { tot += sm.n[s]; if (sm.n[s] > best) { best = sm.n[s]; bs = s; } }
                    sm.total_n = tot; sm.pivot_s = bs;
                    if (tot == 0) sm.use_upper = 1;     // > CAPS duplicates of one key inside a dump: take keys == pivot too
                }
                __syncthreads();
                if (sm.total_n) break;
            }
            // R3: splitters = quantiles of the longest slice; per thread S lower bounds
            {
                const uint32_t ps = sm.pivot_s, pn = sm.n[ps];
                const unsigned long long sp = (tid == 0) ? 0ull : sm.K[ps][sm.lo[ps] + (uint32_t)(((unsigned long long)tid * pn) / M2_NT)];
#pragma unroll
                for (int s = 0; s < M2_S; s++) {
                    const uint32_t lo0 = sm.lo[s], n = sm.n[s];
                    uint32_t lo = 0, hi = n;
                    if (tid == 0) hi = 0;
                    while (lo < hi) {
                        uint32_t mid = (lo + hi) >> 1;
                        if (sm.K[s][lo0 + mid] < sp) lo = mid + 1; else hi = mid;
                    }
                    sm.st[s][tid] = (uint16_t)(lo0 + lo);
                    if (tid == 0) sm.st[s][M2_NT] = (uint16_t)(lo0 + n);
                }
            }
            __syncthreads();
            // R4/R5: pass 1 — sequential merge of this thread's key range, heads in registers
            uint32_t ce[M2_S];
            unsigned long long hk[M2_S];
            int32_t hc[M2_S];
            uint32_t nrec = 0;
#pragma unroll
            for (int s = 0; s < M2_S; s++) {
                uint32_t a = sm.st[s][tid], b = sm.st[s][tid + 1];
                ce[s] = a | (b << 16);
                nrec += b - a;
                hk[s] = (a < b) ? sm.K[s][a] : DK_SENTINEL;
                hc[s] = (a < b) ? sm.C[s][a] : 0;
            }
            uint32_t dummy;
            const uint32_t rowbase = m2_block_excl_scan(nrec, sm, dummy);
            uint32_t rows = 0, npass = 0;
            const int min_ab = P.min_abund, min_rec = P.min_rec;
            for (;;) {
                unsigned long long m = hk[0];
#pragma unroll
                for (int s = 1; s < M2_S; s++) m = hk[s] < m ? hk[s] : m;
                if (m == DK_SENTINEL) break;
                uint32_t mask = 0;
                int rec = 0;
#pragma unroll
                for (int s = 0; s < M2_S; s++) {
                    const bool p = (hk[s] == m);
                    mask |= (p ? 1u : 0u) << s;
                    rec += (p && hc[s] >= min_ab) ? 1 : 0;                 // joinCounts.c:81-83
                    if (p) {
                        const uint32_t c = ce[s] + 1;
                        ce[s] = c;
                        const uint32_t i = c & 0xffffu;
                        const bool more = i != (c >> 16);
                        hk[s] = more ? sm.K[s][i] : DK_SENTINEL;
                        hc[s] = more ? sm.C[s][i] : 0;
                    }
                }
                const uint32_t pass = rec >= min_rec ? 1u : 0u;             // joinCounts.c:104
                sm.rowmask[rowbase + rows] = mask | (pass << 16);
                rows++;
                npass += pass;
            }
            // R6: block scans -> in-order output offsets
            uint32_t tile_pass, tile_rows;
            const uint32_t outoff = m2_block_excl_scan(npass, sm, tile_pass);
            (void)m2_block_excl_scan(rows, sm, tile_rows);
            // R7: chunk allocation for tile_pass rows appended to the partition's chunk sequence
            if (tid == 0) {
                const uint32_t have = (sm.chunk_cur != CHUNK_NONE) ? 1u : 0u;
                const uint32_t slots = have ? (DK_CH - sm.rowpos) : 0u;
                uint32_t nnew = tile_pass > slots ? (tile_pass - slots + DK_CH - 1) / DK_CH : 0u;
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
            // R8: pass 2 — replay the masks, write the surviving rows in key order
            {
                const uint32_t have = sm.have_cur, nb = sm.newbase, rp0 = sm.rowpos, ccur = sm.chunk_cur;
                uint32_t cu[M2_S];
#pragma unroll
                for (int s = 0; s < M2_S; s++) cu[s] = sm.st[s][tid];
                uint32_t q = outoff;
                for (uint32_t j = 0; j < rows; j++) {
                    const uint32_t mk = sm.rowmask[rowbase + j];
                    if (mk & 0x10000u) {
                        uint32_t v = (have ? rp0 : 0u) + q;
                        uint32_t k = v / DK_CH, off = v % DK_CH;
                        uint32_t chunk = (have && k == 0) ? ccur : nb + (k - have);
                        q++;
                        if (chunk < P.max_chunks) {
                            const size_t phys = (size_t)chunk * DK_CH + off;
                            unsigned long long key = 0;
                            int32_t c[M2_S];
#pragma unroll
                            for (int s = 0; s < M2_S; s++) {
                                const bool p = (mk >> s) & 1u;
                                c[s] = p ? sm.C[s][cu[s]] : 0;
                                if (p) key = sm.K[s][cu[s]];
                            }
                            if (S == M2_S) {
                                int4 *dst = reinterpret_cast<int4 *>(P.out_cnts + phys * M2_S);
                                dst[0] = make_int4(c[0], c[1], c[2], c[3]);
                                dst[1] = make_int4(c[4], c[5], c[6], c[7]);
                                dst[2] = make_int4(c[8], c[9], c[10], c[11]);
                                dst[3] = make_int4(c[12], c[13], c[14], c[15]);
                            } else {
#pragma unroll
                                for (int s = 0; s < M2_S; s++) if (s < S) P.out_cnts[phys * S + s] = c[s];
                            }
                            P.out_keys[phys] = key;
                        }
                    }
#pragma unroll
                    for (int s = 0; s < M2_S; s++) cu[s] += (mk >> s) & 1u;
                }
            }
            __syncthreads();
            // R9: advance the partition state
            if (tid < M2_S) sm.cur[tid] += sm.n[tid];
            if (tid == 0) {
                const uint32_t v = (sm.have_cur ? sm.rowpos : 0u) + tile_pass;
                if (sm.nnew) {
                    sm.chunk_cur = sm.newbase + sm.nnew - 1;
                    sm.seq += sm.nnew;
                    sm.rowpos = v - (sm.have_cur + sm.nnew - 1) * DK_CH;
                } else sm.rowpos = v;
                if (sm.chunk_cur != CHUNK_NONE && sm.rowpos == DK_CH) { sm.chunk_cur = CHUNK_NONE; sm.rowpos = 0; }
                sm.prow += tile_pass; sm.puni += tile_rows;
            }
            __syncthreads();
        }
        if (tid == 0) {
            if (sm.chunk_cur != CHUNK_NONE && sm.chunk_cur < P.max_chunks) P.chunk_rows[sm.chunk_cur] = sm.rowpos;
            P.part_nchunks[part] = sm.seq;
            P.part_rows[part] = sm.prow;
            P.part_union[part] = sm.puni;
        }
        __syncthreads();
    }
}
