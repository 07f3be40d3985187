// dk_tags.cu — K10, contig assembly of the surviving k-mers (share/mergeTags/mergeTags.c:68-212),
// stranded mode. For every overlap length i = k-1 .. min_k the reference indexes the i-base left and right ends of
// all current contigs in two hash maps, an end shared by two contigs being void (add_assembly_kmer :56-66), and
// merges X+Y whenever X's right end equals Y's left end, pair by pair in hash-bucket order, pass after pass, until
// nothing merges (:74-205). The link graph of one level is static — merging never creates or frees an end — and every
// contig has at most one successor and one predecessor, so the fixed point of those passes is: every PATH of links
// becomes one contig. That is what is computed here, in parallel:
//   ends    2-bit packed i-mers of both ends of every contig longer than i;
//   maps    radix sort of (end, contig) pairs; an end is usable when it occurs once on its side;
//   links   every usable right end is looked up among the usable left ends (binary search) -> succ / pred;
//   ranks   pointer jumping along pred: offset of every contig inside its path's sequence, path aggregates
//           (number of merged k-mers; representative = the member with the lowest p-value, the rightmost one on
//           ties, which is what the pairwise rule `left.p < right.p ? left : right` (:147-153) folds to);
//   build   new sequences = head + every other member without its first i bases (:183-184).
// A set of contigs linked in a ring is merged by the reference all but one link, WHICH link depending on bucket
// order; here ring members are left unmerged at that level and counted (dk_contigs_rings).
// Checked against the unmodified reference binary (oracle/_ref/mergeTags) on the config-1 tables and on random
// chains with ties, branches and repeats (tests/test_tags.py).
#include "dk_common.cuh"

struct dk_contigs {
    uint64_t n = 0, nchars = 0, rings = 0;
    // host copies (small: one entry per contig)
    std::vector<uint32_t> rep, nb;
    std::vector<uint64_t> off;
    std::vector<char> chars;
};

#define TG_NONE 0xFFFFFFFFu
#define TG_T 256

__device__ __forceinline__ int tg_code(char c)
{
    switch (c) { case 'A': return 0; case 'C': return 1; case 'G': return 2; case 'T': return 3; default: return -1; }
}

__global__ void k_tags_check(const char *chars, uint64_t n, uint32_t *bad)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && tg_code(chars[i]) < 0) atomicOr(bad, 1u);
}

// packed ends of every contig longer than i (others get the all-ones key, which is never usable)
__global__ void k_tags_ends(const char *chars, const uint64_t *off, uint32_t n, int i, uint64_t *lkey, uint64_t *rkey, uint32_t *id0,
                            uint32_t *id1)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint64_t b = off[j], len = off[j + 1] - b;
    uint64_t l = ~0ull, r = ~0ull;
    if (len > (uint64_t)i) {                                   // mergeTags.c:92 (strlen(seq) <= i: not indexed)
        l = 0; r = 0;
        for (int t = 0; t < i; t++) {
            l = (l << 2) | (uint64_t)tg_code(chars[b + t]);
            r = (r << 2) | (uint64_t)tg_code(chars[b + len - i + t]);
        }
    }
    lkey[j] = l; rkey[j] = r; id0[j] = j; id1[j] = j;
}

__device__ __forceinline__ bool tg_usable(const uint64_t *k, uint32_t n, uint32_t t)
{
    const uint64_t v = k[t];
    return v != ~0ull && (t == 0 || k[t - 1] != v) && (t + 1 >= n || k[t + 1] != v);     // occurs once on its side (:56-66)
}

__global__ void k_tags_links(const uint64_t *rk, const uint32_t *rj, const uint64_t *lk, const uint32_t *lj, uint32_t n, uint32_t *succ,
                             uint32_t *pred)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n || !tg_usable(rk, n, t)) return;
    const uint64_t v = rk[t];
    uint32_t lo = 0, hi = n;
    while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (lk[mid] < v) lo = mid + 1; else hi = mid; }
    if (lo >= n || lk[lo] != v || !tg_usable(lk, n, lo)) return;
    const uint32_t x = rj[t], y = lj[lo];
    if (x == y) return;                                        // :136 (a contig never merges with itself)
    succ[x] = y; pred[y] = x;
}

struct TgAgg { uint32_t nb, rep; double p; };

__global__ void k_tags_rank_init(const uint64_t *off, const uint32_t *pred, const uint32_t *nb, const uint32_t *rep, const double *p,
                                 uint32_t n, int i, uint32_t *anc, uint64_t *val, uint32_t *root, TgAgg *agg)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t q = pred[j];
    anc[j] = q; root[j] = j;
    uint64_t v = 0;
    if (q != TG_NONE) {                                        // piece of the predecessor: whole sequence for a head
        const uint64_t lq = off[q + 1] - off[q];
        v = pred[q] == TG_NONE ? lq : lq - (uint64_t)i;
    }
    val[j] = v;
    agg[j].nb = nb[j]; agg[j].rep = rep[j]; agg[j].p = p[j];
}
__global__ void k_tags_rank_step(uint32_t n, const uint32_t *anc, const uint64_t *val, const uint32_t *root, const TgAgg *agg,
                                 uint32_t *anc2, uint64_t *val2, uint32_t *root2, TgAgg *agg2, uint32_t *active)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t a = anc[j];
    uint32_t na = a, nr = root[j];
    uint64_t nv = val[j];
    TgAgg g = agg[j];
    if (a != TG_NONE) {
        const TgAgg l = agg[a];                                // the ancestors' side is on the left
        g.nb += l.nb;
        if (l.p < g.p) { g.p = l.p; g.rep = l.rep; }           // :147-153 left.p < right.p ? left : right
        nv += val[a];
        nr = root[a];
        na = anc[a];
        if (na != TG_NONE) *active = 1u;
    }
    anc2[j] = na; val2[j] = nv; root2[j] = nr; agg2[j] = g;
}
// after the rounds: members of a ring still have an ancestor; they stay unmerged at this level
__global__ void k_tags_classify(uint32_t n, const uint32_t *anc, uint32_t *pred, uint32_t *succ, uint32_t *start, uint32_t *ring)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const bool r = anc[j] != TG_NONE;
    ring[j] = r ? 1u : 0u;
    start[j] = (r || pred[j] == TG_NONE) ? 1u : 0u;            // heads, isolated contigs and ring members open a new contig
}
__global__ void k_tags_newlen(uint32_t n, const uint64_t *off, const uint32_t *pred, const uint32_t *succ, const uint32_t *ring,
                              const uint32_t *root, const uint64_t *val, const TgAgg *agg, const uint32_t *newid_of_start,
                              const uint32_t *nb, const uint32_t *rep, const double *p, int i, uint64_t *newlen, uint32_t *onb,
                              uint32_t *orep, double *op)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint64_t len = off[j + 1] - off[j];
    if (ring[j] || (pred[j] == TG_NONE && succ[j] == TG_NONE)) {                 // carried over unchanged
        const uint32_t id = newid_of_start[j];
        newlen[id] = len; onb[id] = nb[j]; orep[id] = rep[j]; op[id] = p[j];
    } else if (succ[j] == TG_NONE || ring[succ[j]]) {                            // tail of a path: it knows the totals
        const uint32_t id = newid_of_start[root[j]];
        newlen[id] = val[j] + (pred[j] == TG_NONE ? len : len - (uint64_t)i);
        onb[id] = agg[j].nb; orep[id] = agg[j].rep; op[id] = agg[j].p;
    }
}
__global__ void k_tags_copy(uint32_t n, const char *chars, const uint64_t *off, const uint32_t *pred, const uint32_t *ring,
                            const uint32_t *root, const uint64_t *val, const uint32_t *newid_of_start, const uint64_t *newoff, int i,
                            char *out)
{
    const uint32_t j = blockIdx.x;
    if (j >= n) return;
    const uint64_t b = off[j], len = off[j + 1] - b;
    const bool alone = ring[j] || pred[j] == TG_NONE;
    const uint32_t id = newid_of_start[alone ? j : root[j]];
    const uint64_t skip = alone ? 0 : (uint64_t)i;             // :183-184 strcat(merge_seq, &right->seq[i])
    char *dst = out + newoff[id] + (alone ? 0 : val[j]);
    for (uint64_t t = threadIdx.x; t < len - skip; t += blockDim.x) dst[t] = chars[b + skip + t];
}

// seqs: n k-mers of k bases back to back (host); pvalue[n]; rep_id[n] = the caller's id of every k-mer (its line number).
// The caller assembles the up- and the down-regulated k-mers separately (mergeTags.c:296-301), one call each.
extern "C" int dk_merge_tags(dk_ctx *ctx, const char *seqs, uint64_t n, int k, const double *pvalue, const uint32_t *rep_id, int min_k,
                             dk_contigs **out)
{
    if (!ctx || !out || (n && (!seqs || !pvalue || !rep_id))) return dk_fail(ctx, DK_ERR_ARG, "merge_tags: NULL argument");
    if (k < 2 || k > 31) return dk_fail(ctx, DK_ERR_ARG, "merge_tags: k=%d outside [2,31]", k);
    if (n >= 0x7FFFFFFFull) return dk_fail(ctx, DK_ERR_ARG, "merge_tags: too many k-mers");
    if (min_k < 1) min_k = 1;
    *out = nullptr;
    dk_contigs *res = new dk_contigs();
    if (!n) { res->off.push_back(0); *out = res; return DK_OK; }
    cudaStream_t st = ctx->stream;
    uint32_t N = (uint32_t)n;
    uint64_t nchars = n * (uint64_t)k;
    int rc = DK_OK;
#define TG_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess && rc == DK_OK) rc = dk_fail(ctx, DK_ERR_CUDA, "merge_tags: %s: %s", #call, cudaGetErrorString(e__)); } while (0)
#define TG_ALLOC(T, name, count) T *name = (T *)dk_dev_alloc(ctx, sizeof(T) * ((size_t)(count) + 4)); if (!name && rc == DK_OK) rc = DK_ERR_NOMEM
    // current generation
    TG_ALLOC(char, chars, nchars); TG_ALLOC(uint64_t, off, N + 1); TG_ALLOC(uint32_t, nb, N); TG_ALLOC(uint32_t, rep, N); TG_ALLOC(double, p, N);
    // next generation (never more contigs or characters than the current one)
    TG_ALLOC(char, chars2, nchars); TG_ALLOC(uint64_t, off2, N + 1); TG_ALLOC(uint32_t, nb2, N); TG_ALLOC(uint32_t, rep2, N); TG_ALLOC(double, p2, N);
    TG_ALLOC(uint64_t, lk, N); TG_ALLOC(uint64_t, rk, N); TG_ALLOC(uint64_t, lk2, N); TG_ALLOC(uint64_t, rk2, N);
    TG_ALLOC(uint32_t, lj, N); TG_ALLOC(uint32_t, rj, N); TG_ALLOC(uint32_t, lj2, N); TG_ALLOC(uint32_t, rj2, N);
    TG_ALLOC(uint32_t, succ, N); TG_ALLOC(uint32_t, pred, N); TG_ALLOC(uint32_t, anc, N); TG_ALLOC(uint32_t, anc2, N);
    TG_ALLOC(uint32_t, root, N); TG_ALLOC(uint32_t, root2, N); TG_ALLOC(uint64_t, val, N); TG_ALLOC(uint64_t, val2, N);
    TG_ALLOC(TgAgg, agg, N); TG_ALLOC(TgAgg, agg2, N); TG_ALLOC(uint32_t, start, N + 1); TG_ALLOC(uint32_t, newid, N + 2);
    TG_ALLOC(uint32_t, ring, N + 1); TG_ALLOC(uint32_t, ringsum, N + 2); TG_ALLOC(uint64_t, newlen, N + 1); TG_ALLOC(uint32_t, flag, 16);
    std::vector<void *> all = {chars, off, nb, rep, p, chars2, off2, nb2, rep2, p2, lk, rk, lk2, rk2, lj, rj, lj2, rj2, succ, pred, anc, anc2,
                               root, root2, val, val2, agg, agg2, start, newid, ring, ringsum, newlen, flag};
    auto cleanup = [&]() { for (void *q : all) dk_dev_release(ctx, q); };
    if (rc != DK_OK) { cleanup(); delete res; return rc; }
    {
        std::vector<uint64_t> h_off(n + 1);
        std::vector<uint32_t> ones(n, 1u);
        for (uint64_t j = 0; j <= n; j++) h_off[j] = j * (uint64_t)k;
        TG_CUDA(cudaMemcpyAsync(chars, seqs, nchars, cudaMemcpyHostToDevice, st));
        TG_CUDA(cudaMemcpyAsync(off, h_off.data(), 8 * (n + 1), cudaMemcpyHostToDevice, st));
        TG_CUDA(cudaMemcpyAsync(nb, ones.data(), 4 * n, cudaMemcpyHostToDevice, st));
        TG_CUDA(cudaMemcpyAsync(rep, rep_id, 4 * n, cudaMemcpyHostToDevice, st));
        TG_CUDA(cudaMemcpyAsync(p, pvalue, 8 * n, cudaMemcpyHostToDevice, st));
        TG_CUDA(cudaMemsetAsync(flag, 0, 64, st));
        k_tags_check<<<(unsigned)((nchars + TG_T - 1) / TG_T), TG_T, 0, st>>>(chars, nchars, flag); ctx->launches++;
        uint32_t bad = 0;
        TG_CUDA(cudaMemcpyAsync(&bad, flag, 4, cudaMemcpyDeviceToHost, st));
        TG_CUDA(cudaStreamSynchronize(st));                  // host staging vectors die here
        if (rc == DK_OK && bad) rc = dk_fail(ctx, DK_ERR_UNSUPPORTED, "merge_tags: k-mers must be upper-case ACGT");
    }
    uint64_t rings_total = 0;
    for (int i = k - 1; i >= min_k && rc == DK_OK && N > 1; i--) {
        const unsigned g = (N + TG_T - 1) / TG_T;
        k_tags_ends<<<g, TG_T, 0, st>>>(chars, off, N, i, lk, rk, lj, rj); ctx->launches++;
        uint64_t *slk, *srk;
        uint32_t *slj, *srj;
        rc = dk_sort_pairs_u64(ctx, lk, lj, lk2, lj2, N, &slk, &slj);
        if (rc == DK_OK) rc = dk_sort_pairs_u64(ctx, rk, rj, rk2, rj2, N, &srk, &srj);
        if (rc != DK_OK) break;
        TG_CUDA(cudaMemsetAsync(succ, 0xFF, 4 * (size_t)N, st));
        TG_CUDA(cudaMemsetAsync(pred, 0xFF, 4 * (size_t)N, st));
        k_tags_links<<<g, TG_T, 0, st>>>(srk, srj, slk, slj, N, succ, pred); ctx->launches++;
        k_tags_rank_init<<<g, TG_T, 0, st>>>(off, pred, nb, rep, p, N, i, anc, val, root, agg); ctx->launches++;
        uint32_t *a0 = anc, *a1 = anc2, *r0 = root, *r1 = root2;
        uint64_t *v0 = val, *v1 = val2;
        TgAgg *g0 = agg, *g1 = agg2;
        int max_rounds = 2;
        while ((1ull << (max_rounds - 2)) < (uint64_t)N) max_rounds++;
        for (int round = 0; round < max_rounds && rc == DK_OK; round++) {
            TG_CUDA(cudaMemsetAsync(flag, 0, 4, st));
            k_tags_rank_step<<<g, TG_T, 0, st>>>(N, a0, v0, r0, g0, a1, v1, r1, g1, flag); ctx->launches++;
            std::swap(a0, a1); std::swap(v0, v1); std::swap(r0, r1); std::swap(g0, g1);
            uint32_t active = 0;
            TG_CUDA(cudaMemcpyAsync(&active, flag, 4, cudaMemcpyDeviceToHost, st));
            TG_CUDA(cudaStreamSynchronize(st));
            if (!active) break;
        }
        if (rc != DK_OK) break;
        k_tags_classify<<<g, TG_T, 0, st>>>(N, a0, pred, succ, start, ring); ctx->launches++;
        rc = dk_scan_excl_u32(ctx, start, newid, N);
        if (rc == DK_OK) rc = dk_scan_excl_u32(ctx, ring, ringsum, N);
        uint32_t N2 = 0, nring = 0;
        TG_CUDA(cudaMemcpyAsync(&N2, newid + N, 4, cudaMemcpyDeviceToHost, st));
        TG_CUDA(cudaMemcpyAsync(&nring, ringsum + N, 4, cudaMemcpyDeviceToHost, st));
        TG_CUDA(cudaStreamSynchronize(st));
        if (rc != DK_OK) break;
        rings_total += nring;
        if (N2 == N) continue;                                // nothing merged at this overlap
        k_tags_newlen<<<g, TG_T, 0, st>>>(N, off, pred, succ, ring, r0, v0, g0, newid, nb, rep, p, i, newlen, nb2, rep2, p2); ctx->launches++;
        rc = dk_scan_excl_u64(ctx, newlen, off2, N2);
        if (rc != DK_OK) break;
        k_tags_copy<<<N, 64, 0, st>>>(N, chars, off, pred, ring, r0, v0, newid, off2, i, chars2); ctx->launches++;
        TG_CUDA(cudaGetLastError());
        std::swap(chars, chars2); std::swap(off, off2); std::swap(nb, nb2); std::swap(rep, rep2); std::swap(p, p2);
        N = N2;
    }
    if (rc == DK_OK) {
        res->n = N; res->rings = rings_total;
        res->rep.resize(N); res->nb.resize(N); res->off.resize((size_t)N + 1);
        TG_CUDA(cudaMemcpyAsync(res->off.data(), off, 8 * ((size_t)N + 1), cudaMemcpyDeviceToHost, st));
        TG_CUDA(cudaMemcpyAsync(res->rep.data(), rep, 4 * (size_t)N, cudaMemcpyDeviceToHost, st));
        TG_CUDA(cudaMemcpyAsync(res->nb.data(), nb, 4 * (size_t)N, cudaMemcpyDeviceToHost, st));
        TG_CUDA(cudaStreamSynchronize(st));
        if (rc == DK_OK) {
            res->nchars = res->off[N];
            res->chars.resize(res->nchars + 1);
            TG_CUDA(cudaMemcpyAsync(res->chars.data(), chars, res->nchars, cudaMemcpyDeviceToHost, st));
            TG_CUDA(cudaStreamSynchronize(st));
        }
    }
#undef TG_CUDA
#undef TG_ALLOC
    all = {chars, off, nb, rep, p, chars2, off2, nb2, rep2, p2, lk, rk, lk2, rk2, lj, rj, lj2, rj2, succ, pred, anc, anc2,
           root, root2, val, val2, agg, agg2, start, newid, ring, ringsum, newlen, flag};      // the generations were swapped
    cleanup();
    if (rc != DK_OK) { delete res; return rc; }
    *out = res;
    return DK_OK;
}

extern "C" uint64_t dk_contigs_count(const dk_contigs *c) { return c ? c->n : 0; }
extern "C" uint64_t dk_contigs_chars(const dk_contigs *c) { return c ? c->nchars : 0; }
extern "C" uint64_t dk_contigs_rings(const dk_contigs *c) { return c ? c->rings : 0; }
// rep[n], nb[n], off[n+1] (contig j is chars[off[j] .. off[j+1])), chars[dk_contigs_chars]; any may be NULL
extern "C" int dk_contigs_export(const dk_contigs *c, uint32_t *rep, uint32_t *nb, uint64_t *off, char *chars)
{
    if (!c) return DK_ERR_ARG;
    if (rep && c->n) memcpy(rep, c->rep.data(), 4 * c->n);
    if (nb && c->n) memcpy(nb, c->nb.data(), 4 * c->n);
    if (off) memcpy(off, c->off.data(), 8 * (c->n + 1));
    if (chars && c->nchars) memcpy(chars, c->chars.data(), c->nchars);
    return DK_OK;
}
extern "C" void dk_contigs_free(dk_contigs *c) { delete c; }
