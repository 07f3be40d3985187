// dk_api.cu — C-ABI glue: host-buffer entry points, matrix import/export, whole-path driver.
#include <math.h>
#include <stdlib.h>
#include <algorithm>
#include <utility>
#include "dk_common.cuh"

void dk_matrix_release(dk_ctx *ctx, dk_matrix *m);
void dk_ttest_release(dk_ctx *ctx, dk_ttest *t);

extern "C" int dk_join_counts_dev(dk_ctx *ctx, int S, const uint64_t *const *keys_dev, const int32_t *const *cnts_dev,
                                  const uint64_t *n, int min_rec, int min_abund, int emit_stale_tail, dk_matrix **out)
{
    if (!ctx || !keys_dev || !cnts_dev || !n) return dk_fail(ctx, DK_ERR_ARG, "join_counts: NULL argument");
    return dk_join_impl(ctx, S, keys_dev, cnts_dev, n, min_rec, min_abund, emit_stale_tail, out);
}

// upload S host dumps into two device slabs (each sample 256-byte aligned)
static int upload_dumps(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                        unsigned char **slab, std::vector<const uint64_t *> &kd, std::vector<const int32_t *> &cd)
{
    size_t tot = 0;
    std::vector<size_t> ko(S), co(S);
    for (int s = 0; s < S; s++) { ko[s] = tot; tot += dk_round_up(n[s] * 8 + 16, 256); }
    for (int s = 0; s < S; s++) { co[s] = tot; tot += dk_round_up(n[s] * 4 + 16, 256); }
    *slab = (unsigned char *)dk_dev_alloc(ctx, tot);
    if (!*slab) return DK_ERR_NOMEM;
    for (int s = 0; s < S; s++) {
        kd[s] = (const uint64_t *)(*slab + ko[s]);
        cd[s] = (const int32_t *)(*slab + co[s]);
        if (n[s]) {
            DK_CUDA(ctx, cudaMemcpyAsync(*slab + ko[s], keys[s], n[s] * 8, cudaMemcpyHostToDevice, ctx->stream));
            DK_CUDA(ctx, cudaMemcpyAsync(*slab + co[s], cnts[s], n[s] * 4, cudaMemcpyHostToDevice, ctx->stream));
        }
    }
    return DK_OK;
}

// ---- pipelined host join ------------------------------------------------------------
// Large host inputs are cut into key-range slices (rows are independent across key ranges): while slice c
// is merged on the compute stream, slice c+1 is copied host->device on a second stream into the other half
// of a double buffer. The per-slice matrices are then spliced into one (chunk tables re-based; reference
// order == slice order). The H2D copy of 12 B/record is the floor of this entry point; the merge hides under it.
__global__ void k_concat_tables(const uint32_t *order_c, const uint64_t *lbase_c, uint32_t nch, uint32_t chunk_base,
                                uint32_t pos_base, uint64_t row_base, uint32_t *order, uint64_t *lbase, int last)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nch) { order[pos_base + i] = order_c[i] + chunk_base; lbase[pos_base + i] = lbase_c[i] + row_base; }
    if (last && i == 0) lbase[pos_base + nch] = lbase_c[nch] + row_base;
}

static int concat_matrices(dk_ctx *ctx, std::vector<dk_matrix *> &parts, dk_matrix **out)
{
    cudaStream_t st = ctx->stream;
    const int S = parts[0]->S;
    uint64_t nch = 0, R = 0, U = 0;
    for (auto *p : parts) { nch += p->n_chunks; R += p->R; U += p->U; }
    if (nch >= 0xFFFFFFF0ull) return dk_fail(ctx, DK_ERR_NOMEM, "join: too many chunks");
    dk_matrix *m = new dk_matrix();
    m->S = S; m->R = R; m->U = U; m->n_chunks = nch; m->cap_chunks = nch + 1;
    m->keys = (uint64_t *)dk_dev_alloc(ctx, 8 * m->cap_chunks * DK_CH);
    m->cnts = (int32_t *)dk_dev_alloc(ctx, 4 * m->cap_chunks * DK_CH * S);
    m->chunk_rows = (uint32_t *)dk_dev_alloc(ctx, 4 * m->cap_chunks);
    m->order = (uint32_t *)dk_dev_alloc(ctx, 4 * (nch + 2));
    m->lbase = (uint64_t *)dk_dev_alloc(ctx, 8 * (nch + 3));
    if (!m->keys || !m->cnts || !m->chunk_rows || !m->order || !m->lbase) { dk_matrix_release(ctx, m); return DK_ERR_NOMEM; }
    uint64_t cb = 0, rb = 0;
    for (size_t c = 0; c < parts.size(); c++) {
        dk_matrix *p = parts[c];
        const uint32_t n = (uint32_t)p->n_chunks;
        const int last = (c + 1 == parts.size());
        if (n) {
            DK_CUDA(ctx, cudaMemcpyAsync(m->keys + cb * DK_CH, p->keys, 8 * (size_t)n * DK_CH, cudaMemcpyDeviceToDevice, st));
            DK_CUDA(ctx, cudaMemcpyAsync(m->cnts + cb * DK_CH * S, p->cnts, 4 * (size_t)n * DK_CH * S, cudaMemcpyDeviceToDevice, st));
            DK_CUDA(ctx, cudaMemcpyAsync(m->chunk_rows + cb, p->chunk_rows, 4 * (size_t)n, cudaMemcpyDeviceToDevice, st));
        }
        k_concat_tables<<<(n + 256) / 256, 256, 0, st>>>(p->order, p->lbase, n, (uint32_t)cb, (uint32_t)cb, rb, m->order, m->lbase, last);
        DK_LAUNCHED(ctx);
        cb += n; rb += p->R;
    }
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    for (auto *p : parts) dk_matrix_release(ctx, p);
    parts.clear();
    *out = m;
    return DK_OK;
}

// cut[s][c] .. cut[s][c+1] = the records of dump s in key-range slice c. Slice edges are k-mers, so equal k-mers of
// different dumps never straddle a slice.
typedef std::vector<std::vector<uint64_t>> dk_cuts;
static void cuts_from_edges(int S, const uint64_t *const *keys, const uint64_t *n, const std::vector<uint64_t> &edges, dk_cuts &cut)
{
    const int ns = (int)edges.size() + 1;
    cut.assign(S, std::vector<uint64_t>(ns + 1, 0));
    for (int s = 0; s < S; s++) {
        cut[s][ns] = n[s];
        for (int c = 1; c < ns; c++) {
            uint64_t lo = cut[s][c - 1], hi = n[s];
            const uint64_t edge = edges[c - 1];
            while (lo < hi) { uint64_t mid = (lo + hi) >> 1; if (keys[s][mid] < edge) lo = mid + 1; else hi = mid; }
            cut[s][c] = lo;
        }
    }
}

struct dk_h2d_stats { double seconds = 0; uint64_t bytes = 0; };

// Walks the slices: while slice c is joined (and handed to on_part) on the compute stream, slice c+1 is copied
// host->device on a second stream into the other half of a double buffer. on_part owns the matrix it receives.
// first_only: stop after slice 0 (the head of the table, for the reference-mode NF).
template <class F>
static int join_host_slices(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const dk_cuts &cut,
                            int min_rec, int min_abund, bool first_only, dk_h2d_stats *hs, F on_part)
{
    const int nslices = first_only ? 1 : (int)cut[0].size() - 1;
    size_t slab_bytes = 0;
    for (int c = 0; c < nslices; c++) {
        size_t b = 0;
        for (int s = 0; s < S; s++) { uint64_t k = cut[s][c + 1] - cut[s][c]; b += dk_round_up(k * 8 + 16, 256) + dk_round_up(k * 4 + 16, 256); }
        if (b > slab_bytes) slab_bytes = b;
    }
    unsigned char *slab[2] = {(unsigned char *)dk_dev_alloc(ctx, slab_bytes), (unsigned char *)dk_dev_alloc(ctx, slab_bytes)};
    if (!slab[0] || !slab[1]) { dk_dev_release(ctx, slab[0]); dk_dev_release(ctx, slab[1]); return DK_ERR_NOMEM; }
    cudaStream_t cs = nullptr;
    cudaEvent_t up[2] = {nullptr, nullptr}, t0[2] = {nullptr, nullptr};
    DK_CUDA(ctx, cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) { DK_CUDA(ctx, cudaEventCreate(&up[i])); DK_CUDA(ctx, cudaEventCreate(&t0[i])); }
    std::vector<std::vector<const uint64_t *>> kd(2, std::vector<const uint64_t *>(S));
    std::vector<std::vector<const int32_t *>> cd(2, std::vector<const int32_t *>(S));
    std::vector<uint64_t> nn(S);
    auto enqueue = [&](int c) -> int {
        const int b = c & 1;
        size_t off = 0;
        DK_CUDA(ctx, cudaEventRecord(t0[b], cs));
        for (int s = 0; s < S; s++) {
            const uint64_t k0 = cut[s][c], k = cut[s][c + 1] - k0;
            kd[b][s] = (const uint64_t *)(slab[b] + off);
            if (k) DK_CUDA(ctx, cudaMemcpyAsync(slab[b] + off, keys[s] + k0, k * 8, cudaMemcpyHostToDevice, cs));
            off += dk_round_up(k * 8 + 16, 256);
            cd[b][s] = (const int32_t *)(slab[b] + off);
            if (k) DK_CUDA(ctx, cudaMemcpyAsync(slab[b] + off, cnts[s] + k0, k * 4, cudaMemcpyHostToDevice, cs));
            off += dk_round_up(k * 4 + 16, 256);
            if (hs) hs->bytes += k * 12;
        }
        DK_CUDA(ctx, cudaEventRecord(up[b], cs));
        return DK_OK;
    };
    int rc = enqueue(0);
    double ms_join = 0, ms_merge = 0;
    for (int c = 0; c < nslices && rc == DK_OK; c++) {
        // slice c+1 goes into the other buffer: its previous user (slice c-1) has been joined AND consumed by on_part
        if (c + 1 < nslices) rc = enqueue(c + 1);
        if (rc != DK_OK) break;
        if (cudaStreamWaitEvent(ctx->stream, up[c & 1], 0) != cudaSuccess) { rc = dk_fail(ctx, DK_ERR_CUDA, "join: stream wait failed"); break; }
        for (int s = 0; s < S; s++) nn[s] = cut[s][c + 1] - cut[s][c];
        dk_matrix *m = nullptr;
        rc = dk_join_impl(ctx, S, kd[c & 1].data(), cd[c & 1].data(), nn.data(), min_rec, min_abund, 0, &m);
        if (rc == DK_OK) {
            ms_join += ctx->stage_ms[0]; ms_merge += ctx->stage_ms[4];
            if (hs) { float ms = 0; if (cudaEventElapsedTime(&ms, t0[c & 1], up[c & 1]) == cudaSuccess) hs->seconds += ms * 1e-3; }
            rc = on_part(m, c);
        }
    }
    cudaStreamSynchronize(cs);
    cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < 2; i++) { cudaEventDestroy(up[i]); cudaEventDestroy(t0[i]); }
    cudaStreamDestroy(cs);
    dk_dev_release(ctx, slab[0]); dk_dev_release(ctx, slab[1]);
    ctx->stage_ms[0] = ms_join; ctx->stage_ms[4] = ms_merge;
    return rc;
}

// optional copy of the filtered matrix to caller-owned (pinned) host buffers, in reference order, WHILE the join runs:
// the rows of slice c leave on a third stream (device -> host) beside the host -> device copies of later slices
struct dk_host_out { uint64_t *keys = nullptr; int32_t *cnts = nullptr; uint64_t cap = 0, written = 0; };

static int join_host_pipelined(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                               int min_rec, int min_abund, int nslices, dk_matrix **out, dk_host_out *ho = nullptr)
{
    // slice edges: quantiles of the longest dump; per dump a binary search on the host arrays
    int big = 0;
    for (int s = 1; s < S; s++) if (n[s] > n[big]) big = s;
    std::vector<uint64_t> edges;
    for (int c = 1; c < nslices; c++) edges.push_back(keys[big][n[big] / nslices * c]);
    dk_cuts cut;
    cuts_from_edges(S, keys, n, edges, cut);
    std::vector<dk_matrix *> parts;
    cudaStream_t ds = nullptr;
    cudaEvent_t ready[2] = {nullptr, nullptr}, gone[2] = {nullptr, nullptr};
    uint64_t *sk[2] = {nullptr, nullptr}; int32_t *sc[2] = {nullptr, nullptr};
    uint64_t scap[2] = {0, 0};
    bool used[2] = {false, false};
    if (ho) {
        DK_CUDA(ctx, cudaStreamCreateWithFlags(&ds, cudaStreamNonBlocking));
        for (int i = 0; i < 2; i++) { DK_CUDA(ctx, cudaEventCreateWithFlags(&ready[i], cudaEventDisableTiming)); DK_CUDA(ctx, cudaEventCreateWithFlags(&gone[i], cudaEventDisableTiming)); }
    }
    int rc = join_host_slices(ctx, S, keys, cnts, cut, min_rec, min_abund, false, nullptr, [&](dk_matrix *m, int c) -> int {
        parts.push_back(m);
        if (!ho || !m->R) return DK_OK;
        const uint64_t rows = m->R;
        if (ho->written + rows > ho->cap) { ho->written += rows; return DK_OK; }           // keep counting: the caller learns the need
        const int b = c & 1;
        if (used[b] && cudaEventSynchronize(gone[b]) != cudaSuccess) return dk_fail(ctx, DK_ERR_CUDA, "join: host copy-out failed");
        if (scap[b] < rows) {
            dk_dev_release(ctx, sk[b]); dk_dev_release(ctx, sc[b]);
            scap[b] = rows + rows / 8 + 1024;
            sk[b] = (uint64_t *)dk_dev_alloc(ctx, 8 * scap[b]);
            sc[b] = (int32_t *)dk_dev_alloc(ctx, 4 * scap[b] * S);
            if (!sk[b] || !sc[b]) return DK_ERR_NOMEM;
        }
        DK_TRY(dk_matrix_export_dev(ctx, m, 0, rows, sk[b], sc[b]));                       // dense, reference order
        DK_CUDA(ctx, cudaEventRecord(ready[b], ctx->stream));
        DK_CUDA(ctx, cudaStreamWaitEvent(ds, ready[b], 0));
        DK_CUDA(ctx, cudaMemcpyAsync(ho->keys + ho->written, sk[b], 8 * rows, cudaMemcpyDeviceToHost, ds));
        DK_CUDA(ctx, cudaMemcpyAsync(ho->cnts + ho->written * (size_t)S, sc[b], 4 * rows * S, cudaMemcpyDeviceToHost, ds));
        DK_CUDA(ctx, cudaEventRecord(gone[b], ds));
        used[b] = true;
        ho->written += rows;
        return DK_OK;
    });
    const double ms_join = ctx->stage_ms[0], ms_merge = ctx->stage_ms[4];
    if (ho) {
        cudaStreamSynchronize(ds);
        for (int i = 0; i < 2; i++) { cudaEventDestroy(ready[i]); cudaEventDestroy(gone[i]); dk_dev_release(ctx, sk[i]); dk_dev_release(ctx, sc[i]); }
        cudaStreamDestroy(ds);
        if (rc == DK_OK && ho->written > ho->cap)
            rc = dk_fail(ctx, DK_ERR_ARG, "matrix_out: %llu rows, the host buffers hold %llu", (unsigned long long)ho->written, (unsigned long long)ho->cap);
    }
    if (rc != DK_OK) { for (auto *p : parts) dk_matrix_release(ctx, p); return rc; }
    rc = concat_matrices(ctx, parts, out);
    ctx->stage_ms[0] = ms_join; ctx->stage_ms[4] = ms_merge;
    return rc;
}

static int join_counts_host(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                            int min_rec, int min_abund, int emit_stale_tail, dk_matrix **out, dk_host_out *ho);

static int join_counts_host(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                            int min_rec, int min_abund, int emit_stale_tail, dk_matrix **out, dk_host_out *ho)
{
    if (!ctx || !keys || !cnts || !n) return dk_fail(ctx, DK_ERR_ARG, "join_counts: NULL argument");
    if (S < 1 || S > DK_MAX_SAMPLES) return dk_fail(ctx, DK_ERR_ARG, "join: S=%d outside [1,%d]", S, DK_MAX_SAMPLES);
    {   // large inputs: overlap the host->device copies with the merge, slice by slice (DK_JOIN_SLICES=1 disables)
        uint64_t total = 0;
        for (int s = 0; s < S; s++) total += n[s];
        int nslices = total >= (64ull << 20) ? 8 : 1;
        if (const char *e = getenv("DK_JOIN_SLICES")) nslices = atoi(e);
        if (nslices > 64) nslices = 64;
        if (nslices > 1 && !emit_stale_tail && total >= (uint64_t)nslices * 1024)
            return join_host_pipelined(ctx, S, keys, cnts, n, min_rec, min_abund, nslices, out, ho);
    }
    unsigned char *slab = nullptr;
    std::vector<const uint64_t *> kd(S);
    std::vector<const int32_t *> cd(S);
    DK_TRY(upload_dumps(ctx, S, keys, cnts, n, &slab, kd, cd));
    int rc = dk_join_impl(ctx, S, kd.data(), cd.data(), n, min_rec, min_abund, emit_stale_tail, out);
    cudaStreamSynchronize(ctx->stream);
    dk_dev_release(ctx, slab);
    if (rc == DK_OK && ho) {                                     // small input: one plain copy after the join
        ho->written = (*out)->R;
        if (ho->written > ho->cap) {
            dk_matrix_release(ctx, *out); *out = nullptr;
            return dk_fail(ctx, DK_ERR_ARG, "matrix_out: %llu rows, the host buffers hold %llu", (unsigned long long)ho->written, (unsigned long long)ho->cap);
        }
        rc = dk_matrix_export(ctx, *out, 0, ho->written, ho->keys, ho->cnts);
    }
    return rc;
}

extern "C" int dk_join_counts(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts,
                              const uint64_t *n, int min_rec, int min_abund, int emit_stale_tail, dk_matrix **out)
{
    return join_counts_host(ctx, S, keys, cnts, n, min_rec, min_abund, emit_stale_tail, out, nullptr);
}


// ---- count-balanced shard boundaries --------------------------------------------------------------
// Shards are k-mer ranges; real dumps are far from uniform in the key (low-complexity sequence clusters), so equal
// slices of the key SPACE give unequal work. Every dump is sampled at regular positions (a sample stands for the
// n_s / m_s records around it), the samples of all dumps are ordered by key and the boundaries are read off where the
// cumulative record count crosses g/G of the total. The error per boundary is at most one sampling stride per dump.
__global__ void k_sample_keys(const uint64_t *keys, uint64_t n, uint32_t m, uint64_t *out)
{
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < m) out[j] = keys[(uint64_t)(((double)j + 0.5) * (double)n / (double)m)];
}

extern "C" int dk_key_splitters(dk_ctx *ctx, int S, const uint64_t *const *keys, const uint64_t *n, int device_inputs, int G,
                                uint64_t *splitters)
{
    if (!ctx || !keys || !n || (G > 1 && !splitters)) return dk_fail(ctx, DK_ERR_ARG, "key_splitters: NULL argument");
    if (S < 1 || S > DK_MAX_SAMPLES || G < 1 || G > 4096) return dk_fail(ctx, DK_ERR_ARG, "key_splitters: S=%d G=%d out of range", S, G);
    if (G == 1) return DK_OK;
    // samples per dump: a boundary is off by at most one sampling stride per dump, i.e. total/MS records in all,
    // which is G/MS of a shard: 256 samples per shard and dump keep that below 0.4 %
    uint32_t MS = 256u * (uint32_t)G;
    MS = MS < 1024u ? 1024u : (MS > 16384u ? 16384u : MS);
    std::vector<std::pair<uint64_t, double>> smp;
    smp.reserve((size_t)S * MS);
    std::vector<uint64_t> h(MS);
    uint64_t *d = device_inputs ? (uint64_t *)dk_dev_alloc(ctx, sizeof(uint64_t) * MS) : nullptr;
    if (device_inputs && !d) return DK_ERR_NOMEM;
    double total = 0;
    for (int s = 0; s < S; s++) {
        if (!n[s]) continue;
        const uint32_t m = n[s] < MS ? (uint32_t)n[s] : MS;
        if (device_inputs) {
            k_sample_keys<<<(m + 255) / 256, 256, 0, ctx->stream>>>(keys[s], n[s], m, d); DK_LAUNCHED(ctx);
            DK_CUDA(ctx, cudaMemcpyAsync(h.data(), d, sizeof(uint64_t) * m, cudaMemcpyDeviceToHost, ctx->stream));
            DK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        } else
            for (uint32_t j = 0; j < m; j++) h[j] = keys[s][(uint64_t)(((double)j + 0.5) * (double)n[s] / (double)m)];
        const double w = (double)n[s] / (double)m;
        for (uint32_t j = 0; j < m; j++) smp.emplace_back(h[j], w);
        total += (double)n[s];
    }
    dk_dev_release(ctx, d);
    std::sort(smp.begin(), smp.end(), [](const std::pair<uint64_t, double> &a, const std::pair<uint64_t, double> &b) { return a.first < b.first; });
    double cum = 0;
    size_t i = 0;
    for (int g = 1; g < G; g++) {
        const double want = total * (double)g / (double)G;
        while (i < smp.size() && cum + smp[i].second <= want) cum += smp[i++].second;
        // shard g starts at this key: records with key < splitters[g-1] belong to shards < g (equal keys never straddle)
        splitters[g - 1] = i < smp.size() ? smp[i].first : DK_SENTINEL;
    }
    return DK_OK;
}

__global__ void k_identity_chunks(uint32_t *order, uint64_t *lbase, uint32_t *chunk_rows, uint32_t n_chunks, uint64_t R);

// ---- chunk-streamed path for inputs larger than HBM (BASELINE.json configs[3]) ----------------------------------
// The dumps stay in (pinned) host memory. The key space is cut into chunks of about chunk_records records (equal-count
// boundaries, dk_key_splitters); chunk c+1 is copied in on a second stream while chunk c is joined, tested and reduced
// to its CANDIDATE rows (raw p <= p_cut): k-mer, counts, p, log2FC of those rows are appended to compact device arrays
// in k-mer order and the chunk's matrix is dropped. BH needs nothing else (dk_bh.cu: rows with p > p_cut can never
// reach an adjusted p <= p_cut), so the result is an ordinary (matrix, test) pair over the candidate rows that the
// existing BH / filter / export entry points take — with R_global = all filtered rows of all chunks and ranks.
__global__ void k_flag_le(const double *p, uint64_t n, double cut, uint32_t *flag)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) flag[i] = p[i] <= cut ? 1u : 0u;
}
// one warp per flagged row: its key, counts, p and log2FC go to position pos[row] of the compact arrays
__global__ void k_gather_flagged(const uint64_t *keys, const int32_t *cnts, const uint64_t *lbase, const uint32_t *order,
                                 uint32_t n_chunks, int S, uint64_t R, const double *p, const double *fc, const uint32_t *pos,
                                 uint64_t base, uint64_t *okeys, int32_t *ocnts, double *op, double *ofc, uint64_t rownum_base)
{
    const uint64_t row = (uint64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    if (row >= R) return;
    const uint32_t o = pos[row];
    if (pos[row + 1] == o) return;                               // not flagged (pos = exclusive scan of the flags)
    uint32_t lo = 0, hi = n_chunks;
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (lbase[mid] <= row) lo = mid; else hi = mid; }
    const size_t ph = (size_t)order[lo] * DK_CH + (size_t)(row - lbase[lo]);
    const int lane = threadIdx.x & 31;
    const uint64_t d = base + o;
    for (int j = lane; j < S; j += 32) ocnts[d * (size_t)S + j] = cnts[ph * (size_t)S + j];
    if (lane == 0) {
        okeys[d] = p ? keys[ph] : rownum_base + row;             // collection sweep: the row's number instead of its k-mer
        if (p) { op[d] = p[row]; ofc[d] = fc[row]; }
    }
}
// flag[row] = 1 iff every count of the row is > 0 (the rows computeNF uses, computeNF.c:122-130); one warp per row
__global__ void k_flag_allpos(const int32_t *cnts, const uint64_t *lbase, const uint32_t *order, uint32_t n_chunks, int S,
                              uint64_t R, uint32_t *flag)
{
    const uint64_t row = (uint64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    if (row >= R) return;
    uint32_t lo = 0, hi = n_chunks;
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (lbase[mid] <= row) lo = mid; else hi = mid; }
    const size_t ph = (size_t)order[lo] * DK_CH + (size_t)(row - lbase[lo]);
    const int lane = threadIdx.x & 31;
    bool pos = true;
    for (int j = lane; j < S; j += 32) pos = pos && cnts[ph * (size_t)S + j] > 0;
    pos = __all_sync(0xffffffffu, pos);
    if (lane == 0) flag[row] = pos ? 1u : 0u;
}

struct dk_cand_acc {
    uint64_t n = 0, cap = 0;
    uint64_t *keys = nullptr; int32_t *cnts = nullptr; double *p = nullptr, *fc = nullptr;
};
static int acc_reserve(dk_ctx *ctx, dk_cand_acc &a, int S, uint64_t need)
{
    if (need + 2 * DK_CH <= a.cap) return DK_OK;
    uint64_t cap = a.cap ? a.cap : (1u << 16);
    while (cap < need + 2 * DK_CH) cap *= 2;
    cap = dk_round_up(cap, DK_CH);
    uint64_t *k = (uint64_t *)dk_dev_alloc(ctx, 8 * cap);
    int32_t *c = (int32_t *)dk_dev_alloc(ctx, 4 * cap * S);
    double *pp = (double *)dk_dev_alloc(ctx, 8 * cap), *f = (double *)dk_dev_alloc(ctx, 8 * cap);
    if (!k || !c || !pp || !f) { dk_dev_release(ctx, k); dk_dev_release(ctx, c); dk_dev_release(ctx, pp); dk_dev_release(ctx, f); return DK_ERR_NOMEM; }
    if (a.n) {
        cudaStream_t st = ctx->stream;
        DK_CUDA(ctx, cudaMemcpyAsync(k, a.keys, 8 * a.n, cudaMemcpyDeviceToDevice, st));
        DK_CUDA(ctx, cudaMemcpyAsync(c, a.cnts, 4 * a.n * S, cudaMemcpyDeviceToDevice, st));
        DK_CUDA(ctx, cudaMemcpyAsync(pp, a.p, 8 * a.n, cudaMemcpyDeviceToDevice, st));
        DK_CUDA(ctx, cudaMemcpyAsync(f, a.fc, 8 * a.n, cudaMemcpyDeviceToDevice, st));
        DK_CUDA(ctx, cudaStreamSynchronize(st));
    }
    dk_dev_release(ctx, a.keys); dk_dev_release(ctx, a.cnts); dk_dev_release(ctx, a.p); dk_dev_release(ctx, a.fc);
    a.keys = k; a.cnts = c; a.p = pp; a.fc = f; a.cap = cap;
    return DK_OK;
}
static void acc_drop(dk_ctx *ctx, dk_cand_acc &a)
{
    dk_dev_release(ctx, a.keys); dk_dev_release(ctx, a.cnts); dk_dev_release(ctx, a.p); dk_dev_release(ctx, a.fc);
    a = dk_cand_acc();
}

static int stream_cuts(dk_ctx *ctx, int S, const uint64_t *const *keys, const uint64_t *n, uint64_t chunk_records, dk_cuts &cut)
{
    uint64_t total = 0;
    for (int s = 0; s < S; s++) total += n[s];
    if (chunk_records < 4096) chunk_records = 4096;
    uint64_t G = (total + chunk_records - 1) / chunk_records;
    if (G < 1) G = 1;
    if (G > 4096) return dk_fail(ctx, DK_ERR_ARG, "stream: %llu chunks; raise chunk_records", (unsigned long long)G);
    std::vector<uint64_t> spl(G > 1 ? G - 1 : 0), edges;
    if (G > 1) DK_TRY(dk_key_splitters(ctx, S, keys, n, 0, (int)G, spl.data()));
    for (uint64_t e : spl) if (e != DK_SENTINEL && (edges.empty() || e > edges.back())) edges.push_back(e);
    cuts_from_edges(S, keys, n, edges, cut);
    return DK_OK;
}

extern "C" int dk_stream_nf_reference(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                                      int min_rec, int min_abund, uint64_t chunk_records, uint64_t step_rows, int widen,
                                      double *nf_out, uint64_t *info, int *stop_fired)
{
    if (!ctx || !keys || !cnts || !n || !nf_out) return dk_fail(ctx, DK_ERR_ARG, "stream_nf_reference: NULL argument");
    if (S < 1 || S > DK_MAX_SAMPLES) return dk_fail(ctx, DK_ERR_ARG, "stream: S=%d outside [1,%d]", S, DK_MAX_SAMPLES);
    dk_cuts cut;
    DK_TRY(stream_cuts(ctx, S, keys, n, chunk_records, cut));
    const int nch = (int)cut[0].size() - 1;
    uint64_t inf[4] = {0, 0, 0, 0};
    bool fired = false;
    // computeNF.c:138-168 reads the table from its first row until its early stop: a head of the table answers for the
    // whole table only if the stop fired inside it (rows consumed < rows of the head) and outside the reference's
    // undefined zone. The head is the first chunk; with `widen` it grows to 2, 4, 8 ... chunks until the stop fires or
    // these dumps are exhausted (exact when they ARE the whole table; a rank of a sharded run must not widen).
    for (int k = 1;; k *= 2) {
        const int kk = k < nch ? k : nch;
        dk_cuts head(S, std::vector<uint64_t>(2, 0));
        for (int s = 0; s < S; s++) head[s][1] = cut[s][kk];
        uint64_t Rc = 0;
        int rc = join_host_slices(ctx, S, keys, cnts, head, min_rec, min_abund, true, nullptr, [&](dk_matrix *m, int) -> int {
            Rc = m->R;
            int r = m->R ? dk_compute_nf(ctx, m, DK_NF_REFERENCE, step_rows, nf_out, inf)
                         : dk_fail(ctx, DK_ERR_UNSUPPORTED, "stream_nf_reference: the head of the table has no rows");
            dk_matrix_release(ctx, m);
            return r;
        });
        if (rc != DK_OK) return rc;
        fired = inf[0] < Rc && !inf[2];
        if (fired || kk == nch || !widen) break;
    }
    if (info) memcpy(info, inf, sizeof inf);
    if (stop_fired) *stop_fired = fired ? 1 : 0;
    return DK_OK;
}

// first sweep of the two-sweep form: the rows computeNF can use (every count > 0), with their row numbers
extern "C" int dk_stream_positive_rows(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                                       int min_rec, int min_abund, uint64_t chunk_records, dk_stream_stats *stats, dk_matrix **pos_rows)
{
    if (!ctx || !keys || !cnts || !n || !pos_rows) return dk_fail(ctx, DK_ERR_ARG, "stream_positive_rows: NULL argument");
    *pos_rows = nullptr;
    if (S < 1 || S > DK_MAX_SAMPLES) return dk_fail(ctx, DK_ERR_ARG, "stream: S=%d outside [1,%d]", S, DK_MAX_SAMPLES);
    dk_cuts cut;
    DK_TRY(stream_cuts(ctx, S, keys, n, chunk_records, cut));
    dk_stream_stats st;
    memset(&st, 0, sizeof st);
    st.chunks = cut[0].size() - 1;
    dk_cand_acc acc;
    dk_h2d_stats hs;
    int rc = join_host_slices(ctx, S, keys, cnts, cut, min_rec, min_abund, false, &hs, [&](dk_matrix *m, int) -> int {
        int r = DK_OK;
        uint32_t *flag = nullptr;
        const uint64_t R = m->R;
        if (R) {
            flag = (uint32_t *)dk_dev_alloc(ctx, 4 * (R + 2));
            if (!flag) r = DK_ERR_NOMEM;
            if (r == DK_OK) {
                k_flag_allpos<<<(unsigned)((R + 7) / 8), 256, 0, ctx->stream>>>(m->cnts, m->lbase, m->order, (uint32_t)m->n_chunks, S, R, flag);
                ctx->launches++;
                r = dk_scan_excl_u32(ctx, flag, flag, R);
            }
            uint32_t nc = 0;
            if (r == DK_OK && (cudaMemcpyAsync(&nc, flag + R, 4, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
                               cudaStreamSynchronize(ctx->stream) != cudaSuccess)) r = dk_fail(ctx, DK_ERR_CUDA, "stream: positive count");
            if (r == DK_OK && nc) r = acc_reserve(ctx, acc, S, acc.n + nc);
            if (r == DK_OK && nc) {
                k_gather_flagged<<<(unsigned)((R + 7) / 8), 256, 0, ctx->stream>>>(m->keys, m->cnts, m->lbase, m->order, (uint32_t)m->n_chunks,
                                                                                   S, R, nullptr, nullptr, flag, acc.n, acc.keys, acc.cnts,
                                                                                   acc.p, acc.fc, st.R);
                ctx->launches++;
                if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) r = dk_fail(ctx, DK_ERR_CUDA, "stream: gather");
                acc.n += nc;
            }
        }
        st.U += m->U; st.R += R;
        dk_dev_release(ctx, flag);
        dk_matrix_release(ctx, m);
        return r;
    });
    if (rc != DK_OK) { acc_drop(ctx, acc); return rc; }
    st.n_cand = acc.n; st.h2d_bytes = hs.bytes; st.h2d_seconds = hs.seconds;
    st.ms_join = ctx->stage_ms[0]; st.ms_merge = ctx->stage_ms[4];
    if (stats) *stats = st;
    // a dense matrix whose KEY column holds the row numbers (in this rank's filtered table) of the all-positive rows
    dk_matrix *m = new dk_matrix();
    m->S = S; m->R = acc.n; m->U = st.U;
    if (acc.n == 0) DK_TRY(acc_reserve(ctx, acc, S, 1));
    m->n_chunks = (acc.n + DK_CH - 1) / DK_CH; m->cap_chunks = acc.cap / DK_CH;
    m->keys = acc.keys; m->cnts = acc.cnts;
    dk_dev_release(ctx, acc.p); dk_dev_release(ctx, acc.fc);
    m->chunk_rows = (uint32_t *)dk_dev_alloc(ctx, 4 * m->cap_chunks);
    m->order = (uint32_t *)dk_dev_alloc(ctx, 4 * (m->n_chunks + 2));
    m->lbase = (uint64_t *)dk_dev_alloc(ctx, 8 * (m->n_chunks + 3));
    if (!m->chunk_rows || !m->order || !m->lbase) { dk_matrix_release(ctx, m); return DK_ERR_NOMEM; }
    k_identity_chunks<<<(unsigned)((m->n_chunks + 256) / 256), 256, 0, ctx->stream>>>(m->order, m->lbase, m->chunk_rows, (uint32_t)m->n_chunks, m->R);
    ctx->launches++;
    DK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *pos_rows = m;
    return DK_OK;
}

extern "C" int dk_stream_path(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                              int min_rec, int min_abund, uint64_t chunk_records, const double *nf, const int *idxA, int nA,
                              const int *idxB, int nB, double p_cut, dk_stream_stats *stats, dk_matrix **cand_matrix,
                              dk_ttest **cand_test)
{
    if (!ctx || !keys || !cnts || !n || !nf || !idxA || !idxB || !cand_matrix || !cand_test)
        return dk_fail(ctx, DK_ERR_ARG, "stream_path: NULL argument");
    *cand_matrix = nullptr; *cand_test = nullptr;
    if (S < 1 || S > DK_MAX_SAMPLES) return dk_fail(ctx, DK_ERR_ARG, "stream: S=%d outside [1,%d]", S, DK_MAX_SAMPLES);
    dk_cuts cut;
    DK_TRY(stream_cuts(ctx, S, keys, n, chunk_records, cut));
    dk_stream_stats st;
    memset(&st, 0, sizeof st);
    st.chunks = cut[0].size() - 1;
    dk_cand_acc acc;
    dk_h2d_stats hs;
    cudaEvent_t e0, e1;
    DK_CUDA(ctx, cudaEventCreate(&e0)); DK_CUDA(ctx, cudaEventCreate(&e1));
    cudaEventRecord(e0, ctx->stream);
    double ms_ttest = 0;
    int rc = join_host_slices(ctx, S, keys, cnts, cut, min_rec, min_abund, false, &hs, [&](dk_matrix *m, int) -> int {
        st.U += m->U; st.R += m->R;
        int r = DK_OK;
        dk_ttest *t = nullptr;
        uint32_t *flag = nullptr;
        if (m->R) {
            r = dk_ttest_rows(ctx, m, nf, idxA, nA, idxB, nB, &t);
            if (r == DK_OK) {
                ms_ttest += ctx->stage_ms[2];
                const uint64_t R = m->R;
                flag = (uint32_t *)dk_dev_alloc(ctx, 4 * (R + 2));
                if (!flag) r = DK_ERR_NOMEM;
                if (r == DK_OK) {
                    k_flag_le<<<(unsigned)((R + 255) / 256), 256, 0, ctx->stream>>>(t->p_raw, R, p_cut, flag);
                    ctx->launches++;
                    r = dk_scan_excl_u32(ctx, flag, flag, R);                  // flag[R] = candidates of this chunk
                }
                uint32_t nc = 0;
                if (r == DK_OK) {
                    if (cudaMemcpyAsync(&nc, flag + R, 4, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
                        cudaStreamSynchronize(ctx->stream) != cudaSuccess) r = dk_fail(ctx, DK_ERR_CUDA, "stream: candidate count");
                }
                if (r == DK_OK && nc) r = acc_reserve(ctx, acc, S, acc.n + nc);
                if (r == DK_OK && nc) r = dk_ttest_need_log2fc(ctx, t, nullptr, 0);   // the gather reads it at the flagged rows
                if (r == DK_OK && nc) {
                    k_gather_flagged<<<(unsigned)((R + 7) / 8), 256, 0, ctx->stream>>>(m->keys, m->cnts, m->lbase, m->order, (uint32_t)m->n_chunks,
                                                                                       S, R, t->p_raw, t->log2fc, flag, acc.n, acc.keys,
                                                                                       acc.cnts, acc.p, acc.fc, 0);
                    ctx->launches++;
                    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) r = dk_fail(ctx, DK_ERR_CUDA, "stream: gather");
                    acc.n += nc;
                }
            }
        }
        dk_dev_release(ctx, flag);
        dk_ttest_release(ctx, t);
        dk_matrix_release(ctx, m);
        return r;
    });
    cudaEventRecord(e1, ctx->stream);
    cudaStreamSynchronize(ctx->stream);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (rc != DK_OK) { acc_drop(ctx, acc); return rc; }
    st.n_cand = acc.n; st.h2d_bytes = hs.bytes; st.h2d_seconds = hs.seconds;
    st.ms_join = ctx->stage_ms[0]; st.ms_merge = ctx->stage_ms[4]; st.ms_ttest = ms_ttest; st.ms_total = ms;
    ctx->stage_ms[2] = ms_ttest;
    if (stats) *stats = st;
    if (acc.n == 0) { acc_drop(ctx, acc); return DK_OK; }        // no candidate anywhere: *cand_matrix and *cand_test stay NULL
    // the candidate rows as an ordinary dense matrix + test object
    cudaStream_t cs = ctx->stream;
    dk_matrix *m = new dk_matrix();
    m->S = S; m->R = acc.n; m->U = st.U;
    m->n_chunks = (acc.n + DK_CH - 1) / DK_CH; m->cap_chunks = acc.cap / DK_CH;
    m->keys = acc.keys; m->cnts = acc.cnts;
    m->chunk_rows = (uint32_t *)dk_dev_alloc(ctx, 4 * m->cap_chunks);
    m->order = (uint32_t *)dk_dev_alloc(ctx, 4 * (m->n_chunks + 2));
    m->lbase = (uint64_t *)dk_dev_alloc(ctx, 8 * (m->n_chunks + 3));
    dk_ttest *t = new dk_ttest();
    t->m = m; t->R = acc.n; t->S = S; t->nA = nA; t->nB = nB;
    t->p_raw = acc.p; t->log2fc = acc.fc;
    t->nf_dev = (double *)dk_dev_alloc(ctx, sizeof(double) * S);
    t->idx_dev = (int *)dk_dev_alloc(ctx, sizeof(int) * (nA + nB));
    t->tie_counts = (uint64_t *)dk_dev_alloc(ctx, 64);
    if (!m->chunk_rows || !m->order || !m->lbase || !t->nf_dev || !t->idx_dev || !t->tie_counts) {
        dk_ttest_release(ctx, t); dk_matrix_release(ctx, m);
        return DK_ERR_NOMEM;
    }
    std::vector<int> idx(nA + nB);
    memcpy(idx.data(), idxA, sizeof(int) * nA);
    memcpy(idx.data() + nA, idxB, sizeof(int) * nB);
    k_identity_chunks<<<(unsigned)((m->n_chunks + 255) / 256), 256, 0, cs>>>(m->order, m->lbase, m->chunk_rows, (uint32_t)m->n_chunks, m->R);
    ctx->launches++;
    DK_CUDA(ctx, cudaMemcpyAsync(t->nf_dev, nf, sizeof(double) * S, cudaMemcpyHostToDevice, cs));
    DK_CUDA(ctx, cudaMemcpyAsync(t->idx_dev, idx.data(), sizeof(int) * (nA + nB), cudaMemcpyHostToDevice, cs));
    DK_CUDA(ctx, cudaMemsetAsync(t->tie_counts, 0, 64, cs));
    DK_CUDA(ctx, cudaStreamSynchronize(cs));
    *cand_matrix = m; *cand_test = t;
    return DK_OK;
}

// ---- matrix ---------------------------------------------------------------------
__global__ void k_identity_chunks(uint32_t *order, uint64_t *lbase, uint32_t *chunk_rows, uint32_t n_chunks, uint64_t R)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_chunks) {
        order[i] = i;
        lbase[i] = (uint64_t)i * DK_CH;
        uint64_t left = R - (uint64_t)i * DK_CH;
        chunk_rows[i] = left < DK_CH ? (uint32_t)left : DK_CH;
    }
    if (i == 0) lbase[n_chunks] = R;
}

extern "C" int dk_matrix_from_host(dk_ctx *ctx, int S, uint64_t R, const uint64_t *keys, const int32_t *cnts, dk_matrix **out)
{
    if (!ctx || !out || (R && !cnts)) return dk_fail(ctx, DK_ERR_ARG, "matrix_from_host: NULL argument");
    if (S < 1 || S > DK_MAX_SAMPLES) return dk_fail(ctx, DK_ERR_ARG, "matrix_from_host: S=%d outside [1,%d]", S, DK_MAX_SAMPLES);
    cudaStream_t st = ctx->stream;
    dk_matrix *m = new dk_matrix();
    m->S = S; m->R = R; m->U = 0;
    m->n_chunks = (R + DK_CH - 1) / DK_CH;
    m->cap_chunks = m->n_chunks + 1;
    m->has_keys = keys != nullptr;
    m->keys = (uint64_t *)dk_dev_alloc(ctx, 8 * m->cap_chunks * DK_CH);
    m->cnts = (int32_t *)dk_dev_alloc(ctx, 4 * m->cap_chunks * DK_CH * S);
    m->chunk_rows = (uint32_t *)dk_dev_alloc(ctx, 4 * m->cap_chunks);
    m->order = (uint32_t *)dk_dev_alloc(ctx, 4 * (m->n_chunks + 2));
    m->lbase = (uint64_t *)dk_dev_alloc(ctx, 8 * (m->n_chunks + 3));
    if (!m->keys || !m->cnts || !m->chunk_rows || !m->order || !m->lbase) { dk_matrix_release(ctx, m); return DK_ERR_NOMEM; }
    if (R) {
        if (keys) DK_CUDA(ctx, cudaMemcpyAsync(m->keys, keys, 8 * R, cudaMemcpyHostToDevice, st));
        DK_CUDA(ctx, cudaMemcpyAsync(m->cnts, cnts, 4 * R * S, cudaMemcpyHostToDevice, st));
    }
    unsigned g = (unsigned)((m->n_chunks + 255) / 256);
    k_identity_chunks<<<g ? g : 1, 256, 0, st>>>(m->order, m->lbase, m->chunk_rows, (uint32_t)m->n_chunks, R); DK_LAUNCHED(ctx);
    DK_CUDA(ctx, cudaStreamSynchronize(st));
    *out = m;
    return DK_OK;
}
extern "C" uint64_t dk_matrix_rows(const dk_matrix *m) { return m ? m->R : 0; }
extern "C" uint64_t dk_matrix_union(const dk_matrix *m) { return m ? m->U : 0; }
extern "C" int dk_matrix_samples(const dk_matrix *m) { return m ? m->S : 0; }
extern "C" void dk_matrix_free(dk_ctx *ctx, dk_matrix *m) { if (ctx) dk_matrix_release(ctx, m); }

// one thread per (row, 4-byte word): contiguous row-major destination, chunked source
__global__ void k_matrix_gather(const uint64_t *keys, const int32_t *cnts, const uint64_t *lbase, const uint32_t *order,
                                uint32_t n_chunks, int S, uint64_t row0, uint64_t nrows, uint64_t *okeys, int32_t *ocnts)
{
    uint64_t r = (uint64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    if (r >= nrows) return;
    uint64_t row = row0 + r;
    uint32_t lo = 0, hi = n_chunks;
    while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (lbase[mid] <= row) lo = mid; else hi = mid; }
    size_t ph = (size_t)order[lo] * DK_CH + (size_t)(row - lbase[lo]);
    int lane = threadIdx.x & 31;
    if (ocnts) for (int j = lane; j < S; j += 32) ocnts[r * (size_t)S + j] = cnts[ph * (size_t)S + j];
    if (okeys && lane == 0) okeys[r] = keys[ph];
}

extern "C" int dk_matrix_export_dev(dk_ctx *ctx, const dk_matrix *m, uint64_t row0, uint64_t nrows, uint64_t *keys_dev,
                                    int32_t *cnts_dev)
{
    if (!ctx || !m) return dk_fail(ctx, DK_ERR_ARG, "matrix_export: NULL argument");
    if (row0 + nrows > m->R) return dk_fail(ctx, DK_ERR_ARG, "matrix_export: rows [%llu,+%llu) beyond R=%llu",
                                            (unsigned long long)row0, (unsigned long long)nrows, (unsigned long long)m->R);
    if (!nrows) return DK_OK;
    if (keys_dev && !m->has_keys) return dk_fail(ctx, DK_ERR_ARG, "matrix_export: matrix has no keys");
    k_matrix_gather<<<(unsigned)((nrows + 7) / 8), 256, 0, ctx->stream>>>(m->keys, m->cnts, m->lbase, m->order, (uint32_t)m->n_chunks,
                                                                      m->S, row0, nrows, keys_dev, cnts_dev);
    DK_LAUNCHED(ctx);
    return DK_OK;
}
extern "C" int dk_matrix_export(dk_ctx *ctx, const dk_matrix *m, uint64_t row0, uint64_t nrows, uint64_t *keys, int32_t *cnts)
{
    if (!ctx || !m) return dk_fail(ctx, DK_ERR_ARG, "matrix_export: NULL argument");
    if (!nrows) return DK_OK;
    uint64_t *dk = keys ? (uint64_t *)dk_dev_alloc(ctx, 8 * nrows) : nullptr;
    int32_t *dc = cnts ? (int32_t *)dk_dev_alloc(ctx, 4 * nrows * m->S) : nullptr;
    if ((keys && !dk) || (cnts && !dc)) return DK_ERR_NOMEM;
    int rc = dk_matrix_export_dev(ctx, m, row0, nrows, dk, dc);
    if (rc == DK_OK) {
        if (keys) DK_CUDA(ctx, cudaMemcpyAsync(keys, dk, 8 * nrows, cudaMemcpyDeviceToHost, ctx->stream));
        if (cnts) DK_CUDA(ctx, cudaMemcpyAsync(cnts, dc, 4 * nrows * m->S, cudaMemcpyDeviceToHost, ctx->stream));
        DK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    dk_dev_release(ctx, dk); dk_dev_release(ctx, dc);
    return rc;
}

__global__ void k_matrix_checksum(const uint64_t *keys, const int32_t *cnts, const uint32_t *order, const uint64_t *lbase,
                                  uint32_t n_chunks, int S, int has_keys, unsigned long long *out)
{
    uint32_t oc = blockIdx.x;
    if (oc >= n_chunks) return;
    uint32_t rows = (uint32_t)(lbase[oc + 1] - lbase[oc]);
    size_t p0 = (size_t)order[oc] * DK_CH;
    unsigned long long x = 0, s = 0;
    for (uint32_t i = threadIdx.x; i < rows * (uint32_t)S; i += blockDim.x) s += (unsigned long long)(uint32_t)cnts[p0 * S + i];
    if (has_keys) for (uint32_t i = threadIdx.x; i < rows; i += blockDim.x) x ^= keys[p0 + i];
    for (int o = 16; o; o >>= 1) { x ^= __shfl_down_sync(0xffffffffu, x, o); s += __shfl_down_sync(0xffffffffu, s, o); }
    if ((threadIdx.x & 31) == 0) { atomicXor(&out[0], x); atomicAdd(&out[1], s); }
}
extern "C" int dk_matrix_checksum(dk_ctx *ctx, const dk_matrix *m, uint64_t *key_xor, uint64_t *cnt_sum)
{
    if (!ctx || !m) return dk_fail(ctx, DK_ERR_ARG, "matrix_checksum: NULL argument");
    unsigned long long *d = (unsigned long long *)dk_dev_alloc(ctx, 64);
    if (!d) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemsetAsync(d, 0, 64, ctx->stream));
    if (m->n_chunks) {
        k_matrix_checksum<<<(unsigned)m->n_chunks, 128, 0, ctx->stream>>>(m->keys, m->cnts, m->order, m->lbase, (uint32_t)m->n_chunks,
                                                                        m->S, m->has_keys ? 1 : 0, d);
        DK_LAUNCHED(ctx);
    }
    unsigned long long h[2];
    DK_CUDA(ctx, cudaMemcpyAsync(h, d, 16, cudaMemcpyDeviceToHost, ctx->stream));
    DK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (key_xor) *key_xor = h[0];
    if (cnt_sum) *cnt_sum = h[1];
    dk_dev_release(ctx, d);
    return DK_OK;
}

// ---- whole path -------------------------------------------------------------------
// "%f" round trip of the NFs: computeNF prints 6 decimals (computeNF.c:174) and
// TtestFilter reads them back with atof (TtestFilter.c:261), so the t-test never sees
// the full-precision factor. Formatting S numbers is pipeline glue, not compute.
static double nf_text_round(double v)
{
    char buf[64];
    snprintf(buf, sizeof buf, "%f", v);
    return atof(buf);
}

static int run_path_impl(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                         int device_inputs, int min_rec, int min_abund, int nf_mode, const int *idxA, int nA,
                         const int *idxB, int nB, double pmax, double fmin_, int nf_text_roundtrip, bool masked,
                         const uint64_t *mask_keys, uint64_t n_mask, int mask_on_device, dk_path_stats *stats,
                         uint64_t *rows_masked, dk_matrix **out_matrix, dk_ttest **out_test, dk_host_out *ho = nullptr)
{
    if (!ctx || !keys || !cnts || !n) return dk_fail(ctx, DK_ERR_ARG, "run_path: NULL argument");
    if (rows_masked) *rows_masked = 0;
    if (out_matrix) *out_matrix = nullptr;
    if (out_test) *out_test = nullptr;
    dk_matrix *m = nullptr;
    dk_ttest *t = nullptr;
    int rc = device_inputs ? dk_join_counts_dev(ctx, S, keys, cnts, n, min_rec, min_abund, 0, &m)
                           : join_counts_host(ctx, S, keys, cnts, n, min_rec, min_abund, 0, &m, ho);
    if (rc == DK_OK && device_inputs && ho) {                    // device-resident dumps: one copy after the join
        ho->written = m->R;
        rc = ho->written <= ho->cap ? dk_matrix_export(ctx, m, 0, m->R, ho->keys, ho->cnts)
                                    : dk_fail(ctx, DK_ERR_ARG, "matrix_out: %llu rows, the host buffers hold %llu",
                                              (unsigned long long)ho->written, (unsigned long long)ho->cap);
        if (rc != DK_OK) { dk_matrix_release(ctx, m); return rc; }
    }
    if (rc != DK_OK) return rc;
    dk_path_stats st;
    memset(&st, 0, sizeof st);
    st.U = m->U; st.R = m->R; st.ms_join = ctx->stage_ms[0];
    uint64_t info[4] = {0, 0, 0, 0};
    if (m->R == 0) { rc = DK_OK; goto finish; }
    rc = dk_compute_nf(ctx, m, nf_mode, 0, st.nf, info);
    if (rc != DK_OK) goto fail;
    st.R_pos = info[3]; st.ms_nf = ctx->stage_ms[1];
    if (masked) {                       // NF saw the raw matrix (Snakefile:366); everything after sees the masked one
        dk_matrix *mm = nullptr;
        rc = dk_matrix_mask(ctx, m, mask_keys, n_mask, mask_on_device, &mm, rows_masked);
        if (rc != DK_OK) goto fail;
        dk_matrix_release(ctx, m);
        m = mm;
        if (m->R == 0) { rc = DK_OK; goto finish; }
    }
    {
        double nf_used[DK_MAX_SAMPLES];
        for (int s = 0; s < S; s++) nf_used[s] = nf_text_roundtrip ? nf_text_round(st.nf[s]) : st.nf[s];
        rc = dk_ttest_rows(ctx, m, nf_used, idxA, nA, idxB, nB, &t);
        if (rc != DK_OK) goto fail;
        st.ms_ttest = ctx->stage_ms[2];
        cudaEventRecord(ctx->ev[0], ctx->stream);
        rc = dk_ttest_adjust(ctx, t, pmax);
        if (rc != DK_OK) goto fail;
        rc = dk_ttest_filter(ctx, t, pmax, fmin_, &st.D, st.ties);
        if (rc != DK_OK) goto fail;
        cudaEventRecord(ctx->ev[1], ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]);
        st.ms_bh = ms; ctx->stage_ms[3] = ms;
    }
finish:
    if (stats) *stats = st;
    if (out_test) *out_test = t; else dk_ttest_release(ctx, t);
    if (out_matrix) *out_matrix = m;
    else {
        if (out_test && t) { dk_ttest_release(ctx, t); *out_test = nullptr; }   // a test object needs its matrix
        dk_matrix_release(ctx, m);
    }
    return DK_OK;
fail:
    dk_ttest_release(ctx, t);
    dk_matrix_release(ctx, m);
    return rc;
}

extern "C" int dk_run_path(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                           int device_inputs, int min_rec, int min_abund, int nf_mode, const int *idxA, int nA,
                           const int *idxB, int nB, double pmax, double fmin_, int nf_text_roundtrip, dk_path_stats *stats,
                           dk_matrix **out_matrix, dk_ttest **out_test)
{
    return run_path_impl(ctx, S, keys, cnts, n, device_inputs, min_rec, min_abund, nf_mode, idxA, nA, idxB, nB, pmax, fmin_,
                         nf_text_roundtrip, false, nullptr, 0, 0, stats, nullptr, out_matrix, out_test);
}

extern "C" int dk_run_path_masked(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                                  int device_inputs, int min_rec, int min_abund, int nf_mode, const int *idxA, int nA,
                                  const int *idxB, int nB, double pmax, double fmin_, int nf_text_roundtrip,
                                  const uint64_t *mask_keys, uint64_t n_mask, int mask_on_device, dk_path_stats *stats,
                                  uint64_t *rows_masked, dk_matrix **out_matrix, dk_ttest **out_test)
{
    if (n_mask && !mask_keys) return dk_fail(ctx, DK_ERR_ARG, "run_path_masked: NULL mask");
    return run_path_impl(ctx, S, keys, cnts, n, device_inputs, min_rec, min_abund, nf_mode, idxA, nA, idxB, nB, pmax, fmin_,
                         nf_text_roundtrip, true, mask_keys, n_mask, mask_on_device, stats, rows_masked, out_matrix, out_test);
}

extern "C" int dk_run_path_matrix_out(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                                      int device_inputs, int min_rec, int min_abund, int nf_mode, const int *idxA, int nA,
                                      const int *idxB, int nB, double pmax, double fmin_, int nf_text_roundtrip,
                                      uint64_t *matrix_keys_host, int32_t *matrix_cnts_host, uint64_t matrix_cap_rows,
                                      uint64_t *matrix_rows, dk_path_stats *stats, dk_matrix **out_matrix, dk_ttest **out_test)
{
    if (!matrix_keys_host || !matrix_cnts_host) return dk_fail(ctx, DK_ERR_ARG, "run_path_matrix_out: NULL host buffer");
    dk_host_out ho;
    ho.keys = matrix_keys_host; ho.cnts = matrix_cnts_host; ho.cap = matrix_cap_rows;
    const int rc = run_path_impl(ctx, S, keys, cnts, n, device_inputs, min_rec, min_abund, nf_mode, idxA, nA, idxB, nB, pmax, fmin_,
                                 nf_text_roundtrip, false, nullptr, 0, 0, stats, nullptr, out_matrix, out_test, &ho);
    if (matrix_rows) *matrix_rows = ho.written;
    return rc;
}
