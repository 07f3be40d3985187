// dk_comm.cu — the multi-GPU path UNDER the C ABI: an NCCL communicator owned by the context and
// dk_run_path_sharded, the whole path on this rank's k-mer-range shard with the global pieces done by NCCL here.
// One process per GPU. NCCL is resolved with dlopen at dk_comm_init (libnccl.so.2: the copy the process already holds,
// e.g. torch's, else the system one), so the library has no link-time dependency on it and single-GPU callers — the
// five CLIs — never touch it.
#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>
#include "dk_common.cuh"

void dk_matrix_release(dk_ctx *ctx, dk_matrix *m);
void dk_ttest_release(dk_ctx *ctx, dk_ttest *t);

// the part of nccl.h (2.x ABI) this file needs
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { DK_NCCL_INT64 = 4, DK_NCCL_UINT64 = 5, DK_NCCL_FLOAT64 = 8, DK_NCCL_SUM = 0 };
struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Broadcast)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
static NcclApi g_nccl;

static int nccl_load(dk_ctx *ctx)
{
    if (g_nccl.lib) return DK_OK;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return dk_fail(ctx, DK_ERR_UNSUPPORTED, "dk_comm: libnccl.so.2 not found (%s)", dlerror());
    NcclApi a;
    a.lib = h;
    *(void **)&a.GetUniqueId = dlsym(h, "ncclGetUniqueId");
    *(void **)&a.CommInitRank = dlsym(h, "ncclCommInitRank");
    *(void **)&a.CommDestroy = dlsym(h, "ncclCommDestroy");
    *(void **)&a.AllReduce = dlsym(h, "ncclAllReduce");
    *(void **)&a.Broadcast = dlsym(h, "ncclBroadcast");
    *(void **)&a.AllGather = dlsym(h, "ncclAllGather");
    *(void **)&a.GetErrorString = dlsym(h, "ncclGetErrorString");
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllReduce || !a.Broadcast || !a.AllGather || !a.GetErrorString)
        return dk_fail(ctx, DK_ERR_UNSUPPORTED, "dk_comm: libnccl lacks a required symbol");
    g_nccl = a;
    return DK_OK;
}
#define DK_NCCL(ctx, call)                                                                          \
    do {                                                                                            \
        int r__ = (call);                                                                           \
        if (r__ != 0) return dk_fail(ctx, DK_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString(r__)); \
    } while (0)

extern "C" int dk_comm_unique_id(dk_ctx *ctx, unsigned char *id)
{
    if (!ctx || !id) return dk_fail(ctx, DK_ERR_ARG, "comm_unique_id: NULL argument");
    DK_TRY(nccl_load(ctx));
    ncclUniqueId u;
    DK_NCCL(ctx, g_nccl.GetUniqueId(&u));
    memcpy(id, u.internal, DK_COMM_ID_BYTES);
    return DK_OK;
}
extern "C" int dk_comm_init(dk_ctx *ctx, const unsigned char *id, int rank, int world)
{
    if (!ctx || !id) return dk_fail(ctx, DK_ERR_ARG, "comm_init: NULL argument");
    if (world < 1 || rank < 0 || rank >= world) return dk_fail(ctx, DK_ERR_ARG, "comm_init: rank %d of %d", rank, world);
    if (ctx->comm) return dk_fail(ctx, DK_ERR_ARG, "comm_init: the context already owns a communicator");
    DK_TRY(nccl_load(ctx));
    DK_CUDA(ctx, cudaSetDevice(ctx->device));
    ncclUniqueId u;
    memcpy(u.internal, id, DK_COMM_ID_BYTES);
    ncclComm_t c = nullptr;
    DK_NCCL(ctx, g_nccl.CommInitRank(&c, world, u, rank));
    ctx->comm = c; ctx->rank = rank; ctx->world = world;
    return DK_OK;
}
extern "C" void dk_comm_free(dk_ctx *ctx)
{
    if (!ctx || !ctx->comm) return;
    cudaStreamSynchronize(ctx->stream);
    g_nccl.CommDestroy((ncclComm_t)ctx->comm);
    ctx->comm = nullptr; ctx->rank = 0; ctx->world = 1;
}
extern "C" int dk_comm_rank(const dk_ctx *ctx) { return ctx ? ctx->rank : 0; }
extern "C" int dk_comm_world(const dk_ctx *ctx) { return ctx ? ctx->world : 1; }

// ---- small collectives on the context's stream (values travel through a scratch block of the context's cache) ---------
static int comm_allreduce_u64(dk_ctx *ctx, uint64_t *v, int n)         // in place, sum, host values
{
    if (!ctx->comm || ctx->world == 1) return DK_OK;
    uint64_t *d = (uint64_t *)dk_dev_alloc(ctx, 8 * (size_t)n);
    if (!d) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemcpyAsync(d, v, 8 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    DK_NCCL(ctx, g_nccl.AllReduce(d, d, (size_t)n, DK_NCCL_UINT64, DK_NCCL_SUM, (ncclComm_t)ctx->comm, ctx->stream));
    DK_CUDA(ctx, cudaMemcpyAsync(v, d, 8 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    DK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    dk_dev_release(ctx, d);
    return DK_OK;
}
static int comm_bcast_f64(dk_ctx *ctx, double *v, int n, int root)
{
    if (!ctx->comm || ctx->world == 1) return DK_OK;
    double *d = (double *)dk_dev_alloc(ctx, 8 * (size_t)n);
    if (!d) return DK_ERR_NOMEM;
    DK_CUDA(ctx, cudaMemcpyAsync(d, v, 8 * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
    DK_NCCL(ctx, g_nccl.Broadcast(d, d, (size_t)n, DK_NCCL_FLOAT64, root, (ncclComm_t)ctx->comm, ctx->stream));
    DK_CUDA(ctx, cudaMemcpyAsync(v, d, 8 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
    DK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    dk_dev_release(ctx, d);
    return DK_OK;
}
static double nf_text_round_c(double v)
{
    char buf[64];
    snprintf(buf, sizeof buf, "%f", v);                                  // computeNF.c:174 -> TtestFilter.c:261
    return atof(buf);
}

// The whole path on this rank's shard (rank order == k-mer order: dk_key_splitters). No collective on the per-row work.
//   R, U           : one all-reduce
//   NF, reference  : computeNF consumes a PREFIX of the sorted table (computeNF.c:138-168): rank 0 answers if its early
//                    stop fired inside its shard, else every rank fails (use DK_NF_FULL); one broadcast
//   NF, full       : exact global element [size/2] per sample by radix select, one all-reduce of S*256 bins per round
//   BH             : candidate narrowing over GLOBAL counts (one scalar all-reduce per counting pass, dk_bh_next_cut), then an
//                    all-gather of the few candidates that remain; every rank ranks the same list (TtestFilter.c:456-476)
//   D, ties        : one all-reduce
extern "C" int dk_run_path_sharded(dk_ctx *ctx, int S, const uint64_t *const *keys, const int32_t *const *cnts, const uint64_t *n,
                                   int device_inputs, int min_rec, int min_abund, int nf_mode, const int *idxA, int nA,
                                   const int *idxB, int nB, double pmax, double fmin_, int nf_text_roundtrip,
                                   dk_path_stats *stats, uint64_t *local_URD, dk_matrix **out_matrix, dk_ttest **out_test)
{
    if (!ctx || !keys || !cnts || !n || !stats) return dk_fail(ctx, DK_ERR_ARG, "run_path_sharded: NULL argument");
    if (out_matrix) *out_matrix = nullptr;
    if (out_test) *out_test = nullptr;
    const int world = ctx->comm ? ctx->world : 1, rank = ctx->comm ? ctx->rank : 0;
    dk_matrix *m = nullptr;
    dk_ttest *t = nullptr;
    dk_path_stats st;
    memset(&st, 0, sizeof st);
    int rc = device_inputs ? dk_join_counts_dev(ctx, S, keys, cnts, n, min_rec, min_abund, 0, &m)
                           : dk_join_counts(ctx, S, keys, cnts, n, min_rec, min_abund, 0, &m);
    if (rc != DK_OK) return rc;
    st.ms_join = ctx->stage_ms[0];
    const uint64_t R_loc = m->R, U_loc = m->U;
    uint64_t ru[2] = {R_loc, U_loc};
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    double *gath = nullptr, *allp = nullptr;
    uint64_t D_loc = 0, ties[3] = {0, 0, 0};
#define SH_FAIL(x) do { rc = (x); if (rc != DK_OK) goto fail; } while (0)
    SH_FAIL(comm_allreduce_u64(ctx, ru, 2));
    st.R = ru[0]; st.U = ru[1];
    if (st.R == 0) goto finish;
    {   // ---- NF
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0, ctx->stream);
        double msg[DK_MAX_SAMPLES + 1];
        memset(msg, 0, sizeof msg);
        if (nf_mode == DK_NF_REFERENCE) {
            if (rank == 0) {
                uint64_t info[4] = {0, 0, 0, 0};
                int r0 = R_loc ? dk_compute_nf(ctx, m, DK_NF_REFERENCE, 0, msg + 1, info) : DK_ERR_UNSUPPORTED;
                msg[0] = (r0 == DK_OK && (world == 1 || (info[0] < R_loc && !info[2]))) ? 1.0 : 0.0;
                st.R_pos = info[3];
            }
            SH_FAIL(comm_bcast_f64(ctx, msg, S + 1, 0));
            if (msg[0] != 1.0)
                SH_FAIL(dk_fail(ctx, DK_ERR_UNSUPPORTED, "run_path_sharded: computeNF's early stop did not fire inside rank 0's shard; "
                                "use DK_NF_FULL or fewer shards"));
            memcpy(st.nf, msg + 1, sizeof(double) * S);
        } else {
            dk_nf_state *ns = nullptr;
            uint64_t npos = 0;
            SH_FAIL(dk_nf_begin(ctx, m, &ns, &npos));
            uint64_t tot = npos;
            rc = comm_allreduce_u64(ctx, &tot, 1);
            uint64_t *hist = rc == DK_OK ? (uint64_t *)dk_dev_alloc(ctx, 8 * (size_t)S * 256) : nullptr;
            if (rc == DK_OK && !hist) rc = DK_ERR_NOMEM;
            if (rc == DK_OK && tot == 0) rc = dk_fail(ctx, DK_ERR_UNSUPPORTED, "compute_nf: no row has all counts > 0");
            if (rc == DK_OK) rc = dk_nf_set_rank(ctx, ns, tot / 2);               // element [size/2], computeNF.c:43
            for (int round = 0; round < 8 && rc == DK_OK; round++) {
                rc = dk_nf_hist_dev(ctx, ns, hist);
                if (rc == DK_OK && world > 1 &&
                    g_nccl.AllReduce(hist, hist, (size_t)S * 256, DK_NCCL_UINT64, DK_NCCL_SUM, (ncclComm_t)ctx->comm, ctx->stream) != 0)
                    rc = dk_fail(ctx, DK_ERR_CUDA, "run_path_sharded: histogram all-reduce failed");
                if (rc == DK_OK) rc = dk_nf_advance_dev(ctx, ns, hist);
            }
            if (rc == DK_OK) rc = dk_nf_result(ctx, ns, st.nf);
            st.R_pos = tot;
            dk_dev_release(ctx, hist);
            dk_nf_end(ctx, ns);
            if (rc != DK_OK) goto fail;
        }
        cudaEventRecord(e1, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        st.ms_nf = ms; ctx->stage_ms[1] = ms;
    }
    {   // ---- rows, BH, filter
        double nf_used[DK_MAX_SAMPLES];
        for (int s = 0; s < S; s++) nf_used[s] = nf_text_roundtrip ? nf_text_round_c(st.nf[s]) : st.nf[s];
        if (R_loc) { SH_FAIL(dk_ttest_rows(ctx, m, nf_used, idxA, nA, idxB, nB, &t)); st.ms_ttest = ctx->stage_ms[2]; }
        cudaEventRecord(e0, ctx->stream);
        double cut = pmax;
        uint64_t n_known = ~0ull;
        if (pmax < 1.0) {
            for (int it = 0; it < 4; it++) {
                uint64_t c = 0;
                if (t) SH_FAIL(dk_ttest_count_le(ctx, t, cut, &c));
                n_known = c;                                               // this rank's count at `cut`
                SH_FAIL(comm_allreduce_u64(ctx, &c, 1));
                const double next = dk_bh_next_cut(pmax, c, st.R);
                if (c <= (uint64_t)DK_BH_SMALL_LIST * (uint64_t)world || !(next < cut * 0.75)) break;
                cut = next; n_known = ~0ull;
            }
        }
        uint64_t nc = 0;
        const double *cp = nullptr;
        if (t) SH_FAIL(dk_ttest_candidates_known(ctx, t, cut, pmax, n_known, &nc, &cp));
        if (world == 1) {
            if (t) SH_FAIL(dk_ttest_adjust_global(ctx, t, cp, nc, st.R));
        } else {
            // all-gather of the candidate lists: counts first, then the values padded to the longest list
            std::vector<uint64_t> cnt(world, 0);
            cnt[rank] = nc;
            SH_FAIL(comm_allreduce_u64(ctx, cnt.data(), world));
            uint64_t mx = 0, n_all = 0;
            for (int g = 0; g < world; g++) { mx = cnt[g] > mx ? cnt[g] : mx; n_all += cnt[g]; }
            if (n_all) {
                gath = (double *)dk_dev_alloc(ctx, 8 * mx * (size_t)(world + 1));
                allp = (double *)dk_dev_alloc(ctx, 8 * n_all);
                if (!gath || !allp) SH_FAIL(DK_ERR_NOMEM);
                double *mine = gath + mx * (size_t)world;
                if (cudaMemsetAsync(mine, 0, 8 * mx, ctx->stream) != cudaSuccess ||
                    (nc && cudaMemcpyAsync(mine, cp, 8 * nc, cudaMemcpyDeviceToDevice, ctx->stream) != cudaSuccess))
                    SH_FAIL(dk_fail(ctx, DK_ERR_CUDA, "run_path_sharded: staging the candidates failed"));
                if (g_nccl.AllGather(mine, gath, (size_t)mx, DK_NCCL_FLOAT64, (ncclComm_t)ctx->comm, ctx->stream) != 0)
                    SH_FAIL(dk_fail(ctx, DK_ERR_CUDA, "run_path_sharded: candidate all-gather failed"));
                uint64_t off = 0;
                for (int g = 0; g < world; g++) {
                    if (cnt[g] && cudaMemcpyAsync(allp + off, gath + mx * (size_t)g, 8 * cnt[g], cudaMemcpyDeviceToDevice, ctx->stream) != cudaSuccess)
                        SH_FAIL(dk_fail(ctx, DK_ERR_CUDA, "run_path_sharded: compacting the candidates failed"));
                    off += cnt[g];
                }
            }
            if (t) SH_FAIL(dk_ttest_adjust_global(ctx, t, allp, n_all, st.R));
        }
        if (t) SH_FAIL(dk_ttest_filter(ctx, t, pmax, fmin_, &D_loc, ties));
        uint64_t dd[4] = {D_loc, ties[0], ties[1], ties[2]};
        SH_FAIL(comm_allreduce_u64(ctx, dd, 4));
        st.D = dd[0]; st.ties[0] = dd[1]; st.ties[1] = dd[2]; st.ties[2] = dd[3];
        cudaEventRecord(e1, ctx->stream);
        cudaStreamSynchronize(ctx->stream);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        st.ms_bh = ms; ctx->stage_ms[3] = ms;
    }
finish:
    if (e0) { cudaEventDestroy(e0); cudaEventDestroy(e1); }
    dk_dev_release(ctx, gath); dk_dev_release(ctx, allp);
    *stats = st;
    if (local_URD) { local_URD[0] = U_loc; local_URD[1] = R_loc; local_URD[2] = D_loc; }
    if (out_test && out_matrix) { *out_test = t; *out_matrix = m; }
    else { dk_ttest_release(ctx, t); if (out_matrix) *out_matrix = m; else dk_matrix_release(ctx, m); }
    return DK_OK;
fail:
    if (e0) { cudaEventDestroy(e0); cudaEventDestroy(e1); }
    dk_dev_release(ctx, gath); dk_dev_release(ctx, allp);
    dk_ttest_release(ctx, t);
    dk_matrix_release(ctx, m);
    return rc;
#undef SH_FAIL
}
