// dk_common.cuh — context, device-memory cache, matrix object, error plumbing.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <map>
#include <string>
#include <vector>
#include "../../include/dekupl_b200.h"

#define DK_CH 64                 // rows per output chunk of the merge (see DESIGN.md "matrix layout")
#define DK_SENTINEL 0xFFFFFFFFFFFFFFFFull

struct dk_ctx {
    int device = 0;
    int sm_count = 148;
    size_t smem_optin = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    std::string err;
    uint64_t launches = 0;
    double stage_ms[8] = {0};
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    // size-bucketed cache of device blocks: steady-state jobs never call cudaMalloc
    std::multimap<size_t, void *> free_blocks;
    std::map<void *, size_t> live;
    size_t cached_bytes = 0;
    uint64_t chunk_hint = 0;      // matrix chunks the previous join of THIS context needed (sizes the next one)
    // multi-GPU (dk_comm.cu): an NCCL communicator owned by the context; world == 1 without one
    void *comm = nullptr;
    int rank = 0, world = 1;
};

int dk_fail(dk_ctx *ctx, int code, const char *fmt, ...);
void *dk_dev_alloc(dk_ctx *ctx, size_t bytes);           // nullptr on failure (error recorded)
void dk_dev_release(dk_ctx *ctx, void *p);               // back to the cache
void dk_dev_trim(dk_ctx *ctx);                           // cudaFree everything cached

#define DK_CUDA(ctx, call)                                                                  \
    do {                                                                                    \
        cudaError_t e__ = (call);                                                           \
        if (e__ != cudaSuccess)                                                             \
            return dk_fail(ctx, DK_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call,     \
                           cudaGetErrorString(e__));                                        \
    } while (0)

#define DK_LAUNCHED(ctx)                                                                    \
    do {                                                                                    \
        (ctx)->launches++;                                                                  \
        cudaError_t e__ = cudaGetLastError();                                               \
        if (e__ != cudaSuccess)                                                             \
            return dk_fail(ctx, DK_ERR_CUDA, "%s:%d launch: %s", __FILE__, __LINE__,        \
                           cudaGetErrorString(e__));                                        \
    } while (0)

#define DK_TRY(expr)                                                                        \
    do {                                                                                    \
        int rc__ = (expr);                                                                  \
        if (rc__ != DK_OK) return rc__;                                                     \
    } while (0)

// Filtered count matrix, device resident. Physical storage is CHUNKED: the merge
// allocates DK_CH-row chunks from a bump counter as streams produce rows, so
// physical chunk order is arbitrary; `order` lists chunks in reference (sorted
// k-mer) order and `lbase[i]` is the logical index of the first row of order[i].
struct dk_matrix {
    int S = 0;
    uint64_t R = 0, U = 0;
    uint64_t n_chunks = 0;        // chunks in use
    uint64_t cap_chunks = 0;      // chunks allocated
    uint64_t *keys = nullptr;     // [cap_chunks*DK_CH]
    int32_t *cnts = nullptr;      // [cap_chunks*DK_CH*S] row-major
    uint32_t *chunk_rows = nullptr;  // [cap_chunks] rows used in physical chunk
    uint32_t *order = nullptr;    // [n_chunks] physical chunk id at ordered position i
    uint64_t *lbase = nullptr;    // [n_chunks+1] logical row base of ordered chunk i
    bool has_keys = true;
};

struct dk_ttest {
    const dk_matrix *m = nullptr;
    uint64_t R = 0;
    int S = 0, nA = 0, nB = 0;
    double *nf_dev = nullptr;     // [S]
    int *idx_dev = nullptr;       // [nA+nB]
    double *p_raw = nullptr;      // [R] logical order
    double *log2fc = nullptr;     // [R]
    mutable bool fc_lazy = false; // log2fc not written by the row kernel: dk_ttest_need_log2fc fills it before it is read
    double *p_adj = nullptr;      // [R] NaN where not computed
    // candidates (p_raw <= p_cut)
    uint64_t n_cand = 0;
    double p_cut = 0;
    double *cand_p = nullptr;     // [n_cand]
    uint32_t *cand_row = nullptr; // [n_cand] logical row (R < 2^32 per GPU)
    uint64_t cand_cap = 0;
    // survivors
    uint64_t D = 0;
    uint32_t *surv_row = nullptr; // [D] ascending logical rows
    uint64_t *tie_counts = nullptr; // [4] device
};

static inline size_t dk_round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- internal stage entry points (defined in the .cu files) --------------------
int dk_join_impl(dk_ctx *ctx, int S, const uint64_t *const *keys_dev, const int32_t *const *cnts_dev,
                 const uint64_t *n, int min_rec, int min_abund, int emit_stale_tail, dk_matrix **out);
int dk_scan_excl_u64(dk_ctx *ctx, const uint64_t *in, uint64_t *out, size_t n);   // out[n] = total
int dk_scan_excl_u32(dk_ctx *ctx, const uint32_t *in, uint32_t *out, size_t n);
// log2FC of the listed logical rows (rows == NULL: all rows) when the row kernel left it out; no-op otherwise
int dk_ttest_need_log2fc(dk_ctx *ctx, const dk_ttest *t, const uint32_t *rows, uint64_t n);
int dk_ttest_candidates_known(dk_ctx *ctx, dk_ttest *t, double cut, double p_cut, uint64_t n_known, uint64_t *n_cand,
                              const double **cand_p_dev);
#define DK_BH_SMALL_LIST (1u << 20)   // candidate lists up to this size cost less to process than another pass over p_raw
int dk_sort_keys_u64_scratch(dk_ctx *ctx, uint64_t *keys, uint64_t *keys_tmp, size_t n, uint32_t *cnt /* 256*ceil(n/4096) */);
int dk_sort_pairs_u64(dk_ctx *ctx, uint64_t *keys, uint32_t *vals, uint64_t *keys_tmp, uint32_t *vals_tmp,
                      size_t n, uint64_t **keys_sorted, uint32_t **vals_sorted);
