// dk_ctx.cu — context lifetime, error text, cached device allocator, scans.
#include <stdarg.h>
#include "dk_common.cuh"

static thread_local char g_err[512] = "";

int dk_fail(dk_ctx *ctx, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    snprintf(g_err, sizeof g_err, "%s", buf);
    if (ctx) ctx->err = buf;
    return code;
}

extern "C" const char *dk_last_error(const dk_ctx *ctx) { return ctx ? ctx->err.c_str() : g_err; }
extern "C" const char *dk_version(void) { return "dekupl_b200 0.1 (sm_100a)"; }

extern "C" int dk_init(dk_ctx **out, int device)
{
    if (!out) return dk_fail(nullptr, DK_ERR_ARG, "dk_init: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return dk_fail(nullptr, DK_ERR_NODEVICE, "dk_init: no CUDA device (%s); this library has no CPU fallback",
                       e == cudaSuccess ? "count=0" : cudaGetErrorString(e));
    if (device < 0 || device >= n) return dk_fail(nullptr, DK_ERR_ARG, "dk_init: device %d out of range [0,%d)", device, n);
    dk_ctx *ctx = new dk_ctx();
    ctx->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess) {
        delete ctx;
        return dk_fail(nullptr, DK_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    if (prop.major < 10) {
        delete ctx;
        return dk_fail(nullptr, DK_ERR_NODEVICE, "dk_init: device %d is sm_%d%d; this library is sm_100a only", device,
                       prop.major, prop.minor);
    }
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    ctx->own_stream = true;
    for (int i = 0; i < 4; i++) cudaEventCreate(&ctx->ev[i]);
    *out = ctx;
    return DK_OK;
}

extern "C" void dk_free(dk_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    dk_comm_free(ctx);
    cudaStreamSynchronize(ctx->stream);
    dk_dev_trim(ctx);
    for (auto &kv : ctx->live) cudaFree(kv.first);
    for (int i = 0; i < 4; i++) if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
    if (ctx->own_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" int dk_set_stream(dk_ctx *ctx, void *s)
{
    if (!ctx) return DK_ERR_ARG;
    if (ctx->own_stream) { cudaStreamSynchronize(ctx->stream); cudaStreamDestroy(ctx->stream); ctx->own_stream = false; }
    ctx->stream = (cudaStream_t)s;
    return DK_OK;
}
extern "C" int dk_sync(dk_ctx *ctx)
{
    DK_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return DK_OK;
}
extern "C" uint64_t dk_launch_count(const dk_ctx *ctx) { return ctx ? ctx->launches : 0; }
extern "C" double dk_stage_ms(const dk_ctx *ctx, int which) { return (ctx && which >= 0 && which < 8) ? ctx->stage_ms[which] : 0.0; }

// ---- cached allocator -----------------------------------------------------------
void *dk_dev_alloc(dk_ctx *ctx, size_t bytes)
{
    if (bytes == 0) bytes = 256;
    bytes = dk_round_up(bytes, 512);
    auto it = ctx->free_blocks.lower_bound(bytes);
    if (it != ctx->free_blocks.end() && it->first <= bytes + bytes / 4 + 4096) {
        void *p = it->second;
        ctx->live[p] = it->first;
        ctx->cached_bytes -= it->first;
        ctx->free_blocks.erase(it);
        return p;
    }
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        dk_dev_trim(ctx);               // drop the cache and retry once
        e = cudaMalloc(&p, bytes);
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        dk_fail(ctx, DK_ERR_NOMEM, "cudaMalloc(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
        return nullptr;
    }
    ctx->live[p] = bytes;
    return p;
}
void dk_dev_release(dk_ctx *ctx, void *p)
{
    if (!p) return;
    auto it = ctx->live.find(p);
    if (it == ctx->live.end()) return;
    ctx->free_blocks.insert({it->second, p});
    ctx->cached_bytes += it->second;
    ctx->live.erase(it);
}
void dk_dev_trim(dk_ctx *ctx)
{
    cudaStreamSynchronize(ctx->stream);
    for (auto &kv : ctx->free_blocks) cudaFree(kv.second);
    ctx->free_blocks.clear();
    ctx->cached_bytes = 0;
}

// ---- exclusive scans (reduce / scan block sums / apply) --------------------------
// Used only on small bookkeeping arrays (partition tables, chunk tables): the
// per-row data path never needs an ordered scan.
#define SCAN_T 256
#define SCAN_V 8
template <typename T>
__global__ void k_scan_reduce(const T *in, T *bsum, size_t n)
{
    __shared__ T sh[SCAN_T / 32];
    size_t base = (size_t)blockIdx.x * SCAN_T * SCAN_V;
    T v = 0;
    for (int i = 0; i < SCAN_V; i++) {
        size_t k = base + (size_t)i * SCAN_T + threadIdx.x;
        if (k < n) v += in[k];
    }
    for (int o = 16; o; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        T t = 0;
        for (int i = 0; i < SCAN_T / 32; i++) t += sh[i];
        bsum[blockIdx.x] = t;
    }
}
template <typename T>
__global__ void k_scan_bsums(T *bsum, size_t nb, T *total)
{
    // single block, sequential over tiles of SCAN_T
    __shared__ T sh[SCAN_T];
    __shared__ T carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (size_t b0 = 0; b0 < nb; b0 += SCAN_T) {
        size_t k = b0 + threadIdx.x;
        T v = k < nb ? bsum[k] : 0;
        sh[threadIdx.x] = v;
        __syncthreads();
        for (int o = 1; o < SCAN_T; o <<= 1) {
            T t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
            __syncthreads();
            sh[threadIdx.x] += t;
            __syncthreads();
        }
        T incl = sh[threadIdx.x];
        if (k < nb) bsum[k] = carry + incl - v;
        __syncthreads();
        if (threadIdx.x == SCAN_T - 1) carry += incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}
template <typename T>
__global__ void k_scan_apply(const T *in, T *out, const T *bsum, size_t n)
{
    // each thread owns SCAN_V consecutive items; block-level scan of thread sums
    __shared__ T sh[SCAN_T];
    size_t base = (size_t)blockIdx.x * SCAN_T * SCAN_V + (size_t)threadIdx.x * SCAN_V;
    T loc[SCAN_V];
    T s = 0;
    for (int i = 0; i < SCAN_V; i++) {
        loc[i] = (base + i < n) ? in[base + i] : 0;
        s += loc[i];
    }
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 1; o < SCAN_T; o <<= 1) {
        T t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0;
        __syncthreads();
        sh[threadIdx.x] += t;
        __syncthreads();
    }
    T run = bsum[blockIdx.x] + sh[threadIdx.x] - s;
    for (int i = 0; i < SCAN_V; i++) {
        if (base + i < n) out[base + i] = run;
        run += loc[i];
    }
}
// NOTE: k_scan_reduce sums items strided by SCAN_T while k_scan_apply uses a
// blocked layout; both cover the same SCAN_T*SCAN_V window, so block sums agree.

template <typename T>
static int scan_excl(dk_ctx *ctx, const T *in, T *out, size_t n)
{
    if (n == 0) { DK_CUDA(ctx, cudaMemsetAsync(out, 0, sizeof(T), ctx->stream)); return DK_OK; }
    size_t nb = (n + SCAN_T * SCAN_V - 1) / (SCAN_T * SCAN_V);
    T *bsum = (T *)dk_dev_alloc(ctx, sizeof(T) * nb);
    if (!bsum) return DK_ERR_NOMEM;
    k_scan_reduce<T><<<(unsigned)nb, SCAN_T, 0, ctx->stream>>>(in, bsum, n); DK_LAUNCHED(ctx);
    k_scan_bsums<T><<<1, SCAN_T, 0, ctx->stream>>>(bsum, nb, out + n); DK_LAUNCHED(ctx);
    k_scan_apply<T><<<(unsigned)nb, SCAN_T, 0, ctx->stream>>>(in, out, bsum, n); DK_LAUNCHED(ctx);
    dk_dev_release(ctx, bsum);   // stream-ordered reuse: later users enqueue on the same stream
    return DK_OK;
}
int dk_scan_excl_u64(dk_ctx *ctx, const uint64_t *in, uint64_t *out, size_t n) { return scan_excl<uint64_t>(ctx, in, out, n); }
int dk_scan_excl_u32(dk_ctx *ctx, const uint32_t *in, uint32_t *out, size_t n) { return scan_excl<uint32_t>(ctx, in, out, n); }
