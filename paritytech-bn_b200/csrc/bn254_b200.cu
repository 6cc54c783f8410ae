// C-ABI implementation (include/bn254_b200.h): contexts, device buffers, kernel launches.
// Every entry point launches CUDA kernels; there is no CPU arithmetic path in this library.
#include <cuda_runtime.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/bn254_b200.h"
#include "bn254_kernels.cuh"

using namespace bn;

static_assert(sizeof(Fq) == 32 && sizeof(Fq2) == 64 && sizeof(Fq12) == 384, "repr(C) sizes");
static_assert(sizeof(G1) == 96 && sizeof(G2) == 192 && sizeof(EllCoeffs) == 192, "repr(C) sizes");
static_assert(sizeof(AffG2) == 128 && sizeof(AffG1) == 64, "repr(C) sizes");

static thread_local std::string g_err;
static int fail(int code, const char* what, cudaError_t e = cudaSuccess) {
    g_err = what;
    if (e != cudaSuccess) { g_err += ": "; g_err += cudaGetErrorString(e); }
    return code;
}
#define CU(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return fail(BN254_ERR_CUDA, #call, _e); } while (0)

struct bn254_ctx {
    int device;
    cudaStream_t stream;
    std::atomic<uint64_t> launches;
    void* io;          // grow-only device arena for the host-pointer entry points (no cudaMalloc per call)
    size_t io_bytes;
    void* ws;          // grow-only device workspace (line-coefficient tables of the split Miller path)
    size_t ws_bytes;
    int split_miller;  // 1: three-kernel Miller path (default), 0: fused single kernel (BN254_FUSED_MILLER=1)
    int split_fexp;    // 1: easy + hard final-exponentiation kernels (BN254_SPLIT_FEXP=1), 0: one kernel
    cudaStream_t stream2;   // second stream of the pipelined host-buffer pairing path (the other half of the batch)
    cudaEvent_t ev_fork, ev_join;
    void* ws_ovr;      // when set, ctx_workspace hands out this region instead of the arena (one region per half batch)
    size_t ws_ovr_bytes;
    int pipeline;      // 1 (default): bn254_pairing_many runs large batches as two half batches on two streams (BN254_PIPELINE=0 disables)
    int big_blocks;    // 1 (default): 512-thread blocks for batches of more than one wave (BN254_BIG_BLOCKS=0: always at most 448)
    unsigned force_block;   // non-zero: block size of the phase-synchronised kernels regardless of the batch (pipelined halves)
    int sms;           // multiprocessor count (block size of the phase-synchronised kernels: one block per SM)
    int lines_smem;    // 1 (default): G2 line tables on the slot file (k_g2_lines_sync; BN254_LINES_SMEM=0: local-memory kernel)
    int miller_smem;   // 1 (default): Miller loop from the table on the slot file (k_miller_table_sync; BN254_MILLER_SMEM=0: local memory)
    int fexp_smem;     // hard part of the final exponentiation (BN254_FEXP_SMEM): 2 (default) = shared-memory slot file, one
                       // phase-synchronised block per SM (k_final_exp_hard_sync); 1 = slot file, 64-thread blocks; 0 = local memory
    bool zerocopy_out; // pinned output buffers are written by the last kernel directly (BN254_ZEROCOPY_OUT=0 disables)
    bool subgroup_ref; // true: G2 order check with the reference's [r-1]P + P ladder (BN254_SUBGROUP_REF=1); false: psi test
};
struct bn254_g2precomp {
    int device;
    size_t n;
    uint4* tab;     // [(c*12 + w) * n + j]
    AffG2* q_aff;   // [n]
};

// RAII device buffer on the context's device
struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    int alloc(size_t bytes) {
        if (bytes == 0) bytes = 16;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) { p = nullptr; return fail(BN254_ERR_NOMEM, "cudaMalloc", e); }
        return 0;
    }
    template <class T> T* as() { return reinterpret_cast<T*>(p); }
};
#define TRY(x) do { int _r = (x); if (_r != 0) return _r; } while (0)

// Block size of the one-block-per-SM kernels: the batch spread over all SMs in whole warps, at most BN_FEXP_BLOCK threads
// (2^16 pairings on 148 SMs: 448 threads, 147 blocks; 2^15: 224 threads; small batches: one warp per block).
static inline unsigned sync_block(const struct bn254_ctx* ctx, size_t n);
static inline unsigned grid_for(size_t n) { return (unsigned)((n + BN_BLOCK - 1) / BN_BLOCK); }
#define LAUNCH(ctx, kernel, n, ...)                                                        \
    do {                                                                                   \
        if ((n) > 0) {                                                                     \
            kernel<<<grid_for(n), BN_BLOCK, 0, (ctx)->stream>>>(__VA_ARGS__);              \
            (ctx)->launches.fetch_add(1);                                                  \
            CU(cudaGetLastError());                                                        \
        }                                                                                  \
    } while (0)

static inline unsigned sync_block(const bn254_ctx* ctx, size_t n) {
    if (ctx->force_block) return ctx->force_block;
    // more than one wave of 448-thread blocks: 512-thread blocks on 6 slots per thread -- four warps on every sub-partition
    // instead of 4 + 4 + 3 + 3 (DESIGN F14)
    // -- when that saves waves: a 512-thread wave takes 1.10 x a 448-thread one (measured at 2^19: 23.5 against 21.4 ms for
    // the whole path), so 2^19 (7 against 8 waves) and 2^20 (14 against 16) gain 4 %, 2^17 and 2^18 (same wave count) would lose
    if (ctx->big_blocks == 2 && n > BN_SYNC_MAX_BLOCK) return BN_SYNC_MAX_BLOCK;
    if (ctx->big_blocks && n > (size_t)ctx->sms * BN_FEXP_BLOCK) {
        size_t sm = (size_t)ctx->sms;
        size_t w7 = ((n + BN_FEXP_BLOCK - 1) / BN_FEXP_BLOCK + sm - 1) / sm;
        size_t w6 = ((n + BN_SYNC_MAX_BLOCK - 1) / BN_SYNC_MAX_BLOCK + sm - 1) / sm;
        if (w6 * 11 < w7 * 10) return BN_SYNC_MAX_BLOCK;
        return BN_FEXP_BLOCK;
    }
    size_t per_sm = (n + (size_t)ctx->sms - 1) / (size_t)ctx->sms;
    size_t b = ((per_sm + 31) / 32) * 32;
    if (b < 32) b = 32;
    if (b > BN_FEXP_BLOCK) b = BN_FEXP_BLOCK;
    return (unsigned)b;
}
static inline size_t sync_smem(unsigned bs) { return (size_t)bs * (bs > BN_SLOT_BLOCK7 ? BN_SLOT_STRIDE6 : BN_SLOT_STRIDE); }
static int use(bn254_ctx* ctx) {
    if (!ctx) return fail(BN254_ERR_ARG, "null context");
    CU(cudaSetDevice(ctx->device));
    return 0;
}
// device-side alias of a pinned, mapped host buffer (cudaHostAlloc / cudaHostRegister under UVA); nullptr for pageable memory
static void* mapped_device_ptr(const void* host) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return (a.type == cudaMemoryTypeHost) ? a.devicePointer : nullptr;
}
static int h2d(bn254_ctx* ctx, void* d, const void* h, size_t bytes) {
    if (bytes) CU(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return 0;
}
static int d2h(bn254_ctx* ctx, void* h, const void* d, size_t bytes) {
    if (bytes) CU(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return 0;
}

extern "C" {

const char* bn254_last_error(void) { return g_err.c_str(); }
const char* bn254_build_info(void) {
    static char buf[160];
    snprintf(buf, sizeof buf, "bn254_b200 sm_100a block=%d min_blocks=%d cuda=%d", BN_BLOCK, BN_MIN_BLOCKS, CUDART_VERSION);
    return buf;
}

int bn254_ctx_create(int device, bn254_ctx** out) {
    if (!out) return fail(BN254_ERR_ARG, "null out");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return fail(BN254_ERR_CUDA, "no CUDA device (this library has no CPU fallback)", e);
    if (device < 0 || device >= count) return fail(BN254_ERR_ARG, "bad device index");
    CU(cudaSetDevice(device));
    bn254_ctx* c = new (std::nothrow) bn254_ctx();
    if (!c) return fail(BN254_ERR_NOMEM, "ctx alloc");
    c->device = device;
    c->launches = 0;
    c->ws = nullptr;
    c->ws_bytes = 0;
    c->io = nullptr;
    c->io_bytes = 0;
    c->ws_ovr = nullptr;
    c->ws_ovr_bytes = 0;
    c->force_block = 0;
    { const char* e = getenv("BN254_BIG_BLOCKS"); c->big_blocks = e ? atoi(e) : 1; }   // 2: 512-thread blocks for every batch above 512 (tests)
    c->sms = 148;
    cudaDeviceGetAttribute(&c->sms, cudaDevAttrMultiProcessorCount, device);
    if (c->sms <= 0) c->sms = 148;
    c->stream2 = nullptr;
    { const char* e = getenv("BN254_PIPELINE"); c->pipeline = !(e && e[0] == '0'); }
    { const char* e = getenv("BN254_FUSED_MILLER"); c->split_miller = (e && e[0] == '1') ? 0 : 1; }
    { const char* e = getenv("BN254_SPLIT_FEXP"); c->split_fexp = (e && e[0] == '1') ? 1 : 0; }
    { const char* e = getenv("BN254_FEXP_SMEM"); c->fexp_smem = e ? atoi(e) : 2; }
    { const char* e = getenv("BN254_MILLER_SMEM"); c->miller_smem = e ? atoi(e) : 1; }
    { const char* e = getenv("BN254_LINES_SMEM"); c->lines_smem = e ? atoi(e) : 1; }
    if (c->lines_smem) {
        cudaError_t ea = cudaFuncSetAttribute(k_g2_lines_sync, cudaFuncAttributeMaxDynamicSharedMemorySize, BN_FEXP_BLOCK * BN_SLOT_STRIDE);
        if (ea != cudaSuccess) { delete c; return fail(BN254_ERR_CUDA, "cudaFuncSetAttribute(k_g2_lines_sync)", ea); }
    }
    if (c->miller_smem) {
        cudaError_t ea = cudaFuncSetAttribute(k_miller_table_sync, cudaFuncAttributeMaxDynamicSharedMemorySize, BN_FEXP_BLOCK * BN_SLOT_STRIDE);
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(k_miller_product_sync<1, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, BN_FEXP_BLOCK * BN_SLOT_STRIDE);
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(k_miller_product_sync<4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, BN_FEXP_BLOCK * BN_SLOT_STRIDE);
        if (ea == cudaSuccess) ea = cudaFuncSetAttribute(k_miller_product_sync<8, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, BN_FEXP_BLOCK * BN_SLOT_STRIDE);
        if (ea != cudaSuccess) { delete c; return fail(BN254_ERR_CUDA, "cudaFuncSetAttribute(k_miller_table_sync)", ea); }
    }
    if (c->fexp_smem == 1) {
        cudaFuncSetAttribute(k_final_exp_hard_s, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    } else if (c->fexp_smem == 2) {
        cudaError_t ea = cudaFuncSetAttribute(k_final_exp_hard_sync, cudaFuncAttributeMaxDynamicSharedMemorySize, BN_FEXP_BLOCK * BN_SLOT_STRIDE);
        if (ea != cudaSuccess) { delete c; return fail(BN254_ERR_CUDA, "cudaFuncSetAttribute(k_final_exp_hard_sync)", ea); }
    }
    { const char* e = getenv("BN254_SUBGROUP_REF"); c->subgroup_ref = (e && e[0] == '1'); }
    { const char* e = getenv("BN254_ZEROCOPY_OUT"); c->zerocopy_out = !(e && e[0] == '0'); }
    // the main stream outranks the second one: in the pipelined path the first half batch gets ahead of the second, so its
    // results cross PCIe while the second half is still computing
    int prio_least = 0, prio_greatest = 0;
    cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
    e = cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_greatest);
    if (e != cudaSuccess) { delete c; return fail(BN254_ERR_CUDA, "cudaStreamCreate", e); }
    e = cudaStreamCreateWithPriority(&c->stream2, cudaStreamNonBlocking, prio_least);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming);
    if (e != cudaSuccess) { delete c; return fail(BN254_ERR_CUDA, "cudaStreamCreate (second stream)", e); }
    // deep call chains (Fq12 -> Fq6 -> Fq2 noinline bodies) need more than the default stack
    cudaDeviceSetLimit(cudaLimitStackSize, 16 * 1024);
    *out = c;
    return 0;
}
void bn254_ctx_destroy(bn254_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaStreamDestroy(ctx->stream);
    if (ctx->stream2) { cudaStreamSynchronize(ctx->stream2); cudaStreamDestroy(ctx->stream2); cudaEventDestroy(ctx->ev_fork); cudaEventDestroy(ctx->ev_join); }
    if (ctx->ws) cudaFree(ctx->ws);
    if (ctx->io) cudaFree(ctx->io);
    delete ctx;
}
int bn254_ctx_device(const bn254_ctx* ctx) { return ctx ? ctx->device : -1; }
void* bn254_ctx_stream(const bn254_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int bn254_ctx_synchronize(bn254_ctx* ctx) {
    TRY(use(ctx));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
uint64_t bn254_ctx_launch_count(const bn254_ctx* ctx) { return ctx ? ctx->launches.load() : 0; }

}  // extern "C"
// grow-only workspace: stream-ordered reuse is safe because every user enqueues on ctx->stream
static int ctx_workspace(bn254_ctx* ctx, size_t bytes, void** out) {
    if (ctx->ws_ovr) {
        if (bytes > ctx->ws_ovr_bytes) return fail(BN254_ERR_ARG, "internal: workspace region too small");
        *out = ctx->ws_ovr;
        return 0;
    }
    if (bytes > ctx->ws_bytes) {
        if (ctx->ws) { CU(cudaStreamSynchronize(ctx->stream)); cudaFree(ctx->ws); ctx->ws = nullptr; ctx->ws_bytes = 0; }
        cudaError_t e = cudaMalloc(&ctx->ws, bytes);
        if (e != cudaSuccess) { ctx->ws = nullptr; return fail(BN254_ERR_NOMEM, "cudaMalloc workspace", e); }
        ctx->ws_bytes = bytes;
    }
    *out = ctx->ws;
    return 0;
}
// final exponentiation of n values (in place allowed): single kernel, or easy + hard kernels
static int final_exp_dev(bn254_ctx* ctx, const Fq12* d_in, Fq12* d_out, uint8_t* d_some, uint8_t* d_is_one, size_t n) {
    if (!ctx->split_fexp && !ctx->fexp_smem) {
        LAUNCH(ctx, k_final_exp_many, n, d_in, d_out, d_some, d_is_one, n);
        return 0;
    }
    // scratch: easy-part values + flags, placed after the Miller tables' use (stream-ordered, so reuse is safe)
    void* ws = nullptr;
    size_t vals = n * sizeof(Fq12);
    TRY(ctx_workspace(ctx, vals + ((n + 15) & ~(size_t)15), &ws));
    Fq12* t = (Fq12*)ws;
    uint8_t* flags = (uint8_t*)ws + vals;
    LAUNCH(ctx, k_final_exp_easy, n, d_in, t, flags, n);
    if (ctx->fexp_smem == 2) {
        unsigned bs = sync_block(ctx, n), g = (unsigned)((n + bs - 1) / bs);
        k_final_exp_hard_sync<<<g, bs, sync_smem(bs), ctx->stream>>>((const Fq12*)t, (const uint8_t*)flags, d_out, d_some, d_is_one, n);
        ctx->launches.fetch_add(1);
        CU(cudaGetLastError());
        return 0;
    }
    if (ctx->fexp_smem) {
        k_final_exp_hard_s<<<grid_for(n), BN_BLOCK, BN_BLOCK * BN_SLOT_STRIDE, ctx->stream>>>((const Fq12*)t, (const uint8_t*)flags, d_out, d_some, d_is_one, n);
        ctx->launches.fetch_add(1);
        CU(cudaGetLastError());
        return 0;
    }
    LAUNCH(ctx, k_final_exp_hard, n, (const Fq12*)t, (const uint8_t*)flags, d_out, d_some, d_is_one, n);
    return 0;
}
// I/O arena of the synchronous host-pointer entry points: carve `count` 256-byte-aligned regions
static int ctx_io(bn254_ctx* ctx, const size_t* sizes, int count, void** out) {
    size_t total = 0;
    for (int i = 0; i < count; i++) total += (sizes[i] + 255) & ~(size_t)255;
    if (total > ctx->io_bytes) {
        if (ctx->io) { CU(cudaStreamSynchronize(ctx->stream)); cudaFree(ctx->io); ctx->io = nullptr; ctx->io_bytes = 0; }
        cudaError_t e = cudaMalloc(&ctx->io, total);
        if (e != cudaSuccess) { ctx->io = nullptr; return fail(BN254_ERR_NOMEM, "cudaMalloc I/O arena", e); }
        ctx->io_bytes = total;
    }
    char* p = (char*)ctx->io;
    for (int i = 0; i < count; i++) { out[i] = p; p += (sizes[i] + 255) & ~(size_t)255; }
    return 0;
}
// line tables of m affine G2 points (slot layout for v pairs per check): slot-file kernel or the local-memory kernel
static int lines_launch(bn254_ctx* ctx, const AffG2* qa, const uint8_t* ok, uint4* tab, size_t m, size_t v) {
    // one wave only: with more work the 64-thread kernel reaches 16 warps per SM (the slot file allows 14) and is faster
    // (2^17 points: 5.6 against 6.0 ms; 2^16: 3.24 against 3.01 ms)
    if (ctx->lines_smem && m <= (size_t)ctx->sms * BN_FEXP_BLOCK) {
        unsigned bs = sync_block(ctx, m);
        if (bs > BN_FEXP_BLOCK) bs = BN_FEXP_BLOCK;   // this kernel uses all 7 slots
        unsigned g = (unsigned)((m + bs - 1) / bs);
        k_g2_lines_sync<<<g, bs, sync_smem(bs), ctx->stream>>>(qa, ok, tab, m, v);
        ctx->launches.fetch_add(1);
        CU(cudaGetLastError());
        return 0;
    }
    LAUNCH(ctx, k_g2_lines_affine, m, qa, ok, tab, m, v);
    return 0;
}
// Miller loop of m pairs from their line tables: slot-file kernel (one 448-thread block per SM) or the local-memory kernel
static int miller_table_launch(bn254_ctx* ctx, const AffG1* pa, const uint8_t* ok, const uint4* tab, Fq12* f, uint8_t* ok_out, size_t m) {
    if (ctx->miller_smem) {
        unsigned bs = sync_block(ctx, m), g = (unsigned)((m + bs - 1) / bs);
        k_miller_table_sync<<<g, bs, sync_smem(bs), ctx->stream>>>(pa, ok, tab, f, ok_out, m);
        ctx->launches.fetch_add(1);
        CU(cudaGetLastError());
        return 0;
    }
    LAUNCH(ctx, k_miller_table_affine, m, pa, ok, tab, f, ok_out, m);
    return 0;
}
static const size_t MILLER_CHUNK = (size_t)1 << 17;   // pairings per table chunk (2.2 GB of line coefficients)
static size_t miller_ws_bytes(size_t n) {
    size_t chunk = n < MILLER_CHUNK ? n : MILLER_CHUNK;
    return (size_t)BN254_N_COEFFS * 12 * chunk * sizeof(uint4) + chunk * sizeof(AffG2) + chunk * sizeof(AffG1) + ((chunk + 15) & ~(size_t)15);
}
static size_t fexp_ws_bytes(size_t n) { return n * sizeof(Fq12) + ((n + 15) & ~(size_t)15); }
// Miller values of n independent pairs (fused single kernel, or affine -> line table -> Miller from table)
static int miller_many_dev(bn254_ctx* ctx, const G2* d_g2, const G1* d_g1, Fq12* d_f, uint8_t* d_ok, size_t n) {
    if (!ctx->split_miller) {
        LAUNCH(ctx, k_miller_many, n, d_g2, d_g1, d_f, d_ok, n);
        return 0;
    }
    size_t chunk = n < MILLER_CHUNK ? n : MILLER_CHUNK;
    size_t tab_bytes = (size_t)BN254_N_COEFFS * 12 * chunk * sizeof(uint4);
    size_t qa_off = tab_bytes, pa_off = qa_off + chunk * sizeof(AffG2), ok_off = pa_off + chunk * sizeof(AffG1);
    void* ws = nullptr;
    TRY(ctx_workspace(ctx, ok_off + ((chunk + 15) & ~(size_t)15), &ws));
    uint4* tab = (uint4*)ws;
    AffG2* qa = (AffG2*)((char*)ws + qa_off);
    AffG1* pa = (AffG1*)((char*)ws + pa_off);
    uint8_t* ok = (uint8_t*)ws + ok_off;
    for (size_t lo = 0; lo < n; lo += chunk) {
        size_t m = n - lo < chunk ? n - lo : chunk;
        LAUNCH(ctx, k_pairs_to_affine, m, d_g2 + lo, d_g1 + lo, qa, pa, ok, m);
        TRY(lines_launch(ctx, (const AffG2*)qa, (const uint8_t*)ok, tab, m, (size_t)1));
        TRY(miller_table_launch(ctx, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tab, d_f + lo, d_ok ? d_ok + lo : (uint8_t*)nullptr, m));
    }
    return 0;
}
extern "C" {

// ---------------------------------------------------------------------------------------------
// pairing
// ---------------------------------------------------------------------------------------------
int bn254_pairing_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_g2, void* d_gt_out, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!d_g1 || !d_g2 || !d_gt_out) return fail(BN254_ERR_ARG, "null pointer");
    // Miller values are written to the output buffer and exponentiated in place
    TRY(miller_many_dev(ctx, (const G2*)d_g2, (const G1*)d_g1, (Fq12*)d_gt_out, (uint8_t*)nullptr, n));
    TRY(final_exp_dev(ctx, (const Fq12*)d_gt_out, (Fq12*)d_gt_out, (uint8_t*)nullptr, (uint8_t*)nullptr, n));
    return 0;
}
int bn254_pairing_many_dev_timed(bn254_ctx* ctx, const void* d_g1, const void* d_g2, void* d_gt_out, size_t n, float* ms) {
    TRY(use(ctx));
    if (!ms) return fail(BN254_ERR_ARG, "null pointer");
    ms[0] = ms[1] = ms[2] = ms[3] = 0.f;
    if (n == 0) return 0;
    if (!d_g1 || !d_g2 || !d_gt_out) return fail(BN254_ERR_ARG, "null pointer");
    cudaEvent_t ev[5];
    for (int i = 0; i < 5; i++) CU(cudaEventCreate(&ev[i]));
    const G1* g1 = (const G1*)d_g1;
    const G2* g2 = (const G2*)d_g2;
    Fq12* f = (Fq12*)d_gt_out;
    auto body = [&]() -> int {
        if (!ctx->split_miller) {
            CU(cudaEventRecord(ev[0], ctx->stream));
            LAUNCH(ctx, k_miller_many, n, g2, g1, f, (uint8_t*)nullptr, n);
            CU(cudaEventRecord(ev[1], ctx->stream));
            CU(cudaEventSynchronize(ev[1]));
            float t; CU(cudaEventElapsedTime(&t, ev[0], ev[1])); ms[2] = t;
        } else {
            size_t chunk = n < MILLER_CHUNK ? n : MILLER_CHUNK;
            size_t tab_bytes = (size_t)BN254_N_COEFFS * 12 * chunk * sizeof(uint4);
            size_t qa_off = tab_bytes, pa_off = qa_off + chunk * sizeof(AffG2), ok_off = pa_off + chunk * sizeof(AffG1);
            void* ws = nullptr;
            TRY(ctx_workspace(ctx, ok_off + ((chunk + 15) & ~(size_t)15), &ws));
            uint4* tab = (uint4*)ws;
            AffG2* qa = (AffG2*)((char*)ws + qa_off);
            AffG1* pa = (AffG1*)((char*)ws + pa_off);
            uint8_t* ok = (uint8_t*)ws + ok_off;
            for (size_t lo = 0; lo < n; lo += chunk) {
                size_t m = n - lo < chunk ? n - lo : chunk;
                CU(cudaEventRecord(ev[0], ctx->stream));
                LAUNCH(ctx, k_pairs_to_affine, m, g2 + lo, g1 + lo, qa, pa, ok, m);
                CU(cudaEventRecord(ev[1], ctx->stream));
                TRY(lines_launch(ctx, (const AffG2*)qa, (const uint8_t*)ok, tab, m, (size_t)1));
                CU(cudaEventRecord(ev[2], ctx->stream));
                TRY(miller_table_launch(ctx, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tab, f + lo, (uint8_t*)nullptr, m));
                CU(cudaEventRecord(ev[3], ctx->stream));
                CU(cudaEventSynchronize(ev[3]));
                for (int i = 0; i < 3; i++) { float t; CU(cudaEventElapsedTime(&t, ev[i], ev[i + 1])); ms[i] += t; }
            }
        }
        CU(cudaEventRecord(ev[3], ctx->stream));
        TRY(final_exp_dev(ctx, (const Fq12*)f, f, (uint8_t*)nullptr, (uint8_t*)nullptr, n));
        CU(cudaEventRecord(ev[4], ctx->stream));
        CU(cudaEventSynchronize(ev[4]));
        float t; CU(cudaEventElapsedTime(&t, ev[3], ev[4])); ms[3] = t;
        return 0;
    };
    int rc = body();
    for (int i = 0; i < 5; i++) cudaEventDestroy(ev[i]);
    return rc;
}
// one (half) batch of the host-buffer pairing path on ctx->stream: copies in, kernels, results out
static int pairing_many_enqueue(bn254_ctx* ctx, const char* g1, const char* g2, char* gt_out, char* mapped, void* const* d, size_t lo, size_t m) {
    G1* dg1 = (G1*)d[0] + lo;
    G2* dg2 = (G2*)d[1] + lo;
    Fq12* df = (Fq12*)d[2] + lo;
    TRY(h2d(ctx, dg1, g1 + lo * sizeof(G1), m * sizeof(G1)));
    TRY(h2d(ctx, dg2, g2 + lo * sizeof(G2), m * sizeof(G2)));
    TRY(miller_many_dev(ctx, dg2, dg1, df, (uint8_t*)nullptr, m));
    if (mapped) {
        TRY(final_exp_dev(ctx, df, (Fq12*)mapped + lo, (uint8_t*)nullptr, (uint8_t*)nullptr, m));
    } else {
        TRY(final_exp_dev(ctx, df, df, (uint8_t*)nullptr, (uint8_t*)nullptr, m));
        TRY(d2h(ctx, gt_out + lo * sizeof(Fq12), df, m * sizeof(Fq12)));
    }
    return 0;
}
int bn254_pairing_many(bn254_ctx* ctx, const void* g1, const void* g2, void* gt_out, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!g1 || !g2 || !gt_out) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[3] = {n * sizeof(G1), n * sizeof(G2), n * sizeof(Fq12)};
    void* d[3];
    TRY(ctx_io(ctx, sz, 3, d));
    // Pinned (mapped) output buffer: the final-exponentiation kernel stores its 384-byte results straight into host memory.
    char* mapped = ctx->zerocopy_out ? (char*)mapped_device_ptr(gt_out) : nullptr;
    const size_t PIPE_MIN = (size_t)1 << 16;   // below this the halves would leave SMs idle (74 blocks of 448 threads per half at 2^16)
    if (!ctx->pipeline || n < PIPE_MIN || !ctx->split_miller) {
        TRY(pairing_many_enqueue(ctx, (const char*)g1, (const char*)g2, (char*)gt_out, mapped, d, 0, n));
        CU(cudaStreamSynchronize(ctx->stream));
        return 0;
    }
    // Two half batches on two streams.  The one-wave kernels of the slot-file path finish all their blocks together, so a
    // single batch exposes both transfers completely (inputs before the first kernel, results after the last one).  With two
    // halves the second half's inputs travel while the first computes; each half's 448-thread kernels occupy half of the SMs.
    // Measured: +1.5 % end to end at 2^16 (the halves still finish together, so the results' transfer stays exposed).
    // The first half is a whole number of blocks.
    size_t nA = ((n / 2 + BN_FEXP_BLOCK - 1) / BN_FEXP_BLOCK) * BN_FEXP_BLOCK, nB = n - nA;
    size_t wA = miller_ws_bytes(nA), wB = miller_ws_bytes(nB);
    if (fexp_ws_bytes(nA) > wA) wA = fexp_ws_bytes(nA);
    if (fexp_ws_bytes(nB) > wB) wB = fexp_ws_bytes(nB);
    wA = (wA + 255) & ~(size_t)255;
    void* ws = nullptr;
    TRY(ctx_workspace(ctx, wA + wB, &ws));
    cudaStream_t main_stream = ctx->stream;
    CU(cudaEventRecord(ctx->ev_fork, main_stream));
    CU(cudaStreamWaitEvent(ctx->stream2, ctx->ev_fork, 0));
    ctx->ws_ovr = ws; ctx->ws_ovr_bytes = wA;
    // one-wave batches: each half on its own SMs with full blocks (two smaller blocks per SM would interleave their phases);
    // larger batches fill every SM with one block at a time anyway and keep the block size rule of sync_block
    if (n <= (size_t)ctx->sms * BN_FEXP_BLOCK) ctx->force_block = BN_FEXP_BLOCK;
    int rc = pairing_many_enqueue(ctx, (const char*)g1, (const char*)g2, (char*)gt_out, mapped, d, 0, nA);
    if (rc == 0) {
        ctx->stream = ctx->stream2;
        ctx->ws_ovr = (char*)ws + wA; ctx->ws_ovr_bytes = wB;
        rc = pairing_many_enqueue(ctx, (const char*)g1, (const char*)g2, (char*)gt_out, mapped, d, nA, nB);
        ctx->stream = main_stream;
    }
    ctx->ws_ovr = nullptr; ctx->ws_ovr_bytes = 0;
    ctx->force_block = 0;
    cudaEventRecord(ctx->ev_join, ctx->stream2);
    cudaStreamWaitEvent(main_stream, ctx->ev_join, 0);
    cudaError_t es = cudaStreamSynchronize(main_stream);
    if (rc != 0) return rc;
    if (es != cudaSuccess) return fail(BN254_ERR_CUDA, "cudaStreamSynchronize", es);
    return 0;
}

int bn254_miller_loop_many_dev(bn254_ctx* ctx, const void* d_g2, const void* d_g1, void* d_gt_out, uint8_t* d_ok, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!d_g1 || !d_g2 || !d_gt_out) return fail(BN254_ERR_ARG, "null pointer");
    TRY(miller_many_dev(ctx, (const G2*)d_g2, (const G1*)d_g1, (Fq12*)d_gt_out, d_ok, n));
    return 0;
}
int bn254_miller_loop_many(bn254_ctx* ctx, const void* g2, const void* g1, void* gt_out, uint8_t* ok, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!g1 || !g2 || !gt_out) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf a, b, o, k;
    TRY(a.alloc(n * sizeof(G1))); TRY(b.alloc(n * sizeof(G2))); TRY(o.alloc(n * sizeof(Fq12))); TRY(k.alloc(n));
    TRY(h2d(ctx, a.p, g1, n * sizeof(G1)));
    TRY(h2d(ctx, b.p, g2, n * sizeof(G2)));
    TRY(bn254_miller_loop_many_dev(ctx, b.p, a.p, o.p, k.as<uint8_t>(), n));
    TRY(d2h(ctx, gt_out, o.p, n * sizeof(Fq12)));
    if (ok) TRY(d2h(ctx, ok, k.p, n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// product checks (pairing_batch) and miller_loop_batch
// ---------------------------------------------------------------------------------------------
// reduce n Fq12 values in `a` to one product; `b` is scratch of >= (n+1)/2 elements.  Result pointer returned.
static int tree_product(bn254_ctx* ctx, Fq12* a, Fq12* b, size_t n, Fq12** result) {
    Fq12 *src = a, *dst = b;
    while (n > 1) {
        size_t n_out = (n + 1) / 2;
        LAUNCH(ctx, k_fq12_pair_product, n_out, (const Fq12*)src, n, dst);
        n = n_out;
        Fq12* t = src; src = dst; dst = t;
    }
    *result = src;
    return 0;
}

// Miller products of n_checks checks into d_f (n_checks x Fq12); shared tables optional
static int miller_products_dev(bn254_ctx* ctx, const G1* d_g1v, const G2* d_g2v, size_t v, const G1* d_g1s,
                               const bn254_g2precomp* pc, size_t n_checks, Fq12* d_f) {
    size_t m = pc ? pc->n : 0;
    const uint4* tab = pc ? pc->tab : nullptr;
    if (m > 8) return fail(BN254_ERR_ARG, "at most 8 shared tables per check");
    if (v > 8) {
        // many pairs per check: independent Miller loops, then a per-check product (same field element)
        if (m != 0) return fail(BN254_ERR_ARG, "more than 8 variable pairs cannot be combined with shared tables");
        DevBuf tmp;
        TRY(tmp.alloc(n_checks * v * sizeof(Fq12)));
        TRY(miller_many_dev(ctx, d_g2v, d_g1v, tmp.as<Fq12>(), (uint8_t*)nullptr, n_checks * v));
        LAUNCH(ctx, k_fq12_group_product, n_checks, (const Fq12*)tmp.as<Fq12>(), v, d_f, n_checks);
        CU(cudaStreamSynchronize(ctx->stream));   // tmp is freed on return
        return 0;
    }
    if (!ctx->split_miller || v == 0) {
        // single kernel: line coefficients of the variable G2 points produced on the fly
        if (v <= 1 && m <= 4) {
            LAUNCH(ctx, (k_miller_product<1, 4>), n_checks, d_g1v, d_g2v, (int)v, d_g1s, tab, (int)m, d_f, n_checks);
        } else if (v <= 4 && m <= 4) {
            LAUNCH(ctx, (k_miller_product<4, 4>), n_checks, d_g1v, d_g2v, (int)v, d_g1s, tab, (int)m, d_f, n_checks);
        } else {
            LAUNCH(ctx, (k_miller_product<8, 8>), n_checks, d_g1v, d_g2v, (int)v, d_g1s, tab, (int)m, d_f, n_checks);
        }
        return 0;
    }
    // split path: affine conversion and line tables of the variable pairs (slot layout of table_slot), then one
    // product kernel per chunk of checks reading per-check and shared tables
    // checks per chunk: at least 2^16 for up to 4 pairs per check (one full wave of the 448-thread kernel on 148 SMs; 4.4 GB of
    // line tables at v = 4), 2^15 above
    size_t chunk = MILLER_CHUNK / v;
    if (ctx->miller_smem) { size_t want = v <= 4 ? ((size_t)1 << 16) : ((size_t)1 << 15); if (chunk < want) chunk = want; }
    if (chunk > n_checks) chunk = n_checks;
    size_t pairs = chunk * v;
    size_t tab_bytes = (size_t)BN254_N_COEFFS * 12 * pairs * sizeof(uint4);
    size_t qa_off = tab_bytes, pa_off = qa_off + pairs * sizeof(AffG2), ok_off = pa_off + pairs * sizeof(AffG1);
    void* ws = nullptr;
    TRY(ctx_workspace(ctx, ok_off + ((pairs + 15) & ~(size_t)15), &ws));
    uint4* tabv = (uint4*)ws;
    AffG2* qa = (AffG2*)((char*)ws + qa_off);
    AffG1* pa = (AffG1*)((char*)ws + pa_off);
    uint8_t* ok = (uint8_t*)ws + ok_off;
    for (size_t lo = 0; lo < n_checks; lo += chunk) {
        size_t nc = n_checks - lo < chunk ? n_checks - lo : chunk;
        size_t np = nc * v;
        LAUNCH(ctx, k_pairs_to_affine, np, d_g2v + lo * v, d_g1v + lo * v, qa, pa, ok, np);
        TRY(lines_launch(ctx, (const AffG2*)qa, (const uint8_t*)ok, tabv, np, v));
        const G1* gs = d_g1s ? d_g1s + lo * m : nullptr;
        if (ctx->miller_smem) {
            unsigned bs = sync_block(ctx, nc), g = (unsigned)((nc + bs - 1) / bs);
            size_t smem = sync_smem(bs);
            if (v <= 1 && m <= 4) k_miller_product_sync<1, 4><<<g, bs, smem, ctx->stream>>>((const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
            else if (v <= 4 && m <= 4) k_miller_product_sync<4, 4><<<g, bs, smem, ctx->stream>>>((const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
            else k_miller_product_sync<8, 8><<<g, bs, smem, ctx->stream>>>((const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
            ctx->launches.fetch_add(1);
            CU(cudaGetLastError());
        } else if (v <= 1 && m <= 4) {
            LAUNCH(ctx, (k_miller_product_tables<1, 4>), nc, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
        } else if (v <= 4 && m <= 4) {
            LAUNCH(ctx, (k_miller_product_tables<4, 4>), nc, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
        } else {
            LAUNCH(ctx, (k_miller_product_tables<8, 8>), nc, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
        }
    }
    return 0;
}

int bn254_pairing_product_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_g2, size_t pairs_per_check,
                                   size_t n_checks, void* d_gt_out, uint8_t* d_is_one) {
    TRY(use(ctx));
    if (n_checks == 0) return 0;
    if (pairs_per_check && (!d_g1 || !d_g2)) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf f;
    Fq12* d_f = (Fq12*)d_gt_out;
    if (!d_f) { TRY(f.alloc(n_checks * sizeof(Fq12))); d_f = f.as<Fq12>(); }
    TRY(miller_products_dev(ctx, (const G1*)d_g1, (const G2*)d_g2, pairs_per_check, nullptr, nullptr, n_checks, d_f));
    TRY(final_exp_dev(ctx, (const Fq12*)d_f, d_gt_out ? d_f : (Fq12*)nullptr, (uint8_t*)nullptr, d_is_one, n_checks));
    if (!d_gt_out) CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int bn254_pairing_product_many(bn254_ctx* ctx, const void* g1, const void* g2, size_t pairs_per_check, size_t n_checks,
                               void* gt_out, uint8_t* is_one) {
    TRY(use(ctx));
    if (n_checks == 0) return 0;
    size_t n = pairs_per_check * n_checks;
    if (n && (!g1 || !g2)) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[4] = {n * sizeof(G1), n * sizeof(G2), n_checks * sizeof(Fq12), n_checks};
    void* d[4];
    TRY(ctx_io(ctx, sz, 4, d));
    TRY(h2d(ctx, d[0], g1, sz[0]));
    TRY(h2d(ctx, d[1], g2, sz[1]));
    TRY(bn254_pairing_product_many_dev(ctx, d[0], d[1], pairs_per_check, n_checks, d[2], (uint8_t*)d[3]));
    if (gt_out) TRY(d2h(ctx, gt_out, d[2], sz[2]));
    if (is_one) TRY(d2h(ctx, is_one, d[3], n_checks));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int bn254_miller_loop_batch(bn254_ctx* ctx, const void* g2, const void* g1, size_t n, void* gt_out) {
    TRY(use(ctx));
    if (!gt_out) return fail(BN254_ERR_ARG, "null pointer");
    if (n == 0) {   // empty product: Fq12::one(), src/groups/mod.rs:610
        DevBuf o;
        TRY(o.alloc(sizeof(Fq12)));
        LAUNCH(ctx, (k_miller_product<1, 4>), 1, (const G1*)nullptr, (const G2*)nullptr, 0, (const G1*)nullptr,
               (const uint4*)nullptr, 0, o.as<Fq12>(), (size_t)1);
        TRY(d2h(ctx, gt_out, o.p, sizeof(Fq12)));
        CU(cudaStreamSynchronize(ctx->stream));
        return 0;
    }
    if (!g1 || !g2) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf a, b, f, s, k;
    TRY(a.alloc(n * sizeof(G1))); TRY(b.alloc(n * sizeof(G2))); TRY(f.alloc(n * sizeof(Fq12)));
    TRY(s.alloc(((n + 1) / 2) * sizeof(Fq12))); TRY(k.alloc(n));
    TRY(h2d(ctx, a.p, g1, n * sizeof(G1)));
    TRY(h2d(ctx, b.p, g2, n * sizeof(G2)));
    LAUNCH(ctx, k_miller_many, n, b.as<G2>(), a.as<G1>(), f.as<Fq12>(), k.as<uint8_t>(), n);
    std::vector<uint8_t> ok(n);
    TRY(d2h(ctx, ok.data(), k.p, n));
    CU(cudaStreamSynchronize(ctx->stream));
    for (size_t i = 0; i < n; i++)
        if (!ok[i]) return BN254_ERR_TO_AFFINE;   // Err(CurveError::ToAffineConversion), src/lib.rs:629-630
    Fq12* res = nullptr;
    TRY(tree_product(ctx, f.as<Fq12>(), s.as<Fq12>(), n, &res));
    TRY(d2h(ctx, gt_out, res, sizeof(Fq12)));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int bn254_miller_product(bn254_ctx* ctx, const void* g1, const void* g2, size_t n, void* gt_out) {
    TRY(use(ctx));
    if (!gt_out) return fail(BN254_ERR_ARG, "null pointer");
    if (n && (!g1 || !g2)) return fail(BN254_ERR_ARG, "null pointer");
    size_t cnt = n ? n : 1;
    size_t sz[4] = {cnt * sizeof(G1), cnt * sizeof(G2), cnt * sizeof(Fq12), ((cnt + 1) / 2) * sizeof(Fq12)};
    void* d[4];
    TRY(ctx_io(ctx, sz, 4, d));
    Fq12* res = (Fq12*)d[2];
    if (n == 0) {
        LAUNCH(ctx, (k_miller_product<1, 4>), 1, (const G1*)nullptr, (const G2*)nullptr, 0, (const G1*)nullptr,
               (const uint4*)nullptr, 0, res, (size_t)1);
    } else {
        TRY(h2d(ctx, d[0], g1, sz[0]));
        TRY(h2d(ctx, d[1], g2, sz[1]));
        TRY(miller_many_dev(ctx, (const G2*)d[1], (const G1*)d[0], (Fq12*)d[2], (uint8_t*)nullptr, n));
        TRY(tree_product(ctx, (Fq12*)d[2], (Fq12*)d[3], n, &res));
    }
    TRY(d2h(ctx, gt_out, res, sizeof(Fq12)));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Gt
// ---------------------------------------------------------------------------------------------
int bn254_final_exp_many_dev(bn254_ctx* ctx, const void* d_in, void* d_out, uint8_t* d_some, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!d_in || !d_out) return fail(BN254_ERR_ARG, "null pointer");
    TRY(final_exp_dev(ctx, (const Fq12*)d_in, (Fq12*)d_out, d_some, (uint8_t*)nullptr, n));
    return 0;
}
int bn254_final_exp_many(bn254_ctx* ctx, const void* gt_in, void* gt_out, uint8_t* some, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!gt_in || !gt_out) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf a, o, k;
    TRY(a.alloc(n * sizeof(Fq12))); TRY(o.alloc(n * sizeof(Fq12))); TRY(k.alloc(n));
    TRY(h2d(ctx, a.p, gt_in, n * sizeof(Fq12)));
    TRY(bn254_final_exp_many_dev(ctx, a.p, o.p, k.as<uint8_t>(), n));
    TRY(d2h(ctx, gt_out, o.p, n * sizeof(Fq12)));
    if (some) TRY(d2h(ctx, some, k.p, n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int bn254_gt_mul_many(bn254_ctx* ctx, const void* a, const void* b, void* out, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!a || !b || !out) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf x, y, o;
    TRY(x.alloc(n * sizeof(Fq12))); TRY(y.alloc(n * sizeof(Fq12))); TRY(o.alloc(n * sizeof(Fq12)));
    TRY(h2d(ctx, x.p, a, n * sizeof(Fq12))); TRY(h2d(ctx, y.p, b, n * sizeof(Fq12)));
    LAUNCH(ctx, k_gt_mul, n, x.as<Fq12>(), y.as<Fq12>(), o.as<Fq12>(), n);
    TRY(d2h(ctx, out, o.p, n * sizeof(Fq12)));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int bn254_gt_pow_many(bn254_ctx* ctx, const void* a, const void* fr, void* out, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!a || !fr || !out) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf x, y, o;
    TRY(x.alloc(n * sizeof(Fq12))); TRY(y.alloc(n * sizeof(Fq))); TRY(o.alloc(n * sizeof(Fq12)));
    TRY(h2d(ctx, x.p, a, n * sizeof(Fq12))); TRY(h2d(ctx, y.p, fr, n * sizeof(Fq)));
    LAUNCH(ctx, k_gt_pow, n, x.as<Fq12>(), y.as<Fq>(), o.as<Fq12>(), n);
    TRY(d2h(ctx, out, o.p, n * sizeof(Fq12)));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int bn254_gt_inverse_many(bn254_ctx* ctx, const void* a, void* out, uint8_t* some, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!a || !out) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf x, o, k;
    TRY(x.alloc(n * sizeof(Fq12))); TRY(o.alloc(n * sizeof(Fq12))); TRY(k.alloc(n));
    TRY(h2d(ctx, x.p, a, n * sizeof(Fq12)));
    LAUNCH(ctx, k_gt_inverse, n, x.as<Fq12>(), o.as<Fq12>(), k.as<uint8_t>(), n);
    TRY(d2h(ctx, out, o.p, n * sizeof(Fq12)));
    if (some) TRY(d2h(ctx, some, k.p, n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int bn254_gt_eq_many(bn254_ctx* ctx, const void* a, const void* b, uint8_t* eq, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!a || !b || !eq) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf x, y, k;
    TRY(x.alloc(n * sizeof(Fq12))); TRY(y.alloc(n * sizeof(Fq12))); TRY(k.alloc(n));
    TRY(h2d(ctx, x.p, a, n * sizeof(Fq12))); TRY(h2d(ctx, y.p, b, n * sizeof(Fq12)));
    LAUNCH(ctx, k_gt_eq, n, x.as<Fq12>(), y.as<Fq12>(), k.as<uint8_t>(), n);
    TRY(d2h(ctx, eq, k.p, n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// G1 * Fr, G2 * Fr, normalize
// ---------------------------------------------------------------------------------------------
int bn254_g1_mul_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_fr, void* d_out, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!d_g1 || !d_fr || !d_out) return fail(BN254_ERR_ARG, "null pointer");
    LAUNCH(ctx, k_jac_mul<FqOps>, n, (const G1*)d_g1, (const Fq*)d_fr, (G1*)d_out, n);
    return 0;
}
int bn254_g2_mul_many_dev(bn254_ctx* ctx, const void* d_g2, const void* d_fr, void* d_out, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!d_g2 || !d_fr || !d_out) return fail(BN254_ERR_ARG, "null pointer");
    LAUNCH(ctx, k_jac_mul<Fq2Ops>, n, (const G2*)d_g2, (const Fq*)d_fr, (G2*)d_out, n);
    return 0;
}
}  // extern "C"
template <class P>
static int scalar_mul_host(bn254_ctx* ctx, const void* p, const void* fr, void* out, size_t n,
                           int (*dev)(bn254_ctx*, const void*, const void*, void*, size_t)) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!p || !fr || !out) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf a, b, o;
    TRY(a.alloc(n * sizeof(P))); TRY(b.alloc(n * sizeof(Fq))); TRY(o.alloc(n * sizeof(P)));
    TRY(h2d(ctx, a.p, p, n * sizeof(P))); TRY(h2d(ctx, b.p, fr, n * sizeof(Fq)));
    TRY(dev(ctx, a.p, b.p, o.p, n));
    TRY(d2h(ctx, out, o.p, n * sizeof(P)));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
extern "C" {
int bn254_g1_mul_many(bn254_ctx* ctx, const void* g1, const void* fr, void* out, size_t n) {
    return scalar_mul_host<G1>(ctx, g1, fr, out, n, bn254_g1_mul_many_dev);
}
int bn254_g2_mul_many(bn254_ctx* ctx, const void* g2, const void* fr, void* out, size_t n) {
    return scalar_mul_host<G2>(ctx, g2, fr, out, n, bn254_g2_mul_many_dev);
}
int bn254_g1_normalize_many(bn254_ctx* ctx, const void* g1, void* out, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!g1 || !out) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf a, o;
    TRY(a.alloc(n * sizeof(G1))); TRY(o.alloc(n * sizeof(G1)));
    TRY(h2d(ctx, a.p, g1, n * sizeof(G1)));
    LAUNCH(ctx, k_jac_normalize<FqOps>, n, a.as<G1>(), o.as<G1>(), n);
    TRY(d2h(ctx, out, o.p, n * sizeof(G1)));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int bn254_g2_normalize_many(bn254_ctx* ctx, const void* g2, void* out, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!g2 || !out) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf a, o;
    TRY(a.alloc(n * sizeof(G2))); TRY(o.alloc(n * sizeof(G2)));
    TRY(h2d(ctx, a.p, g2, n * sizeof(G2)));
    LAUNCH(ctx, k_jac_normalize<Fq2Ops>, n, a.as<G2>(), o.as<G2>(), n);
    TRY(d2h(ctx, out, o.p, n * sizeof(G2)));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

static int affine_check_dev(bn254_ctx* ctx, const void* d_xy, uint8_t* d_status, size_t n, bool g2) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!d_xy || !d_status) return fail(BN254_ERR_ARG, "null pointer");
    if (g2) {
        LAUNCH(ctx, k_g2_affine_check, n, (const AffG2*)d_xy, d_status, n, ctx->subgroup_ref);
    } else {
        LAUNCH(ctx, k_g1_affine_check, n, (const AffG1*)d_xy, d_status, n);
    }
    return 0;
}
static int affine_check_host(bn254_ctx* ctx, const void* xy, uint8_t* status, size_t n, bool g2) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!xy || !status) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[2] = {n * (g2 ? sizeof(AffG2) : sizeof(AffG1)), n};
    void* d[2];
    TRY(ctx_io(ctx, sz, 2, d));
    TRY(h2d(ctx, d[0], xy, sz[0]));
    TRY(affine_check_dev(ctx, d[0], (uint8_t*)d[1], n, g2));
    TRY(d2h(ctx, status, d[1], n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int bn254_g1_affine_check_many(bn254_ctx* ctx, const void* xy, uint8_t* status, size_t n) {
    return affine_check_host(ctx, xy, status, n, false);
}
int bn254_g2_affine_check_many(bn254_ctx* ctx, const void* xy, uint8_t* status, size_t n) {
    return affine_check_host(ctx, xy, status, n, true);
}
int bn254_g1_affine_check_many_dev(bn254_ctx* ctx, const void* d_xy, uint8_t* d_status, size_t n) {
    return affine_check_dev(ctx, d_xy, d_status, n, false);
}
int bn254_g2_affine_check_many_dev(bn254_ctx* ctx, const void* d_xy, uint8_t* d_status, size_t n) {
    return affine_check_dev(ctx, d_xy, d_status, n, true);
}

static int from_compressed_dev(bn254_ctx* ctx, const uint8_t* d_enc, void* d_out, uint8_t* d_status, size_t n, bool g2) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!d_enc || !d_out || !d_status) return fail(BN254_ERR_ARG, "null pointer");
    if (g2) {
        LAUNCH(ctx, k_g2_from_compressed, n, d_enc, (G2*)d_out, d_status, n, ctx->subgroup_ref);
    } else {
        LAUNCH(ctx, k_g1_from_compressed, n, d_enc, (G1*)d_out, d_status, n);
    }
    return 0;
}
static int from_compressed_host(bn254_ctx* ctx, const uint8_t* enc, void* out, uint8_t* status, size_t n, bool g2) {
    TRY(use(ctx));
    if (n == 0) return 0;
    if (!enc || !out || !status) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[3] = {n * (g2 ? 65 : 33), n * (g2 ? sizeof(G2) : sizeof(G1)), n};
    void* d[3];
    TRY(ctx_io(ctx, sz, 3, d));
    TRY(h2d(ctx, d[0], enc, sz[0]));
    TRY(from_compressed_dev(ctx, (const uint8_t*)d[0], d[1], (uint8_t*)d[2], n, g2));
    TRY(d2h(ctx, out, d[1], sz[1]));
    TRY(d2h(ctx, status, d[2], n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int bn254_g1_from_compressed_many(bn254_ctx* ctx, const uint8_t* enc, void* g1_out, uint8_t* status, size_t n) {
    return from_compressed_host(ctx, enc, g1_out, status, n, false);
}
int bn254_g2_from_compressed_many(bn254_ctx* ctx, const uint8_t* enc, void* g2_out, uint8_t* status, size_t n) {
    return from_compressed_host(ctx, enc, g2_out, status, n, true);
}
int bn254_g1_from_compressed_many_dev(bn254_ctx* ctx, const uint8_t* d_enc, void* d_g1_out, uint8_t* d_status, size_t n) {
    return from_compressed_dev(ctx, d_enc, d_g1_out, d_status, n, false);
}
int bn254_g2_from_compressed_many_dev(bn254_ctx* ctx, const uint8_t* d_enc, void* d_g2_out, uint8_t* d_status, size_t n) {
    return from_compressed_dev(ctx, d_enc, d_g2_out, d_status, n, true);
}

// ---------------------------------------------------------------------------------------------
// shared G2Precomp tables
// ---------------------------------------------------------------------------------------------
void bn254_g2precomp_destroy(bn254_g2precomp* pc) {
    if (!pc) return;
    cudaSetDevice(pc->device);
    if (pc->tab) cudaFree(pc->tab);
    if (pc->q_aff) cudaFree(pc->q_aff);
    delete pc;
}
size_t bn254_g2precomp_count(const bn254_g2precomp* pc) { return pc ? pc->n : 0; }
int bn254_g2_precompute(bn254_ctx* ctx, const void* g2, size_t n, bn254_g2precomp** out) {
    TRY(use(ctx));
    if (!out || !g2 || n == 0) return fail(BN254_ERR_ARG, "bad argument");
    bn254_g2precomp* pc = new (std::nothrow) bn254_g2precomp();
    if (!pc) return fail(BN254_ERR_NOMEM, "precomp alloc");
    pc->device = ctx->device; pc->n = n; pc->tab = nullptr; pc->q_aff = nullptr;
    DevBuf a, bad;
    int rc = a.alloc(n * sizeof(G2));
    if (rc == 0) rc = bad.alloc(sizeof(int));
    if (rc == 0 && cudaMalloc((void**)&pc->tab, (size_t)BN254_N_COEFFS * 12 * n * sizeof(uint4)) != cudaSuccess) rc = fail(BN254_ERR_NOMEM, "cudaMalloc table");
    if (rc == 0 && cudaMalloc((void**)&pc->q_aff, n * sizeof(AffG2)) != cudaSuccess) rc = fail(BN254_ERR_NOMEM, "cudaMalloc q");
    if (rc != 0) { bn254_g2precomp_destroy(pc); return rc; }
    int h_bad = 0;
    auto body = [&]() -> int {
        CU(cudaMemsetAsync(bad.p, 0, sizeof(int), ctx->stream));
        TRY(h2d(ctx, a.p, g2, n * sizeof(G2)));
        LAUNCH(ctx, k_g2_precompute, n, a.as<G2>(), pc->tab, pc->q_aff, bad.as<int>(), n);
        TRY(d2h(ctx, &h_bad, bad.p, sizeof(int)));
        CU(cudaStreamSynchronize(ctx->stream));
        return 0;
    };
    rc = body();
    if (rc == 0 && h_bad) rc = BN254_ERR_TO_AFFINE;
    if (rc != 0) { bn254_g2precomp_destroy(pc); return rc; }
    *out = pc;
    return 0;
}
int bn254_g2precomp_export(bn254_ctx* ctx, const bn254_g2precomp* pc, size_t i, void* out) {
    TRY(use(ctx));
    if (!pc || !out || i >= pc->n) return fail(BN254_ERR_ARG, "bad argument");
    uint8_t* o = (uint8_t*)out;
    TRY(d2h(ctx, o, pc->q_aff + i, sizeof(AffG2)));
    // gather the strided 16-byte words of table i: one 2D copy (pitch = n*16 bytes)
    CU(cudaMemcpy2DAsync(o + sizeof(AffG2), sizeof(uint4), pc->tab + i, pc->n * sizeof(uint4), sizeof(uint4),
                         (size_t)BN254_N_COEFFS * 12, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int bn254_pairing_product_precomp_many_dev(bn254_ctx* ctx, const bn254_g2precomp* pc, const void* d_g1s, const void* d_g1v,
                                           const void* d_g2v, size_t v, size_t n_checks, void* d_gt_out, uint8_t* d_is_one) {
    TRY(use(ctx));
    if (n_checks == 0) return 0;
    if (pc && pc->device != ctx->device) return fail(BN254_ERR_ARG, "precomp tables live on another device");
    if ((pc && !d_g1s) || (v && (!d_g1v || !d_g2v))) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf f;
    Fq12* d_f = (Fq12*)d_gt_out;
    if (!d_f) { TRY(f.alloc(n_checks * sizeof(Fq12))); d_f = f.as<Fq12>(); }
    TRY(miller_products_dev(ctx, (const G1*)d_g1v, (const G2*)d_g2v, v, (const G1*)d_g1s, pc, n_checks, d_f));
    TRY(final_exp_dev(ctx, (const Fq12*)d_f, d_gt_out ? d_f : (Fq12*)nullptr, (uint8_t*)nullptr, d_is_one, n_checks));
    if (!d_gt_out) CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
int bn254_pairing_product_precomp_many(bn254_ctx* ctx, const bn254_g2precomp* pc, const void* g1_shared, const void* g1_var,
                                       const void* g2_var, size_t v, size_t n_checks, void* gt_out, uint8_t* is_one) {
    TRY(use(ctx));
    if (n_checks == 0) return 0;
    size_t m = pc ? pc->n : 0;
    if ((m && !g1_shared) || (v && (!g1_var || !g2_var))) return fail(BN254_ERR_ARG, "null pointer");
    DevBuf s, a, b, o, k;
    TRY(s.alloc(n_checks * m * sizeof(G1))); TRY(a.alloc(n_checks * v * sizeof(G1))); TRY(b.alloc(n_checks * v * sizeof(G2)));
    TRY(o.alloc(n_checks * sizeof(Fq12))); TRY(k.alloc(n_checks));
    TRY(h2d(ctx, s.p, g1_shared, n_checks * m * sizeof(G1)));
    TRY(h2d(ctx, a.p, g1_var, n_checks * v * sizeof(G1)));
    TRY(h2d(ctx, b.p, g2_var, n_checks * v * sizeof(G2)));
    TRY(bn254_pairing_product_precomp_many_dev(ctx, pc, s.p, a.p, b.p, v, n_checks, o.p, k.as<uint8_t>()));
    if (gt_out) TRY(d2h(ctx, gt_out, o.p, n_checks * sizeof(Fq12)));
    if (is_one) TRY(d2h(ctx, is_one, k.p, n_checks));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// one product sharded over several devices: 384-byte partials gathered by peer copy (no NCCL)
// ---------------------------------------------------------------------------------------------
int bn254_pairing_product_sharded(bn254_ctx** ctxs, size_t n_ctx, const void* g1, const void* g2, size_t n, void* gt_out,
                                  uint8_t* is_one) {
    if (!ctxs || n_ctx == 0) return fail(BN254_ERR_ARG, "no contexts");
    for (size_t d = 0; d < n_ctx; d++) if (!ctxs[d]) return fail(BN254_ERR_ARG, "null context");
    if (n && (!g1 || !g2)) return fail(BN254_ERR_ARG, "null pointer");
    struct Shard { DevBuf a, b, f, s; Fq12* res = nullptr; size_t lo = 0, cnt = 0; };
    std::vector<Shard> sh(n_ctx);
    bn254_ctx* c0 = ctxs[0];
    TRY(use(c0));
    DevBuf partials, scratch, out, flag;
    TRY(partials.alloc(n_ctx * sizeof(Fq12))); TRY(scratch.alloc(n_ctx * sizeof(Fq12)));
    TRY(out.alloc(sizeof(Fq12))); TRY(flag.alloc(1));
    // phase 1: every device reduces its index range to one Miller product (all launches asynchronous)
    for (size_t d = 0; d < n_ctx; d++) {
        bn254_ctx* c = ctxs[d];
        Shard& s = sh[d];
        s.lo = n * d / n_ctx;
        s.cnt = n * (d + 1) / n_ctx - s.lo;
        TRY(use(c));
        size_t cnt = s.cnt ? s.cnt : 1;
        TRY(s.a.alloc(cnt * sizeof(G1))); TRY(s.b.alloc(cnt * sizeof(G2)));
        TRY(s.f.alloc(cnt * sizeof(Fq12))); TRY(s.s.alloc(((cnt + 1) / 2) * sizeof(Fq12)));
        if (s.cnt == 0) {   // empty shard contributes one
            LAUNCH(c, (k_miller_product<1, 4>), 1, (const G1*)nullptr, (const G2*)nullptr, 0, (const G1*)nullptr,
                   (const uint4*)nullptr, 0, s.f.as<Fq12>(), (size_t)1);
            s.res = s.f.as<Fq12>();
            continue;
        }
        TRY(h2d(c, s.a.p, (const uint8_t*)g1 + s.lo * sizeof(G1), s.cnt * sizeof(G1)));
        TRY(h2d(c, s.b.p, (const uint8_t*)g2 + s.lo * sizeof(G2), s.cnt * sizeof(G2)));
        LAUNCH(c, k_miller_many, s.cnt, s.b.as<G2>(), s.a.as<G1>(), s.f.as<Fq12>(), (uint8_t*)nullptr, s.cnt);
        TRY(tree_product(c, s.f.as<Fq12>(), s.s.as<Fq12>(), s.cnt, &s.res));
    }
    // phase 2: gather the partials on device 0 (peer copy over NVLink; staged by the driver if no P2P)
    for (size_t d = 0; d < n_ctx; d++) {
        bn254_ctx* c = ctxs[d];
        TRY(use(c));
        CU(cudaMemcpyPeerAsync(partials.as<Fq12>() + d, c0->device, sh[d].res, c->device, sizeof(Fq12), c->stream));
        CU(cudaStreamSynchronize(c->stream));
    }
    // phase 3: <= n_ctx-1 Fq12 multiplications and ONE final exponentiation on device 0
    TRY(use(c0));
    Fq12* res = nullptr;
    TRY(tree_product(c0, partials.as<Fq12>(), scratch.as<Fq12>(), n_ctx, &res));
    LAUNCH(c0, k_final_exp_many, 1, (const Fq12*)res, out.as<Fq12>(), (uint8_t*)nullptr, flag.as<uint8_t>(), (size_t)1);
    if (gt_out) TRY(d2h(c0, gt_out, out.p, sizeof(Fq12)));
    if (is_one) TRY(d2h(c0, is_one, flag.p, 1));
    CU(cudaStreamSynchronize(c0->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// field-level test hook and roofline microbenchmarks
// ---------------------------------------------------------------------------------------------
int bn254_field_op_many(bn254_ctx* ctx, int op, const void* a, const void* b, void* out, size_t n) {
    TRY(use(ctx));
    if (n == 0) return 0;
    static const size_t unit_of[16] = {32, 32, 32, 64, 64, 192, 384, 384, 384, 32, 64, 192, 384, 384, 384, 192};
    if (op < 0 || op > 15 || !a || !out) return fail(BN254_ERR_ARG, "bad argument");
    bool binary = (op == 0 || op == 1 || op == 2 || op == 3 || op == 5 || op == 6 || op == 11 || op == 12 || op == 15);
    if (binary && !b) return fail(BN254_ERR_ARG, "null pointer");
    size_t bytes = unit_of[op] * n;
    DevBuf x, y, o;
    TRY(x.alloc(bytes)); TRY(y.alloc(bytes)); TRY(o.alloc(bytes));
    TRY(h2d(ctx, x.p, a, bytes));
    if (binary) TRY(h2d(ctx, y.p, b, bytes));
    if (op >= 11) {
        k_slot_op<<<grid_for(n), BN_BLOCK, BN_BLOCK * BN_SLOT_STRIDE, ctx->stream>>>(op, x.as<uint4>(), y.as<uint4>(), o.as<uint4>(), n);
        ctx->launches.fetch_add(1);
        CU(cudaGetLastError());
    } else {
        LAUNCH(ctx, k_field_op, n, op, x.as<uint4>(), y.as<uint4>(), o.as<uint4>(), n);
    }
    TRY(d2h(ctx, out, o.p, bytes));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int bn254_microbench(bn254_ctx* ctx, int kind, int grid, int block, int iters, double* ms, double* ops) {
    TRY(use(ctx));
    if (kind < 0 || kind > 2 || grid <= 0 || block <= 0 || block > 256 || iters <= 0 || !ms || !ops) return fail(BN254_ERR_ARG, "bad argument");
    DevBuf sink;
    TRY(sink.alloc(16));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
    for (int rep = 0; rep < 2; rep++) {   // first pass warms up
        CU(cudaEventRecord(e0, ctx->stream));
        if (kind == 0) k_microbench<0><<<grid, block, 0, ctx->stream>>>(sink.as<uint32_t>(), iters, 12345u);
        else if (kind == 1) k_microbench<1><<<grid, block, 0, ctx->stream>>>(sink.as<uint32_t>(), iters, 12345u);
        else k_microbench<2><<<grid, block, 0, ctx->stream>>>(sink.as<uint32_t>(), iters, 12345u);
        ctx->launches.fetch_add(1);
        CU(cudaGetLastError());
        CU(cudaEventRecord(e1, ctx->stream));
        CU(cudaEventSynchronize(e1));
    }
    float t = 0;
    CU(cudaEventElapsedTime(&t, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *ms = t;
    double threads = (double)grid * block;
    *ops = threads * iters * (kind == 2 ? 8.0 : 128.0);   // kinds 0/1: 16 x 8 mads per iteration; kind 2: 8 Fq muls
    return 0;
}

}  // extern "C"
