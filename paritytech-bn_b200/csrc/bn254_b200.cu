// C-ABI implementation (include/bn254_b200.h): contexts, device arenas, kernel launches.
// Every entry point launches CUDA kernels; there is no CPU arithmetic path in this library.
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "../../include/bn254_b200.h"
#include "bn254_kernels.cuh"

using namespace bn;


static_assert(sizeof(Fq) == 32 && sizeof(Fq2) == 64 && sizeof(Fq12) == 384, "repr(C) sizes");
static_assert(sizeof(G1) == 96 && sizeof(G2) == 192 && sizeof(EllCoeffs) == 192, "repr(C) sizes");
static_assert(sizeof(AffG2) == 128 && sizeof(AffG1) == 64, "repr(C) sizes");

// thread-local error text in a fixed buffer: recording an error never allocates, so it cannot throw
static thread_local char g_err[320] = "";
static int fail(int code, const char* what, cudaError_t e = cudaSuccess) noexcept {
    if (e != cudaSuccess) snprintf(g_err, sizeof g_err, "%s: %s", what, cudaGetErrorString(e));
    else snprintf(g_err, sizeof g_err, "%s", what);
    return code;
}
#define CU(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return fail(BN254_ERR_CUDA, #call, _e); } while (0)
#define TRY(x) do { int _r = (x); if (_r != 0) return _r; } while (0)
// every extern "C" body runs inside this guard: no C++ exception crosses the boundary
#define API_BEGIN try {
#define API_END } catch (const std::bad_alloc&) { return fail(BN254_ERR_NOMEM, "out of host memory"); } \
                  catch (...) { return fail(BN254_ERR_CUDA, "unexpected C++ exception"); }

// grow-only device arena (stream-ordered reuse: every user enqueues on the owning context's streams)
struct Arena {
    void* p = nullptr;
    size_t bytes = 0;
};

#define BN_MAX_LANES 8   /* part batches of the pipelined host-buffer pairing path */
struct bn254_ctx {
    int device = 0;
    int sms = 148;     // multiprocessor count (the phase-synchronised kernels run one block per SM)
    cudaStream_t stream = nullptr;    // all work of the context; outranks stream2
    cudaStream_t stream2 = nullptr;   // second lane of the pipelined host-buffer pairing path
    cudaStream_t stream34[BN_MAX_LANES - 2] = {};   // lanes 3 .. 8 (pipeline = 3 .. 8)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_done = nullptr;
    cudaEvent_t ev_join34[BN_MAX_LANES - 2] = {};
    std::atomic<uint64_t> launches{0};
    Arena io;          // device copies of the host-pointer entry points' arguments and results
    Arena ws;          // kernel workspace: line-coefficient tables of the split Miller path, easy-part values
    Arena aux;         // temporaries of the device-pointer entry points (NULL outputs, > 8 pairs per check, partials)
    bn254_ctx_options opt;   // resolved switches (no -1 left)
    std::recursive_mutex mu; // every entry point holds it: one context serialises its callers
};
// worker threads of the device-list entry points: joined on every exit path (a std::thread destroyed while joinable terminates
// the process; creating the k-th thread can throw after k - 1 are running)
struct Workers {
    std::vector<std::thread> th;
    ~Workers() { join(); }
    void join() { for (auto& t : th) if (t.joinable()) t.join(); }
};
struct bn254_g2precomp {
    int device;
    size_t n;
    uint4* tab;     // [(c*12 + w) * n + j]
    AffG2* q_aff;   // [n]
};

// Where one batch runs: a stream, a workspace region and the block size of the phase-synchronised kernels.  The pipelined
// host path runs two lanes of one context side by side; nothing in the context is modified per call.
struct Lane {
    bn254_ctx* ctx;
    cudaStream_t stream;
    void* ws;             // non-null: fixed workspace region of ws_bytes; null: the context's grow-only ws arena
    size_t ws_bytes;
    unsigned force_block; // non-zero: block size of the phase-synchronised kernels regardless of the batch
};
static inline Lane main_lane(bn254_ctx* ctx) { return Lane{ctx, ctx->stream, nullptr, 0, 0}; }

static inline unsigned grid_for(size_t n) { return (unsigned)((n + BN_BLOCK - 1) / BN_BLOCK); }
#define LAUNCH(L, kernel, n, ...)                                                          \
    do {                                                                                   \
        if ((n) > 0) {                                                                     \
            kernel<<<grid_for(n), BN_BLOCK, 0, (L).stream>>>(__VA_ARGS__);                 \
            (L).ctx->launches.fetch_add(1);                                                \
            CU(cudaGetLastError());                                                        \
        }                                                                                  \
    } while (0)
#define LAUNCH_SYNC(L, kernel, grid, block, smem, ...)                                     \
    do {                                                                                   \
        kernel<<<(grid), (block), (smem), (L).stream>>>(__VA_ARGS__);                      \
        (L).ctx->launches.fetch_add(1);                                                    \
        CU(cudaGetLastError());                                                            \
    } while (0)

// Block size of the one-block-per-SM kernels: the batch spread over all SMs in whole warps, at most BN_FEXP_BLOCK threads
// (2^16 pairings on 148 SMs: 448 threads, 147 blocks; 2^15: 224 threads; small batches: one warp per block).
static inline unsigned sync_block(const Lane& L, size_t n) {
    const bn254_ctx* ctx = L.ctx;
    if (L.force_block) return L.force_block;
    // more than one wave of 448-thread blocks: 512-thread blocks on 6 slots per thread -- four warps on every sub-partition
    // instead of 4 + 4 + 3 + 3 (DESIGN F14) -- when that saves waves: a 512-thread wave takes 1.10 x a 448-thread one
    // (measured at 2^19: 23.5 against 21.4 ms for the whole path), so 2^19 (7 against 8 waves) and 2^20 (14 against 16)
    // gain 4 %, 2^17 and 2^18 (same wave count) would lose
    if (ctx->opt.big_blocks == 2 && n > BN_SYNC_MAX_BLOCK) return BN_SYNC_MAX_BLOCK;
    if (ctx->opt.big_blocks && n > (size_t)ctx->sms * BN_FEXP_BLOCK) {
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
static inline bool misaligned(const void* p) { return ((uintptr_t)p & 15u) != 0; }
#define CHECK_ALIGNED(...)                                                                                   \
    do {                                                                                                     \
        const void* _ps[] = {__VA_ARGS__};                                                                   \
        for (const void* _p : _ps) if (misaligned(_p)) return fail(BN254_ERR_ARG, "device pointer not 16-byte aligned"); \
    } while (0)
// device-side alias of a pinned, mapped host buffer (cudaHostAlloc / cudaHostRegister under UVA) whose registration covers
// [host, host + bytes): both ends must be mapped and alias one contiguous device range; nullptr for pageable memory
static void* mapped_device_ptr(const void* host, size_t bytes) {
    if (!host || bytes == 0) return nullptr;
    cudaPointerAttributes a, b;
    if (cudaPointerGetAttributes(&a, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (a.type != cudaMemoryTypeHost || !a.devicePointer) return nullptr;
    if (cudaPointerGetAttributes(&b, (const char*)host + bytes - 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (b.type != cudaMemoryTypeHost || (char*)b.devicePointer != (char*)a.devicePointer + (bytes - 1)) return nullptr;
    return a.devicePointer;
}
// Host <-> device copies.  CUDA refuses ONE copy whose host range is partly page-locked (cudaHostRegister of a part of the
// buffer) and partly pageable with cudaErrorInvalidValue; such a range is split in halves until every piece lies on one side of
// the registration boundary, which CUDA tracks to the byte (cuts are page aligned while the piece is larger than a page; at most
// ~log2(bytes) refused calls).
static int copy_split(const Lane& L, char* dst, const char* src, size_t bytes, cudaMemcpyKind kind) {
    if (bytes == 0) return 0;
    cudaError_t e = cudaMemcpyAsync(dst, src, bytes, kind, L.stream);
    if (e == cudaSuccess) return 0;
    if (e != cudaErrorInvalidValue || bytes <= 1) return fail(BN254_ERR_CUDA, "cudaMemcpyAsync", e);
    cudaGetLastError();
    const char* host = kind == cudaMemcpyHostToDevice ? src : dst;
    size_t half = bytes / 2;
    size_t cut = half;
    if (bytes > 8192) {   // the first piece ends on a page boundary
        cut = (((uintptr_t)host + half + 4095) & ~(uintptr_t)4095) - (uintptr_t)host;
        if (cut == 0 || cut >= bytes) cut = half;
    }
    TRY(copy_split(L, dst, src, cut, kind));
    return copy_split(L, dst + cut, src + cut, bytes - cut, kind);
}
static int h2d(const Lane& L, void* d, const void* h, size_t bytes) { return copy_split(L, (char*)d, (const char*)h, bytes, cudaMemcpyHostToDevice); }
static int d2h(const Lane& L, void* h, const void* d, size_t bytes) { return copy_split(L, (char*)h, (const char*)d, bytes, cudaMemcpyDeviceToHost); }
// grow-only arena: a larger request than any before synchronises the context's streams once and reallocates
static int arena_reserve(bn254_ctx* ctx, Arena& a, size_t bytes, const char* what) {
    if (bytes <= a.bytes) return 0;
    if (a.p) {
        CU(cudaStreamSynchronize(ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream2));
        for (int k = 0; k < BN_MAX_LANES - 2; k++) if (ctx->stream34[k]) CU(cudaStreamSynchronize(ctx->stream34[k]));
        cudaFree(a.p);
        a.p = nullptr; a.bytes = 0;
    }
    cudaError_t e = cudaMalloc(&a.p, bytes);
    if (e != cudaSuccess) { a.p = nullptr; return fail(BN254_ERR_NOMEM, what, e); }
    a.bytes = bytes;
    return 0;
}
// carve `count` 256-byte-aligned regions out of an arena
static int arena_carve(bn254_ctx* ctx, Arena& a, const size_t* sizes, int count, void** out, const char* what) {
    size_t total = 0;
    for (int i = 0; i < count; i++) total += (sizes[i] + 255) & ~(size_t)255;
    if (total == 0) total = 256;
    TRY(arena_reserve(ctx, a, total, what));
    char* p = (char*)a.p;
    for (int i = 0; i < count; i++) { out[i] = p; p += (sizes[i] + 255) & ~(size_t)255; }
    return 0;
}
static int ctx_io(bn254_ctx* ctx, const size_t* sizes, int count, void** out) { return arena_carve(ctx, ctx->io, sizes, count, out, "cudaMalloc I/O arena"); }
static int ctx_aux(bn254_ctx* ctx, const size_t* sizes, int count, void** out) { return arena_carve(ctx, ctx->aux, sizes, count, out, "cudaMalloc aux arena"); }
static int lane_workspace(const Lane& L, size_t bytes, void** out) {
    if (L.ws) {
        if (bytes > L.ws_bytes) return fail(BN254_ERR_ARG, "internal: workspace region too small");
        *out = L.ws;
        return 0;
    }
    TRY(arena_reserve(L.ctx, L.ctx->ws, bytes, "cudaMalloc workspace"));
    *out = L.ctx->ws.p;
    return 0;
}

static int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
}

extern "C" {

const char* bn254_last_error(void) { return g_err; }
const char* bn254_build_info(void) {
    static char buf[160];
    snprintf(buf, sizeof buf, "bn254_b200 sm_100a block=%d min_blocks=%d cuda=%d", BN_BLOCK, BN_MIN_BLOCKS, CUDART_VERSION);
    return buf;
}

void bn254_ctx_options_default(bn254_ctx_options* o) {
    if (!o) return;
    o->struct_size = (uint32_t)sizeof(bn254_ctx_options);
    o->pipeline = env_int("BN254_PIPELINE", 4);
    o->big_blocks = env_int("BN254_BIG_BLOCKS", 1);
    o->fused_miller = env_int("BN254_FUSED_MILLER", 0);
    o->split_fexp = env_int("BN254_SPLIT_FEXP", 0);
    o->fexp_smem = env_int("BN254_FEXP_SMEM", 2);
    o->miller_smem = env_int("BN254_MILLER_SMEM", 1);
    o->lines_smem = env_int("BN254_LINES_SMEM", 1);
    o->subgroup_ref = env_int("BN254_SUBGROUP_REF", 0);
    o->zerocopy_out = env_int("BN254_ZEROCOPY_OUT", 1);
    o->coop_max_n = env_int("BN254_COOP_MAX_N", BN_COOP_DEFAULT_MAX_N);
    o->stack_bytes = env_int("BN254_STACK_BYTES", 0);
    o->mul_compact = env_int("BN254_MUL_COMPACT", 1);
    o->coop4_max_n = env_int("BN254_COOP4_MAX_N", BN_COOP4_DEFAULT_MAX_N);
}

void bn254_ctx_destroy(bn254_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) { cudaStreamSynchronize(ctx->stream); cudaStreamDestroy(ctx->stream); }
    if (ctx->stream2) { cudaStreamSynchronize(ctx->stream2); cudaStreamDestroy(ctx->stream2); }
    for (int k = 0; k < BN_MAX_LANES - 2; k++) {
        if (ctx->stream34[k]) { cudaStreamSynchronize(ctx->stream34[k]); cudaStreamDestroy(ctx->stream34[k]); }
        if (ctx->ev_join34[k]) cudaEventDestroy(ctx->ev_join34[k]);
    }
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    if (ctx->ev_done) cudaEventDestroy(ctx->ev_done);
    if (ctx->ws.p) cudaFree(ctx->ws.p);
    if (ctx->io.p) cudaFree(ctx->io.p);
    if (ctx->aux.p) cudaFree(ctx->aux.p);
    delete ctx;
}

static void noalloc_setup(int device);   // no-alloc entry point (end of this file): resolves its static scratch, never fails the context
static int ctx_init(bn254_ctx* c, int device, const bn254_ctx_options* user) {
    c->device = device;
    bn254_ctx_options_default(&c->opt);
    if (user) {
        // field-wise merge up to the size the caller's header knew; -1 keeps the default
        const int32_t* src = &user->pipeline;
        int32_t* dst = &c->opt.pipeline;
        size_t known = user->struct_size >= sizeof(uint32_t) ? (user->struct_size - sizeof(uint32_t)) / sizeof(int32_t) : 0;
        size_t have = (sizeof(bn254_ctx_options) - sizeof(uint32_t)) / sizeof(int32_t);
        for (size_t i = 0; i < known && i < have; i++) if (src[i] >= 0) dst[i] = src[i];
    }
    cudaDeviceGetAttribute(&c->sms, cudaDevAttrMultiProcessorCount, device);
    if (c->sms <= 0) c->sms = 148;
    const int smem7 = BN_FEXP_BLOCK * BN_SLOT_STRIDE;
    if (c->opt.lines_smem) CU(cudaFuncSetAttribute(k_g2_lines_sync, cudaFuncAttributeMaxDynamicSharedMemorySize, smem7));
    if (c->opt.miller_smem) {
        CU(cudaFuncSetAttribute(k_miller_table_sync, cudaFuncAttributeMaxDynamicSharedMemorySize, smem7));
        CU(cudaFuncSetAttribute(k_miller_product_sync<1, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem7));
        CU(cudaFuncSetAttribute(k_miller_product_sync<4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem7));
        CU(cudaFuncSetAttribute(k_miller_product_sync<8, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem7));
    }
    if (c->opt.fexp_smem == 1) {
        cudaFuncSetAttribute(k_final_exp_hard_s, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    } else if (c->opt.fexp_smem == 2) {
        CU(cudaFuncSetAttribute(k_final_exp_hard_sync, cudaFuncAttributeMaxDynamicSharedMemorySize, smem7));
    }
    CU(cudaFuncSetAttribute(k_coop_pairing, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BN_COOP_SMEM(BN_COOP_MAX_WARPS)));
    CU(cudaFuncSetAttribute(k_jac_mul_compact<FqOps>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CmulLayout<FqOps>::BYTES));
    CU(cudaFuncSetAttribute(k_jac_mul_compact<Fq2Ops>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CmulLayout<Fq2Ops>::BYTES));
    CU(cudaFuncSetAttribute(k_coop_gt_pow, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BN_COOP_SMEM(BN_COOP_MAX_WARPS)));
    CU(cudaFuncSetAttribute(k_coop4_pairing, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BN_COOP4_SMEM(BN_COOP4_MAX_GROUPS)));
    CU(cudaFuncSetAttribute(k_coop4_gt_pow, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BN_COOP4_SMEM(BN_COOP4_MAX_GROUPS)));
    // the main stream outranks the second one: in the pipelined path the first half batch gets ahead of the second
    int prio_least = 0, prio_greatest = 0;
    cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
    CU(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_greatest));
    CU(cudaStreamCreateWithPriority(&c->stream2, cudaStreamNonBlocking, prio_least));
    for (int k = 0; k < BN_MAX_LANES - 2; k++) {
        CU(cudaStreamCreateWithPriority(&c->stream34[k], cudaStreamNonBlocking, prio_least));
        CU(cudaEventCreateWithFlags(&c->ev_join34[k], cudaEventDisableTiming));
    }
    CU(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming));
    noalloc_setup(device);
    // The call graph is static (no recursion, no indirect calls; largest frame 8 KB in ptxas.log), so the driver sizes the
    // stack of every launch itself.  Only an explicit request touches the device-wide limit, and only upwards.
    if (c->opt.stack_bytes > 0) {
        size_t cur = 0;
        CU(cudaDeviceGetLimit(&cur, cudaLimitStackSize));
        if (cur < (size_t)c->opt.stack_bytes) CU(cudaDeviceSetLimit(cudaLimitStackSize, (size_t)c->opt.stack_bytes));
    }
    return 0;
}
int bn254_ctx_create_ex(int device, const bn254_ctx_options* opt, bn254_ctx** out) {
    API_BEGIN
    if (!out) return fail(BN254_ERR_ARG, "null out");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return fail(BN254_ERR_CUDA, "no CUDA device (this library has no CPU fallback)", e);
    if (device < 0 || device >= count) return fail(BN254_ERR_ARG, "bad device index");
    CU(cudaSetDevice(device));
    bn254_ctx* c = new (std::nothrow) bn254_ctx();
    if (!c) return fail(BN254_ERR_NOMEM, "ctx alloc");
    int rc = ctx_init(c, device, opt);
    if (rc != 0) { bn254_ctx_destroy(c); return rc; }   // releases whatever was created (streams, events)
    *out = c;
    return 0;
    API_END
}
int bn254_ctx_create(int device, bn254_ctx** out) { return bn254_ctx_create_ex(device, nullptr, out); }
int bn254_ctx_device(const bn254_ctx* ctx) { return ctx ? ctx->device : -1; }
void* bn254_ctx_stream(const bn254_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
int bn254_ctx_synchronize(bn254_ctx* ctx) {
    API_BEGIN
    TRY(use(ctx));
    std::lock_guard<std::recursive_mutex> g(ctx->mu);
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
    API_END
}
uint64_t bn254_ctx_launch_count(const bn254_ctx* ctx) { return ctx ? ctx->launches.load() : 0; }

}  // extern "C"

#define ENTER(ctx) TRY(use(ctx)); std::lock_guard<std::recursive_mutex> _guard((ctx)->mu)

// Warp-cooperative latency path (bn_coop.cuh): one warp per pairing, for batches too small to fill the machine with one
// thread per pairing.  Same bytes as the throughput path.
static inline bool use_coop(const bn254_ctx* ctx, size_t n) { return ctx->opt.coop_max_n > 0 && n <= (size_t)ctx->opt.coop_max_n; }
// mode 0: pairing, 1: Miller value (+ ok flags), 2: final exponentiation (g1 = the Gt input; ok = Option flags)
// quad-lane executor (bn_coop4.cuh): four warps per unit, for batches of at most coop4_max_n units (a subset of the above)
static inline bool use_coop4(const bn254_ctx* ctx, size_t n) { return use_coop(ctx, n) && n <= (size_t)(ctx->opt.coop4_max_n > 0 ? ctx->opt.coop4_max_n : 0); }
// units per block and blocks of a quad-lane launch: the batch spread over all SMs, at most BN_COOP4_MAX_GROUPS per block
static inline void coop4_shape(const bn254_ctx* ctx, size_t n, unsigned& groups, unsigned& grid) {
    size_t per_sm = (n + (size_t)ctx->sms - 1) / (size_t)ctx->sms;
    groups = per_sm < 1 ? 1u : per_sm > BN_COOP4_MAX_GROUPS ? (unsigned)BN_COOP4_MAX_GROUPS : (unsigned)per_sm;
    grid = (unsigned)((n + groups - 1) / groups);
}
static int coop_pairing_dev(const Lane& L, const void* d_g1, const G2* d_g2, Fq12* d_out, uint8_t* d_ok, size_t n, int mode,
                            uint8_t* d_is_one = nullptr) {
    if (use_coop4(L.ctx, n)) {
        unsigned groups, g;
        coop4_shape(L.ctx, n, groups, g);
        LAUNCH_SYNC(L, k_coop4_pairing, g, groups * BN_COOP4_GROUP, BN_COOP4_SMEM(groups), d_g1, d_g2, d_out, d_ok, d_is_one, n, mode);
        return 0;
    }
    size_t per_sm = (n + (size_t)L.ctx->sms - 1) / (size_t)L.ctx->sms;
    unsigned warps = per_sm < 1 ? 1u : per_sm > BN_COOP_MAX_WARPS ? (unsigned)BN_COOP_MAX_WARPS : (unsigned)per_sm;
    unsigned g = (unsigned)((n + warps - 1) / warps);
    LAUNCH_SYNC(L, k_coop_pairing, g, warps * 32, BN_COOP_SMEM(warps), d_g1, d_g2, d_out, d_ok, d_is_one, n, mode);
    return 0;
}

// final exponentiation of n values (in place allowed): single kernel, or easy + hard kernels
static int final_exp_dev(const Lane& L, const Fq12* d_in, Fq12* d_out, uint8_t* d_some, uint8_t* d_is_one, size_t n) {
    const bn254_ctx_options& o = L.ctx->opt;
    if (use_coop(L.ctx, n)) return coop_pairing_dev(L, d_in, (const G2*)nullptr, d_out, d_some, n, 2, d_is_one);
    if (!o.split_fexp && !o.fexp_smem) {
        LAUNCH(L, k_final_exp_many, n, d_in, d_out, d_some, d_is_one, n);
        return 0;
    }
    // scratch: easy-part values + flags, placed after the Miller tables' use (stream-ordered, so reuse is safe)
    void* ws = nullptr;
    size_t vals = n * sizeof(Fq12);
    TRY(lane_workspace(L, vals + ((n + 15) & ~(size_t)15), &ws));
    Fq12* t = (Fq12*)ws;
    uint8_t* flags = (uint8_t*)ws + vals;
    LAUNCH(L, k_final_exp_easy, n, d_in, t, flags, n);
    if (o.fexp_smem == 2) {
        unsigned bs = sync_block(L, n), g = (unsigned)((n + bs - 1) / bs);
        LAUNCH_SYNC(L, k_final_exp_hard_sync, g, bs, sync_smem(bs), (const Fq12*)t, (const uint8_t*)flags, d_out, d_some, d_is_one, n);
        return 0;
    }
    if (o.fexp_smem) {
        LAUNCH_SYNC(L, k_final_exp_hard_s, grid_for(n), BN_BLOCK, BN_BLOCK * BN_SLOT_STRIDE, (const Fq12*)t, (const uint8_t*)flags, d_out, d_some, d_is_one, n);
        return 0;
    }
    LAUNCH(L, k_final_exp_hard, n, (const Fq12*)t, (const uint8_t*)flags, d_out, d_some, d_is_one, n);
    return 0;
}
// line tables of m affine G2 points (slot layout for v pairs per check): slot-file kernel or the local-memory kernel
static int lines_launch(const Lane& L, const AffG2* qa, const uint8_t* ok, uint4* tab, size_t m, size_t v) {
    // one wave only: with more work the 64-thread kernel reaches 16 warps per SM (the slot file allows 14) and is faster
    // (2^17 points: 5.6 against 6.0 ms; 2^16: 3.24 against 3.01 ms)
    if (L.ctx->opt.lines_smem && m <= (size_t)L.ctx->sms * BN_FEXP_BLOCK) {
        unsigned bs = sync_block(L, m);
        if (bs > BN_FEXP_BLOCK) bs = BN_FEXP_BLOCK;   // this kernel uses all 7 slots
        unsigned g = (unsigned)((m + bs - 1) / bs);
        LAUNCH_SYNC(L, k_g2_lines_sync, g, bs, sync_smem(bs), qa, ok, tab, m, v);
        return 0;
    }
    LAUNCH(L, k_g2_lines_affine, m, qa, ok, tab, m, v);
    return 0;
}
// Miller loop of m pairs from their line tables: slot-file kernel (one 448-thread block per SM) or the local-memory kernel
static int miller_table_launch(const Lane& L, const AffG1* pa, const uint8_t* ok, const uint4* tab, Fq12* f, uint8_t* ok_out, size_t m) {
    if (L.ctx->opt.miller_smem) {
        unsigned bs = sync_block(L, m), g = (unsigned)((m + bs - 1) / bs);
        LAUNCH_SYNC(L, k_miller_table_sync, g, bs, sync_smem(bs), pa, ok, tab, f, ok_out, m);
        return 0;
    }
    LAUNCH(L, k_miller_table_affine, m, pa, ok, tab, f, ok_out, m);
    return 0;
}
static const size_t MILLER_CHUNK = (size_t)1 << 17;   // pairings per table chunk (2.2 GB of line coefficients)
static size_t miller_ws_bytes(size_t n) {
    size_t chunk = n < MILLER_CHUNK ? n : MILLER_CHUNK;
    return (size_t)BN254_N_COEFFS * 12 * chunk * sizeof(uint4) + chunk * sizeof(AffG2) + chunk * sizeof(AffG1) + ((chunk + 15) & ~(size_t)15);
}
static size_t fexp_ws_bytes(size_t n) { return n * sizeof(Fq12) + ((n + 15) & ~(size_t)15); }
struct MillerWs { uint4* tab; AffG2* qa; AffG1* pa; uint8_t* ok; size_t chunk; };
static int miller_ws(const Lane& L, size_t n, MillerWs& w) {
    w.chunk = n < MILLER_CHUNK ? n : MILLER_CHUNK;
    size_t tab_bytes = (size_t)BN254_N_COEFFS * 12 * w.chunk * sizeof(uint4);
    size_t qa_off = tab_bytes, pa_off = qa_off + w.chunk * sizeof(AffG2), ok_off = pa_off + w.chunk * sizeof(AffG1);
    void* ws = nullptr;
    TRY(lane_workspace(L, ok_off + ((w.chunk + 15) & ~(size_t)15), &ws));
    w.tab = (uint4*)ws;
    w.qa = (AffG2*)((char*)ws + qa_off);
    w.pa = (AffG1*)((char*)ws + pa_off);
    w.ok = (uint8_t*)ws + ok_off;
    return 0;
}
// Miller values of n independent pairs (fused single kernel, or affine -> line table -> Miller from table).
// ev (optional, 4 events): recorded before each of the three kernels and after the last one of the LAST chunk; ms (optional)
// accumulates the per-kernel times of every chunk (synchronises per chunk: profiling twin only).
static int miller_many_dev(const Lane& L, const G2* d_g2, const G1* d_g1, Fq12* d_f, uint8_t* d_ok, size_t n,
                           cudaEvent_t* ev = nullptr, float* ms = nullptr) {
    if (!ev && use_coop(L.ctx, n)) return coop_pairing_dev(L, d_g1, d_g2, d_f, d_ok, n, 1);
    if (L.ctx->opt.fused_miller) {
        if (ev) CU(cudaEventRecord(ev[0], L.stream));
        LAUNCH(L, k_miller_many, n, d_g2, d_g1, d_f, d_ok, n);
        if (ev) {
            CU(cudaEventRecord(ev[1], L.stream));
            CU(cudaEventSynchronize(ev[1]));
            float t; CU(cudaEventElapsedTime(&t, ev[0], ev[1])); ms[2] += t;
        }
        return 0;
    }
    MillerWs w;
    TRY(miller_ws(L, n, w));
    for (size_t lo = 0; lo < n; lo += w.chunk) {
        size_t m = n - lo < w.chunk ? n - lo : w.chunk;
        if (ev) CU(cudaEventRecord(ev[0], L.stream));
        LAUNCH(L, k_pairs_to_affine, m, d_g2 + lo, d_g1 + lo, w.qa, w.pa, w.ok, m);
        if (ev) CU(cudaEventRecord(ev[1], L.stream));
        TRY(lines_launch(L, (const AffG2*)w.qa, (const uint8_t*)w.ok, w.tab, m, (size_t)1));
        if (ev) CU(cudaEventRecord(ev[2], L.stream));
        TRY(miller_table_launch(L, (const AffG1*)w.pa, (const uint8_t*)w.ok, (const uint4*)w.tab, d_f + lo, d_ok ? d_ok + lo : (uint8_t*)nullptr, m));
        if (ev) {
            CU(cudaEventRecord(ev[3], L.stream));
            CU(cudaEventSynchronize(ev[3]));
            for (int i = 0; i < 3; i++) { float t; CU(cudaEventElapsedTime(&t, ev[i], ev[i + 1])); ms[i] += t; }
        }
    }
    return 0;
}
// reduce n Fq12 values in `a` to one product; `b` is scratch of >= (n+1)/2 elements.  Result pointer returned.
static int tree_product(const Lane& L, Fq12* a, Fq12* b, size_t n, Fq12** result) {
    Fq12 *src = a, *dst = b;
    while (n > 1) {
        size_t n_out = (n + 1) / 2;
        LAUNCH(L, k_fq12_pair_product, n_out, (const Fq12*)src, n, dst);
        n = n_out;
        Fq12* t = src; src = dst; dst = t;
    }
    *result = src;
    return 0;
}
// un-exponentiated product of the Miller values of n pairs (pairing_batch skip semantics): d_f (n values) and d_s ((n+1)/2
// values) are scratch; the result lands in *res (one of the two).  n = 0 writes one into d_f[0].
static int miller_product_dev(const Lane& L, const G1* d_g1, const G2* d_g2, size_t n, Fq12* d_f, Fq12* d_s, uint8_t* d_ok, Fq12** res) {
    if (n == 0) {
        LAUNCH(L, k_fq12_set_one, 1, d_f, (size_t)1);
        *res = d_f;
        return 0;
    }
    TRY(miller_many_dev(L, d_g2, d_g1, d_f, d_ok, n));
    TRY(tree_product(L, d_f, d_s, n, res));
    return 0;
}

extern "C" {

// ---------------------------------------------------------------------------------------------
// pairing
// ---------------------------------------------------------------------------------------------
int bn254_pairing_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_g2, void* d_gt_out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    if (!d_g1 || !d_g2 || !d_gt_out) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_g1, d_g2, d_gt_out);
    Lane L = main_lane(ctx);
    if (use_coop(ctx, n)) return coop_pairing_dev(L, d_g1, (const G2*)d_g2, (Fq12*)d_gt_out, (uint8_t*)nullptr, n, 0);
    // Miller values are written to the output buffer and exponentiated in place
    TRY(miller_many_dev(L, (const G2*)d_g2, (const G1*)d_g1, (Fq12*)d_gt_out, (uint8_t*)nullptr, n));
    TRY(final_exp_dev(L, (const Fq12*)d_gt_out, (Fq12*)d_gt_out, (uint8_t*)nullptr, (uint8_t*)nullptr, n));
    return 0;
    API_END
}
int bn254_pairing_many_dev_timed(bn254_ctx* ctx, const void* d_g1, const void* d_g2, void* d_gt_out, size_t n, float* ms) {
    API_BEGIN
    ENTER(ctx);
    if (!ms) return fail(BN254_ERR_ARG, "null pointer");
    for (int i = 0; i < 5; i++) ms[i] = 0.f;
    if (n == 0) return 0;
    if (!d_g1 || !d_g2 || !d_gt_out) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_g1, d_g2, d_gt_out);
    cudaEvent_t ev[6];
    for (int i = 0; i < 6; i++) CU(cudaEventCreate(&ev[i]));
    Lane L = main_lane(ctx);
    Fq12* f = (Fq12*)d_gt_out;
    auto body = [&]() -> int {
        TRY(miller_many_dev(L, (const G2*)d_g2, (const G1*)d_g1, f, (uint8_t*)nullptr, n, ev, ms));
        // the final exponentiation with an event between its easy and hard kernels (same launches as final_exp_dev)
        const bn254_ctx_options& o = ctx->opt;
        CU(cudaEventRecord(ev[3], L.stream));
        if (o.fexp_smem == 2) {
            void* ws = nullptr;
            size_t vals = n * sizeof(Fq12);
            TRY(lane_workspace(L, vals + ((n + 15) & ~(size_t)15), &ws));
            Fq12* t = (Fq12*)ws;
            uint8_t* flags = (uint8_t*)ws + vals;
            LAUNCH(L, k_final_exp_easy, n, (const Fq12*)f, t, flags, n);
            CU(cudaEventRecord(ev[4], L.stream));
            unsigned bs = sync_block(L, n), g = (unsigned)((n + bs - 1) / bs);
            LAUNCH_SYNC(L, k_final_exp_hard_sync, g, bs, sync_smem(bs), (const Fq12*)t, (const uint8_t*)flags, f, (uint8_t*)nullptr, (uint8_t*)nullptr, n);
        } else {
            CU(cudaEventRecord(ev[4], L.stream));
            TRY(final_exp_dev(L, (const Fq12*)f, f, (uint8_t*)nullptr, (uint8_t*)nullptr, n));
        }
        CU(cudaEventRecord(ev[5], L.stream));
        CU(cudaEventSynchronize(ev[5]));
        float t;
        CU(cudaEventElapsedTime(&t, ev[3], ev[4])); ms[3] = t;
        CU(cudaEventElapsedTime(&t, ev[4], ev[5])); ms[4] = t;
        return 0;
    };
    int rc = body();
    for (int i = 0; i < 6; i++) cudaEventDestroy(ev[i]);
    return rc;
    API_END
}
}  // extern "C"

// one (part of a) batch of the host-buffer pairing path on lane L: copies in, kernels, results out
static int pairing_many_enqueue(const Lane& L, const char* g1, const char* g2, char* gt_out, char* mapped, void* const* d, size_t lo, size_t m) {
    G1* dg1 = (G1*)d[0] + lo;
    G2* dg2 = (G2*)d[1] + lo;
    Fq12* df = (Fq12*)d[2] + lo;
    TRY(h2d(L, dg1, g1 + lo * sizeof(G1), m * sizeof(G1)));
    TRY(h2d(L, dg2, g2 + lo * sizeof(G2), m * sizeof(G2)));
    if (use_coop(L.ctx, m)) {
        TRY(coop_pairing_dev(L, dg1, dg2, mapped ? (Fq12*)mapped + lo : df, (uint8_t*)nullptr, m, 0));
        if (!mapped) TRY(d2h(L, gt_out + lo * sizeof(Fq12), df, m * sizeof(Fq12)));
        return 0;
    }
    TRY(miller_many_dev(L, dg2, dg1, df, (uint8_t*)nullptr, m));
    if (mapped) {
        TRY(final_exp_dev(L, df, (Fq12*)mapped + lo, (uint8_t*)nullptr, (uint8_t*)nullptr, m));
    } else {
        TRY(final_exp_dev(L, df, df, (uint8_t*)nullptr, (uint8_t*)nullptr, m));
        TRY(d2h(L, gt_out + lo * sizeof(Fq12), df, m * sizeof(Fq12)));
    }
    return 0;
}
// the host-buffer pairing path of one context (caller holds the context)
static int pairing_many_host(bn254_ctx* ctx, const void* g1, const void* g2, void* gt_out, size_t n) {
    if (n == 0) return 0;
    if (!g1 || !g2 || !gt_out) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[3] = {n * sizeof(G1), n * sizeof(G2), n * sizeof(Fq12)};
    void* d[3];
    TRY(ctx_io(ctx, sz, 3, d));
    // Pinned (mapped) output buffer: the final-exponentiation kernel stores its 384-byte results straight into host memory.
    char* mapped = ctx->opt.zerocopy_out ? (char*)mapped_device_ptr(gt_out, n * sizeof(Fq12)) : nullptr;
    Lane A = main_lane(ctx);
    const size_t PIPE_MIN = (size_t)1 << 16;   // below this the halves would leave SMs idle (74 blocks of 448 threads per half at 2^16)
    if (!ctx->opt.pipeline || n < PIPE_MIN || ctx->opt.fused_miller) {
        TRY(pairing_many_enqueue(A, (const char*)g1, (const char*)g2, (char*)gt_out, mapped, d, 0, n));
        CU(cudaStreamSynchronize(ctx->stream));
        return 0;
    }
    // K part batches on K lanes (pipeline = 2, 3 or 4; default 4: +0.3 % at 2^16, +1.8 % at 2^18, +0.8 % at N = 8 over two).  The one-wave kernels of the slot-file path finish all their
    // blocks together, so a single batch exposes both transfers completely (inputs before the first kernel, results after the
    // last one).  With K parts the later parts' inputs travel while the first computes; each part's 448-thread kernels occupy
    // 1/K of the SMs.  Every part but the last is a whole number of blocks.
    int K = ctx->opt.pipeline < 2 ? 2 : ctx->opt.pipeline > BN_MAX_LANES ? BN_MAX_LANES : ctx->opt.pipeline;
    // Beyond one wave a part must itself be a full wave of 448-thread blocks: smaller parts run as 224-thread blocks, two kernels
    // of different lanes share every SM and evict each other's code (2^17 pairs: 45.4 ms on four lanes against 42.3 ms on two,
    // scripts/bench_single_call_lanes.py).  Within one wave (block size forced below) the parts take disjoint SMs.
    const size_t wave = (size_t)ctx->sms * BN_FEXP_BLOCK;
    if (n > wave) {
        const size_t whole = n / ((size_t)1 << 16);
        if ((size_t)K > whole) K = whole < 2 ? 2 : (int)whole;
    }
    size_t part = (((n + K - 1) / K + BN_FEXP_BLOCK - 1) / BN_FEXP_BLOCK) * BN_FEXP_BLOCK;
    size_t lo[BN_MAX_LANES], cnt[BN_MAX_LANES], woff[BN_MAX_LANES], wsz[BN_MAX_LANES], wtot = 0;
    int parts = 0;
    for (size_t at = 0; at < n && parts < K; parts++) {
        size_t m = (parts == K - 1 || n - at < part) ? n - at : part;
        lo[parts] = at; cnt[parts] = m; at += m;
        size_t w = miller_ws_bytes(m);
        if (fexp_ws_bytes(m) > w) w = fexp_ws_bytes(m);
        w = (w + 255) & ~(size_t)255;
        woff[parts] = wtot; wsz[parts] = w; wtot += w;
    }
    TRY(arena_reserve(ctx, ctx->ws, wtot, "cudaMalloc workspace"));
    // one-wave batches: each part on its own SMs with full blocks (smaller blocks sharing an SM would interleave their phases);
    // larger batches fill every SM with one block at a time anyway and keep the block size rule of sync_block
    unsigned fb = n <= (size_t)ctx->sms * BN_FEXP_BLOCK ? BN_FEXP_BLOCK : 0;
    cudaStream_t lanes[BN_MAX_LANES] = {ctx->stream, ctx->stream2};
    cudaEvent_t joins[BN_MAX_LANES] = {nullptr, ctx->ev_join};
    for (int k = 2; k < BN_MAX_LANES; k++) { lanes[k] = ctx->stream34[k - 2]; joins[k] = ctx->ev_join34[k - 2]; }
    CU(cudaEventRecord(ctx->ev_fork, ctx->stream));
    for (int k = 1; k < parts; k++) CU(cudaStreamWaitEvent(lanes[k], ctx->ev_fork, 0));
    int rc = 0;
    for (int k = 0; k < parts && rc == 0; k++) {
        Lane Lk{ctx, lanes[k], (char*)ctx->ws.p + woff[k], wsz[k], fb};
        rc = pairing_many_enqueue(Lk, (const char*)g1, (const char*)g2, (char*)gt_out, mapped, d, lo[k], cnt[k]);
    }
    for (int k = 1; k < parts; k++) {
        cudaEventRecord(joins[k], lanes[k]);
        cudaStreamWaitEvent(ctx->stream, joins[k], 0);
    }
    cudaError_t es = cudaStreamSynchronize(ctx->stream);
    if (rc != 0) return rc;
    if (es != cudaSuccess) return fail(BN254_ERR_CUDA, "cudaStreamSynchronize", es);
    return 0;
}

extern "C" {
int bn254_pairing_many(bn254_ctx* ctx, const void* g1, const void* g2, void* gt_out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return pairing_many_host(ctx, g1, g2, gt_out, n);
    API_END
}

// n independent pairs over a device list: contiguous index ranges, one host thread per context
int bn254_pairing_many_sharded(bn254_ctx** ctxs, size_t n_ctx, const void* g1, const void* g2, void* gt_out, size_t n) {
    API_BEGIN
    if (!ctxs || n_ctx == 0) return fail(BN254_ERR_ARG, "no contexts");
    for (size_t d = 0; d < n_ctx; d++) if (!ctxs[d]) return fail(BN254_ERR_ARG, "null context");
    if (n == 0) return 0;
    if (!g1 || !g2 || !gt_out) return fail(BN254_ERR_ARG, "null pointer");
    if (n_ctx == 1) { ENTER(ctxs[0]); return pairing_many_host(ctxs[0], g1, g2, gt_out, n); }
    struct Result { int rc = 0; char err[320] = ""; };
    std::vector<Result> res(n_ctx);
    Workers workers;
    std::vector<std::thread>& th = workers.th;
    th.reserve(n_ctx);
    for (size_t d = 0; d < n_ctx; d++) {
        size_t lo = n * d / n_ctx, hi = n * (d + 1) / n_ctx;
        th.emplace_back([=, &res]() {
            bn254_ctx* c = ctxs[d];
            int rc = 0;
            try {
                rc = use(c);
                if (rc == 0) {
                    std::lock_guard<std::recursive_mutex> g(c->mu);
                    rc = pairing_many_host(c, (const char*)g1 + lo * sizeof(G1), (const char*)g2 + lo * sizeof(G2),
                                           (char*)gt_out + lo * sizeof(Fq12), hi - lo);
                }
            } catch (...) { rc = fail(BN254_ERR_NOMEM, "exception in a device worker"); }
            res[d].rc = rc;
            if (rc != 0) snprintf(res[d].err, sizeof res[d].err, "device %d: %s", c->device, g_err);   // g_err is this worker's
        });
    }
    workers.join();
    for (size_t d = 0; d < n_ctx; d++) if (res[d].rc != 0) return fail(res[d].rc, res[d].err);
    return 0;
    API_END
}

int bn254_miller_loop_many_dev(bn254_ctx* ctx, const void* d_g2, const void* d_g1, void* d_gt_out, uint8_t* d_ok, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    if (!d_g1 || !d_g2 || !d_gt_out) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_g1, d_g2, d_gt_out);
    TRY(miller_many_dev(main_lane(ctx), (const G2*)d_g2, (const G1*)d_g1, (Fq12*)d_gt_out, d_ok, n));
    return 0;
    API_END
}
int bn254_miller_loop_many(bn254_ctx* ctx, const void* g2, const void* g1, void* gt_out, uint8_t* ok, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    if (!g1 || !g2 || !gt_out) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[4] = {n * sizeof(G1), n * sizeof(G2), n * sizeof(Fq12), n};
    void* d[4];
    TRY(ctx_io(ctx, sz, 4, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], g1, sz[0]));
    TRY(h2d(L, d[1], g2, sz[1]));
    TRY(miller_many_dev(L, (const G2*)d[1], (const G1*)d[0], (Fq12*)d[2], (uint8_t*)d[3], n));
    TRY(d2h(L, gt_out, d[2], sz[2]));
    if (ok) TRY(d2h(L, ok, d[3], n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
    API_END
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// product checks (pairing_batch) and miller_loop_batch
// ---------------------------------------------------------------------------------------------
// product checks that run as independent Miller loops + a per-check product: more than 8 pairs per check, or a batch small
// enough for the warp-cooperative path (no shared tables)
static inline bool indep_products(const bn254_ctx* ctx, size_t v, size_t m, size_t n_checks) {
    return v > 8 || (m == 0 && v > 0 && use_coop(ctx, n_checks * v));
}
// Miller products of n_checks checks into d_f (n_checks x Fq12); shared tables optional
static int miller_products_dev(const Lane& L, const G1* d_g1v, const G2* d_g2v, size_t v, const G1* d_g1s,
                               const bn254_g2precomp* pc, size_t n_checks, Fq12* d_f) {
    bn254_ctx* ctx = L.ctx;
    size_t m = pc ? pc->n : 0;
    const uint4* tab = pc ? pc->tab : nullptr;
    if (m > 8) return fail(BN254_ERR_ARG, "at most 8 shared tables per check");
    if (indep_products(ctx, v, m, n_checks)) {
        // many pairs per check, or a batch small enough for the warp-cooperative path: independent Miller loops, then a
        // per-check product (same field element)
        if (m != 0) return fail(BN254_ERR_ARG, "more than 8 variable pairs cannot be combined with shared tables");
        size_t sz[1] = {n_checks * v * sizeof(Fq12)};
        void* tmp[1];
        TRY(ctx_aux(ctx, sz, 1, tmp));
        TRY(miller_many_dev(L, d_g2v, d_g1v, (Fq12*)tmp[0], (uint8_t*)nullptr, n_checks * v));
        LAUNCH(L, k_fq12_group_product, n_checks, (const Fq12*)tmp[0], v, d_f, n_checks);
        return 0;
    }
    if (ctx->opt.fused_miller || v == 0) {
        // single kernel: line coefficients of the variable G2 points produced on the fly
        if (v <= 1 && m <= 4) {
            LAUNCH(L, (k_miller_product<1, 4>), n_checks, d_g1v, d_g2v, (int)v, d_g1s, tab, (int)m, d_f, n_checks);
        } else if (v <= 4 && m <= 4) {
            LAUNCH(L, (k_miller_product<4, 4>), n_checks, d_g1v, d_g2v, (int)v, d_g1s, tab, (int)m, d_f, n_checks);
        } else {
            LAUNCH(L, (k_miller_product<8, 8>), n_checks, d_g1v, d_g2v, (int)v, d_g1s, tab, (int)m, d_f, n_checks);
        }
        return 0;
    }
    // split path: affine conversion and line tables of the variable pairs (slot layout of table_slot), then one
    // product kernel per chunk of checks reading per-check and shared tables
    // checks per chunk: at least 2^16 for up to 4 pairs per check (one full wave of the 448-thread kernel on 148 SMs; 4.4 GB of
    // line tables at v = 4), 2^15 above
    size_t chunk = MILLER_CHUNK / v;
    if (ctx->opt.miller_smem) { size_t want = v <= 4 ? ((size_t)1 << 16) : ((size_t)1 << 15); if (chunk < want) chunk = want; }
    if (chunk > n_checks) chunk = n_checks;
    size_t pairs = chunk * v;
    size_t tab_bytes = (size_t)BN254_N_COEFFS * 12 * pairs * sizeof(uint4);
    size_t qa_off = tab_bytes, pa_off = qa_off + pairs * sizeof(AffG2), ok_off = pa_off + pairs * sizeof(AffG1);
    void* ws = nullptr;
    TRY(lane_workspace(L, ok_off + ((pairs + 15) & ~(size_t)15), &ws));
    uint4* tabv = (uint4*)ws;
    AffG2* qa = (AffG2*)((char*)ws + qa_off);
    AffG1* pa = (AffG1*)((char*)ws + pa_off);
    uint8_t* ok = (uint8_t*)ws + ok_off;
    for (size_t lo = 0; lo < n_checks; lo += chunk) {
        size_t nc = n_checks - lo < chunk ? n_checks - lo : chunk;
        size_t np = nc * v;
        LAUNCH(L, k_pairs_to_affine, np, d_g2v + lo * v, d_g1v + lo * v, qa, pa, ok, np);
        TRY(lines_launch(L, (const AffG2*)qa, (const uint8_t*)ok, tabv, np, v));
        const G1* gs = d_g1s ? d_g1s + lo * m : nullptr;
        if (ctx->opt.miller_smem) {
            unsigned bs = sync_block(L, nc), g = (unsigned)((nc + bs - 1) / bs);
            size_t smem = sync_smem(bs);
            if (v <= 1 && m <= 4) LAUNCH_SYNC(L, (k_miller_product_sync<1, 4>), g, bs, smem, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
            else if (v <= 4 && m <= 4) LAUNCH_SYNC(L, (k_miller_product_sync<4, 4>), g, bs, smem, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
            else LAUNCH_SYNC(L, (k_miller_product_sync<8, 8>), g, bs, smem, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
        } else if (v <= 1 && m <= 4) {
            LAUNCH(L, (k_miller_product_tables<1, 4>), nc, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
        } else if (v <= 4 && m <= 4) {
            LAUNCH(L, (k_miller_product_tables<4, 4>), nc, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
        } else {
            LAUNCH(L, (k_miller_product_tables<8, 8>), nc, (const AffG1*)pa, (const uint8_t*)ok, (const uint4*)tabv, (int)v, gs, tab, (int)m, d_f + lo, nc);
        }
    }
    return 0;
}
// product checks on device pointers; d_gt_out may be NULL (then the Gt values go to an aux-arena temporary and the call
// synchronises before returning so that the next call cannot reuse it too early)
static int product_checks_dev(bn254_ctx* ctx, const bn254_g2precomp* pc, const void* d_g1s, const void* d_g1v, const void* d_g2v,
                              size_t v, size_t n_checks, void* d_gt_out, uint8_t* d_is_one) {
    Lane L = main_lane(ctx);
    Fq12* d_f = (Fq12*)d_gt_out;
    if (!d_f) {
        // the v > 8 path carves the aux arena as well: take both regions in one carve
        size_t sz[2] = {indep_products(ctx, v, pc ? pc->n : 0, n_checks) ? n_checks * v * sizeof(Fq12) : 0, n_checks * sizeof(Fq12)};
        void* t[2];
        TRY(ctx_aux(ctx, sz, 2, t));
        d_f = (Fq12*)t[1];
    }
    TRY(miller_products_dev(L, (const G1*)d_g1v, (const G2*)d_g2v, v, (const G1*)d_g1s, pc, n_checks, d_f));
    TRY(final_exp_dev(L, (const Fq12*)d_f, d_gt_out ? d_f : (Fq12*)nullptr, (uint8_t*)nullptr, d_is_one, n_checks));
    if (!d_gt_out) CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" {

int bn254_pairing_product_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_g2, size_t pairs_per_check,
                                   size_t n_checks, void* d_gt_out, uint8_t* d_is_one) {
    API_BEGIN
    ENTER(ctx);
    if (n_checks == 0) return 0;
    if (pairs_per_check && (!d_g1 || !d_g2)) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_g1, d_g2, d_gt_out);
    return product_checks_dev(ctx, nullptr, nullptr, d_g1, d_g2, pairs_per_check, n_checks, d_gt_out, d_is_one);
    API_END
}
int bn254_pairing_product_many(bn254_ctx* ctx, const void* g1, const void* g2, size_t pairs_per_check, size_t n_checks,
                               void* gt_out, uint8_t* is_one) {
    API_BEGIN
    ENTER(ctx);
    if (n_checks == 0) return 0;
    size_t n = pairs_per_check * n_checks;
    if (n && (!g1 || !g2)) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[4] = {n * sizeof(G1), n * sizeof(G2), n_checks * sizeof(Fq12), n_checks};
    void* d[4];
    TRY(ctx_io(ctx, sz, 4, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], g1, sz[0]));
    TRY(h2d(L, d[1], g2, sz[1]));
    TRY(product_checks_dev(ctx, nullptr, nullptr, d[0], d[1], pairs_per_check, n_checks, d[2], (uint8_t*)d[3]));
    if (gt_out) TRY(d2h(L, gt_out, d[2], sz[2]));
    if (is_one) TRY(d2h(L, is_one, d[3], n_checks));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
    API_END
}

// miller_loop_batch(&[(G2, G1)]) -> Result<Gt, CurveError>: n independent Miller loops on the split path, the infinity
// check reduced on the device to one flag, then the tree product
int bn254_miller_loop_batch(bn254_ctx* ctx, const void* g2, const void* g1, size_t n, void* gt_out) {
    API_BEGIN
    ENTER(ctx);
    if (!gt_out) return fail(BN254_ERR_ARG, "null pointer");
    if (n && (!g1 || !g2)) return fail(BN254_ERR_ARG, "null pointer");
    size_t cnt = n ? n : 1;
    size_t sz[6] = {cnt * sizeof(G1), cnt * sizeof(G2), cnt * sizeof(Fq12), ((cnt + 1) / 2) * sizeof(Fq12), cnt, 16};
    void* d[6];
    TRY(ctx_io(ctx, sz, 6, d));
    Lane L = main_lane(ctx);
    Fq12* res = nullptr;
    int h_bad = 0;
    if (n) {
        TRY(h2d(L, d[0], g1, n * sizeof(G1)));
        TRY(h2d(L, d[1], g2, n * sizeof(G2)));
        CU(cudaMemsetAsync(d[5], 0, sizeof(int), L.stream));
    }
    TRY(miller_product_dev(L, (const G1*)d[0], (const G2*)d[1], n, (Fq12*)d[2], (Fq12*)d[3], (uint8_t*)d[4], &res));
    if (n) {
        LAUNCH(L, k_any_zero_flag, n, (const uint8_t*)d[4], (int*)d[5], n);
        TRY(d2h(L, &h_bad, d[5], sizeof(int)));
        CU(cudaStreamSynchronize(ctx->stream));
        if (h_bad) return BN254_ERR_TO_AFFINE;   // Err(CurveError::ToAffineConversion), src/lib.rs:629-630; gt_out untouched
    }
    TRY(d2h(L, gt_out, res, sizeof(Fq12)));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
    API_END
}

int bn254_miller_product_dev(bn254_ctx* ctx, const void* d_g1, const void* d_g2, size_t n, void* d_gt_out) {
    API_BEGIN
    ENTER(ctx);
    if (!d_gt_out) return fail(BN254_ERR_ARG, "null pointer");
    if (n && (!d_g1 || !d_g2)) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_g1, d_g2, d_gt_out);
    size_t cnt = n ? n : 1;
    size_t sz[2] = {cnt * sizeof(Fq12), ((cnt + 1) / 2) * sizeof(Fq12)};
    void* t[2];
    TRY(ctx_aux(ctx, sz, 2, t));
    Lane L = main_lane(ctx);
    Fq12* res = nullptr;
    TRY(miller_product_dev(L, (const G1*)d_g1, (const G2*)d_g2, n, (Fq12*)t[0], (Fq12*)t[1], (uint8_t*)nullptr, &res));
    CU(cudaMemcpyAsync(d_gt_out, res, sizeof(Fq12), cudaMemcpyDeviceToDevice, L.stream));
    return 0;
    API_END
}
int bn254_miller_product(bn254_ctx* ctx, const void* g1, const void* g2, size_t n, void* gt_out) {
    API_BEGIN
    ENTER(ctx);
    if (!gt_out) return fail(BN254_ERR_ARG, "null pointer");
    if (n && (!g1 || !g2)) return fail(BN254_ERR_ARG, "null pointer");
    size_t cnt = n ? n : 1;
    size_t sz[4] = {cnt * sizeof(G1), cnt * sizeof(G2), cnt * sizeof(Fq12), ((cnt + 1) / 2) * sizeof(Fq12)};
    void* d[4];
    TRY(ctx_io(ctx, sz, 4, d));
    Lane L = main_lane(ctx);
    Fq12* res = nullptr;
    if (n) {
        TRY(h2d(L, d[0], g1, n * sizeof(G1)));
        TRY(h2d(L, d[1], g2, n * sizeof(G2)));
    }
    TRY(miller_product_dev(L, (const G1*)d[0], (const G2*)d[1], n, (Fq12*)d[2], (Fq12*)d[3], (uint8_t*)nullptr, &res));
    TRY(d2h(L, gt_out, res, sizeof(Fq12)));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
    API_END
}

// ---------------------------------------------------------------------------------------------
// Gt
// ---------------------------------------------------------------------------------------------
int bn254_final_exp_many_dev(bn254_ctx* ctx, const void* d_in, void* d_out, uint8_t* d_some, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    if (!d_in || !d_out) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_in, d_out);
    TRY(final_exp_dev(main_lane(ctx), (const Fq12*)d_in, (Fq12*)d_out, d_some, (uint8_t*)nullptr, n));
    return 0;
    API_END
}
int bn254_final_exp_many(bn254_ctx* ctx, const void* gt_in, void* gt_out, uint8_t* some, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    if (!gt_in || !gt_out) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[3] = {n * sizeof(Fq12), n * sizeof(Fq12), n};
    void* d[3];
    TRY(ctx_io(ctx, sz, 3, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], gt_in, sz[0]));
    TRY(final_exp_dev(L, (const Fq12*)d[0], (Fq12*)d[1], (uint8_t*)d[2], (uint8_t*)nullptr, n));
    TRY(d2h(L, gt_out, d[1], sz[1]));
    if (some) TRY(d2h(L, some, d[2], n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
    API_END
}
}  // extern "C"

// host-pointer wrapper of an elementwise Gt kernel: up to two inputs (a: Fq12, b: Fq12 or Fr), outputs Fq12 and/or a flag byte
template <class Launch>
static int gt_host(bn254_ctx* ctx, const void* a, const void* b, size_t b_unit, void* out, uint8_t* flag, size_t n, Launch launch) {
    if (n == 0) return 0;
    size_t sz[4] = {n * sizeof(Fq12), b ? n * b_unit : 0, n * sizeof(Fq12), n};
    void* d[4];
    TRY(ctx_io(ctx, sz, 4, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], a, sz[0]));
    if (b) TRY(h2d(L, d[1], b, sz[1]));
    TRY(launch(L, d));
    if (out) TRY(d2h(L, out, d[2], sz[2]));
    if (flag) TRY(d2h(L, flag, d[3], n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" {
int bn254_gt_mul_many(bn254_ctx* ctx, const void* a, const void* b, void* out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n && (!a || !b || !out)) return fail(BN254_ERR_ARG, "null pointer");
    return gt_host(ctx, a, b, sizeof(Fq12), out, nullptr, n, [&](const Lane& L, void** d) -> int {
        LAUNCH(L, k_gt_mul, n, (const Fq12*)d[0], (const Fq12*)d[1], (Fq12*)d[2], n);
        return 0;
    });
    API_END
}
int bn254_gt_pow_many(bn254_ctx* ctx, const void* a, const void* fr, void* out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n && (!a || !fr || !out)) return fail(BN254_ERR_ARG, "null pointer");
    return gt_host(ctx, a, fr, sizeof(Fq), out, nullptr, n, [&](const Lane& L, void** d) -> int {
        if (use_coop4(ctx, n)) {  // smallest batches: four warps per element (k_coop4_gt_pow)
            unsigned groups, g;
            coop4_shape(ctx, n, groups, g);
            LAUNCH_SYNC(L, k_coop4_gt_pow, g, groups * BN_COOP4_GROUP, BN_COOP4_SMEM(groups), (const Fq12*)d[0], (const Fq*)d[1], (Fq12*)d[2], n);
            return 0;
        }
        if (use_coop(ctx, n)) {   // small batches: one warp per element (k_coop_gt_pow)
            size_t per_sm = (n + (size_t)ctx->sms - 1) / (size_t)ctx->sms;
            unsigned warps = per_sm > BN_COOP_MAX_WARPS ? (unsigned)BN_COOP_MAX_WARPS : (unsigned)per_sm;
            unsigned g = (unsigned)((n + warps - 1) / warps);
            LAUNCH_SYNC(L, k_coop_gt_pow, g, warps * 32, BN_COOP_SMEM(warps), (const Fq12*)d[0], (const Fq*)d[1], (Fq12*)d[2], n);
            return 0;
        }
        LAUNCH(L, k_gt_pow, n, (const Fq12*)d[0], (const Fq*)d[1], (Fq12*)d[2], n);
        return 0;
    });
    API_END
}
int bn254_gt_inverse_many(bn254_ctx* ctx, const void* a, void* out, uint8_t* some, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n && (!a || !out)) return fail(BN254_ERR_ARG, "null pointer");
    return gt_host(ctx, a, nullptr, 0, out, some, n, [&](const Lane& L, void** d) -> int {
        LAUNCH(L, k_gt_inverse, n, (const Fq12*)d[0], (Fq12*)d[2], (uint8_t*)d[3], n);
        return 0;
    });
    API_END
}
int bn254_gt_eq_many(bn254_ctx* ctx, const void* a, const void* b, uint8_t* eq, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n && (!a || !b || !eq)) return fail(BN254_ERR_ARG, "null pointer");
    return gt_host(ctx, a, b, sizeof(Fq12), nullptr, eq, n, [&](const Lane& L, void** d) -> int {
        k_gt_eq<<<grid_for(n), BN_BLOCK, 0, L.stream>>>((const Fq12*)d[0], (const Fq12*)d[1], (uint8_t*)d[3], n);
        L.ctx->launches.fetch_add(1);
        CU(cudaGetLastError());
        return 0;
    });
    API_END
}
}  // extern "C"

// ---------------------------------------------------------------------------------------------
// G1 * Fr, G2 * Fr, normalize
// ---------------------------------------------------------------------------------------------
// (p * k).normalize() by a signed 4-bit window: uniform control flow, 64 table additions instead of ~254 executed ones
template <class F>
static int mul_normalized_dev(const Lane& L, const Jac<F>* d_p, const Fq* d_fr, Jac<F>* d_out, size_t n) {
    LAUNCH(L, k_jac_mul_window<F>, n, d_p, d_fr, d_out, n);
    return 0;
}
template <class P, class DevFn>
static int scalar_mul_host(bn254_ctx* ctx, const void* p, const void* fr, void* out, size_t n, DevFn dev) {
    if (n == 0) return 0;
    if (!p || !fr || !out) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[3] = {n * sizeof(P), n * sizeof(Fq), n * sizeof(P)};
    void* d[3];
    TRY(ctx_io(ctx, sz, 3, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], p, sz[0]));
    TRY(h2d(L, d[1], fr, sz[1]));
    TRY(dev(L, (const P*)d[0], (const Fq*)d[1], (P*)d[2], n));
    TRY(d2h(L, out, d[2], sz[2]));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
// G * Fr on the reference's exact sequence.  mul_compact = 1 (default): G1 batches of at least one 256-thread block take the
// block-compacted additions (k_jac_mul_compact: +6 %, 10.0 -> 10.6 M/s; the per-bit chain double -> barrier -> add -> barrier is
// latency-bound with two such blocks per SM, so halving the executed additions buys little, and nothing for G2 -- DESIGN F21);
// 2: G2 as well (tests); 0: always one thread per scalar (k_jac_mul).
template <class F>
static int mul_ref_dev(const Lane& L, const Jac<F>* p, const Fq* fr, Jac<F>* out, size_t n) {
    const bool g1 = sizeof(typename F::T) == sizeof(Fq);
    const int mode = L.ctx->opt.mul_compact;
    if ((mode == 2 || (mode == 1 && g1)) && n >= (size_t)BN_CMUL_BLOCK) {
        unsigned g = (unsigned)((n + BN_CMUL_BLOCK - 1) / BN_CMUL_BLOCK);
        LAUNCH_SYNC(L, k_jac_mul_compact<F>, g, BN_CMUL_BLOCK, CmulLayout<F>::BYTES, p, fr, out, n);
        return 0;
    }
    LAUNCH(L, k_jac_mul<F>, n, p, fr, out, n);
    return 0;
}
static int g1_mul_ref_dev(const Lane& L, const G1* p, const Fq* fr, G1* out, size_t n) { return mul_ref_dev<FqOps>(L, p, fr, out, n); }
static int g2_mul_ref_dev(const Lane& L, const G2* p, const Fq* fr, G2* out, size_t n) { return mul_ref_dev<Fq2Ops>(L, p, fr, out, n); }

extern "C" {
int bn254_g1_mul_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_fr, void* d_out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    if (!d_g1 || !d_fr || !d_out) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_g1, d_fr, d_out);
    return g1_mul_ref_dev(main_lane(ctx), (const G1*)d_g1, (const Fq*)d_fr, (G1*)d_out, n);
    API_END
}
int bn254_g2_mul_many_dev(bn254_ctx* ctx, const void* d_g2, const void* d_fr, void* d_out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    if (!d_g2 || !d_fr || !d_out) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_g2, d_fr, d_out);
    return g2_mul_ref_dev(main_lane(ctx), (const G2*)d_g2, (const Fq*)d_fr, (G2*)d_out, n);
    API_END
}
int bn254_g1_mul_many(bn254_ctx* ctx, const void* g1, const void* fr, void* out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return scalar_mul_host<G1>(ctx, g1, fr, out, n, g1_mul_ref_dev);
    API_END
}
int bn254_g2_mul_many(bn254_ctx* ctx, const void* g2, const void* fr, void* out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return scalar_mul_host<G2>(ctx, g2, fr, out, n, g2_mul_ref_dev);
    API_END
}
int bn254_g1_mul_normalized_many_dev(bn254_ctx* ctx, const void* d_g1, const void* d_fr, void* d_out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    if (!d_g1 || !d_fr || !d_out) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_g1, d_fr, d_out);
    return mul_normalized_dev<FqOps>(main_lane(ctx), (const G1*)d_g1, (const Fq*)d_fr, (G1*)d_out, n);
    API_END
}
int bn254_g2_mul_normalized_many_dev(bn254_ctx* ctx, const void* d_g2, const void* d_fr, void* d_out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    if (!d_g2 || !d_fr || !d_out) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_g2, d_fr, d_out);
    return mul_normalized_dev<Fq2Ops>(main_lane(ctx), (const G2*)d_g2, (const Fq*)d_fr, (G2*)d_out, n);
    API_END
}
int bn254_g1_mul_normalized_many(bn254_ctx* ctx, const void* g1, const void* fr, void* out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return scalar_mul_host<G1>(ctx, g1, fr, out, n, mul_normalized_dev<FqOps>);
    API_END
}
int bn254_g2_mul_normalized_many(bn254_ctx* ctx, const void* g2, const void* fr, void* out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return scalar_mul_host<G2>(ctx, g2, fr, out, n, mul_normalized_dev<Fq2Ops>);
    API_END
}
}  // extern "C"

template <class P, class F>
static int normalize_host(bn254_ctx* ctx, const void* p, void* out, size_t n) {
    if (n == 0) return 0;
    if (!p || !out) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[2] = {n * sizeof(P), n * sizeof(P)};
    void* d[2];
    TRY(ctx_io(ctx, sz, 2, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], p, sz[0]));
    LAUNCH(L, k_jac_normalize<F>, n, (const P*)d[0], (P*)d[1], n);
    TRY(d2h(L, out, d[1], sz[1]));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// host-pointer wrapper of the elementwise group operations (k_jac_elementwise): op 0 add, 1 double, 2 neg, 3 eq
template <class P, class F>
static int group_op_host(bn254_ctx* ctx, int op, const void* a, const void* b, void* out, uint8_t* eq, size_t n) {
    if (n == 0) return 0;
    const bool binary = op == 0 || op == 3;
    if (!a || (binary && !b) || (op == 3 ? !eq : !out)) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[4] = {n * sizeof(P), binary ? n * sizeof(P) : 0, op == 3 ? 0 : n * sizeof(P), op == 3 ? n : 0};
    void* d[4];
    TRY(ctx_io(ctx, sz, 4, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], a, sz[0]));
    if (binary) TRY(h2d(L, d[1], b, sz[1]));
    LAUNCH(L, k_jac_elementwise<F>, n, op, (const P*)d[0], (const P*)d[1], (P*)d[2], (uint8_t*)d[3], n);
    if (op == 3) TRY(d2h(L, eq, d[3], n)); else TRY(d2h(L, out, d[2], sz[2]));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
#define GROUP_OP_API(name, P, F, op, A, B, OUT, EQ)                                   \
    API_BEGIN                                                                         \
    ENTER(ctx);                                                                       \
    return group_op_host<P, F>(ctx, op, A, B, OUT, EQ, n);                            \
    API_END
extern "C" {
int bn254_g1_add_many(bn254_ctx* ctx, const void* a, const void* b, void* out, size_t n) { GROUP_OP_API(add, G1, FqOps, 0, a, b, out, nullptr) }
int bn254_g2_add_many(bn254_ctx* ctx, const void* a, const void* b, void* out, size_t n) { GROUP_OP_API(add, G2, Fq2Ops, 0, a, b, out, nullptr) }
int bn254_g1_double_many(bn254_ctx* ctx, const void* a, void* out, size_t n) { GROUP_OP_API(dbl, G1, FqOps, 1, a, nullptr, out, nullptr) }
int bn254_g2_double_many(bn254_ctx* ctx, const void* a, void* out, size_t n) { GROUP_OP_API(dbl, G2, Fq2Ops, 1, a, nullptr, out, nullptr) }
int bn254_g1_neg_many(bn254_ctx* ctx, const void* a, void* out, size_t n) { GROUP_OP_API(neg, G1, FqOps, 2, a, nullptr, out, nullptr) }
int bn254_g2_neg_many(bn254_ctx* ctx, const void* a, void* out, size_t n) { GROUP_OP_API(neg, G2, Fq2Ops, 2, a, nullptr, out, nullptr) }
int bn254_g1_eq_many(bn254_ctx* ctx, const void* a, const void* b, uint8_t* eq, size_t n) { GROUP_OP_API(eq, G1, FqOps, 3, a, b, nullptr, eq) }
int bn254_g2_eq_many(bn254_ctx* ctx, const void* a, const void* b, uint8_t* eq, size_t n) { GROUP_OP_API(eq, G2, Fq2Ops, 3, a, b, nullptr, eq) }
}  // extern "C"

static int affine_check_dev(bn254_ctx* ctx, const void* d_xy, uint8_t* d_status, size_t n, bool g2) {
    if (n == 0) return 0;
    if (!d_xy || !d_status) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_xy);
    Lane L = main_lane(ctx);
    if (g2) {
        LAUNCH(L, k_g2_affine_check, n, (const AffG2*)d_xy, d_status, n, ctx->opt.subgroup_ref != 0);
    } else {
        LAUNCH(L, k_g1_affine_check, n, (const AffG1*)d_xy, d_status, n);
    }
    return 0;
}
static int affine_check_host(bn254_ctx* ctx, const void* xy, uint8_t* status, size_t n, bool g2) {
    if (n == 0) return 0;
    if (!xy || !status) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[2] = {n * (g2 ? sizeof(AffG2) : sizeof(AffG1)), n};
    void* d[2];
    TRY(ctx_io(ctx, sz, 2, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], xy, sz[0]));
    TRY(affine_check_dev(ctx, d[0], (uint8_t*)d[1], n, g2));
    TRY(d2h(L, status, d[1], n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}
static int from_compressed_dev(bn254_ctx* ctx, const uint8_t* d_enc, void* d_out, uint8_t* d_status, size_t n, bool g2) {
    if (n == 0) return 0;
    if (!d_enc || !d_out || !d_status) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_out);
    Lane L = main_lane(ctx);
    if (g2) {
        LAUNCH(L, k_g2_from_compressed, n, d_enc, (G2*)d_out, d_status, n, ctx->opt.subgroup_ref != 0);
    } else {
        LAUNCH(L, k_g1_from_compressed, n, d_enc, (G1*)d_out, d_status, n);
    }
    return 0;
}
static int from_compressed_host(bn254_ctx* ctx, const uint8_t* enc, void* out, uint8_t* status, size_t n, bool g2) {
    if (n == 0) return 0;
    if (!enc || !out || !status) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[3] = {n * (g2 ? 65 : 33), n * (g2 ? sizeof(G2) : sizeof(G1)), n};
    void* d[3];
    TRY(ctx_io(ctx, sz, 3, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], enc, sz[0]));
    TRY(from_compressed_dev(ctx, (const uint8_t*)d[0], d[1], (uint8_t*)d[2], n, g2));
    TRY(d2h(L, out, d[1], sz[1]));
    TRY(d2h(L, status, d[2], n));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" {
int bn254_g1_normalize_many(bn254_ctx* ctx, const void* g1, void* out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return normalize_host<G1, FqOps>(ctx, g1, out, n);
    API_END
}
int bn254_g2_normalize_many(bn254_ctx* ctx, const void* g2, void* out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return normalize_host<G2, Fq2Ops>(ctx, g2, out, n);
    API_END
}
int bn254_g1_affine_check_many(bn254_ctx* ctx, const void* xy, uint8_t* status, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return affine_check_host(ctx, xy, status, n, false);
    API_END
}
int bn254_g2_affine_check_many(bn254_ctx* ctx, const void* xy, uint8_t* status, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return affine_check_host(ctx, xy, status, n, true);
    API_END
}
int bn254_g1_affine_check_many_dev(bn254_ctx* ctx, const void* d_xy, uint8_t* d_status, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return affine_check_dev(ctx, d_xy, d_status, n, false);
    API_END
}
int bn254_g2_affine_check_many_dev(bn254_ctx* ctx, const void* d_xy, uint8_t* d_status, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return affine_check_dev(ctx, d_xy, d_status, n, true);
    API_END
}
int bn254_g1_from_compressed_many(bn254_ctx* ctx, const uint8_t* enc, void* g1_out, uint8_t* status, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return from_compressed_host(ctx, enc, g1_out, status, n, false);
    API_END
}
int bn254_g2_from_compressed_many(bn254_ctx* ctx, const uint8_t* enc, void* g2_out, uint8_t* status, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return from_compressed_host(ctx, enc, g2_out, status, n, true);
    API_END
}
int bn254_g1_from_compressed_many_dev(bn254_ctx* ctx, const uint8_t* d_enc, void* d_g1_out, uint8_t* d_status, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return from_compressed_dev(ctx, d_enc, d_g1_out, d_status, n, false);
    API_END
}
int bn254_g2_from_compressed_many_dev(bn254_ctx* ctx, const uint8_t* d_enc, void* d_g2_out, uint8_t* d_status, size_t n) {
    API_BEGIN
    ENTER(ctx);
    return from_compressed_dev(ctx, d_enc, d_g2_out, d_status, n, true);
    API_END
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
// The tables are long-lived objects owned by the handle, so they are allocated once here (not from an arena).
int bn254_g2_precompute(bn254_ctx* ctx, const void* g2, size_t n, bn254_g2precomp** out) {
    API_BEGIN
    ENTER(ctx);
    if (!out || !g2 || n == 0) return fail(BN254_ERR_ARG, "bad argument");
    bn254_g2precomp* pc = new (std::nothrow) bn254_g2precomp();
    if (!pc) return fail(BN254_ERR_NOMEM, "precomp alloc");
    pc->device = ctx->device; pc->n = n; pc->tab = nullptr; pc->q_aff = nullptr;
    int rc = 0;
    if (cudaMalloc((void**)&pc->tab, (size_t)BN254_N_COEFFS * 12 * n * sizeof(uint4)) != cudaSuccess) rc = fail(BN254_ERR_NOMEM, "cudaMalloc table");
    if (rc == 0 && cudaMalloc((void**)&pc->q_aff, n * sizeof(AffG2)) != cudaSuccess) rc = fail(BN254_ERR_NOMEM, "cudaMalloc q");
    int h_bad = 0;
    auto body = [&]() -> int {
        size_t sz[2] = {n * sizeof(G2), 16};
        void* d[2];
        TRY(ctx_io(ctx, sz, 2, d));
        Lane L = main_lane(ctx);
        CU(cudaMemsetAsync(d[1], 0, sizeof(int), L.stream));
        TRY(h2d(L, d[0], g2, sz[0]));
        LAUNCH(L, k_g2_precompute, n, (const G2*)d[0], pc->tab, pc->q_aff, (int*)d[1], n);
        TRY(d2h(L, &h_bad, d[1], sizeof(int)));
        CU(cudaStreamSynchronize(ctx->stream));
        return 0;
    };
    if (rc == 0) rc = body();
    if (rc == 0 && h_bad) rc = BN254_ERR_TO_AFFINE;
    if (rc != 0) { bn254_g2precomp_destroy(pc); return rc; }
    *out = pc;
    return 0;
    API_END
}
int bn254_g2precomp_export(bn254_ctx* ctx, const bn254_g2precomp* pc, size_t i, void* out) {
    API_BEGIN
    ENTER(ctx);
    if (!pc || !out || i >= pc->n) return fail(BN254_ERR_ARG, "bad argument");
    uint8_t* o = (uint8_t*)out;
    Lane L = main_lane(ctx);
    TRY(d2h(L, o, pc->q_aff + i, sizeof(AffG2)));
    // gather the strided 16-byte words of table i: one 2D copy (pitch = n*16 bytes)
    CU(cudaMemcpy2DAsync(o + sizeof(AffG2), sizeof(uint4), pc->tab + i, pc->n * sizeof(uint4), sizeof(uint4),
                         (size_t)BN254_N_COEFFS * 12, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
    API_END
}
int bn254_pairing_product_precomp_many_dev(bn254_ctx* ctx, const bn254_g2precomp* pc, const void* d_g1s, const void* d_g1v,
                                           const void* d_g2v, size_t v, size_t n_checks, void* d_gt_out, uint8_t* d_is_one) {
    API_BEGIN
    ENTER(ctx);
    if (n_checks == 0) return 0;
    if (pc && pc->device != ctx->device) return fail(BN254_ERR_ARG, "precomp tables live on another device");
    if ((pc && !d_g1s) || (v && (!d_g1v || !d_g2v))) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_g1s, d_g1v, d_g2v, d_gt_out);
    return product_checks_dev(ctx, pc, d_g1s, d_g1v, d_g2v, v, n_checks, d_gt_out, d_is_one);
    API_END
}
int bn254_pairing_product_precomp_many(bn254_ctx* ctx, const bn254_g2precomp* pc, const void* g1_shared, const void* g1_var,
                                       const void* g2_var, size_t v, size_t n_checks, void* gt_out, uint8_t* is_one) {
    API_BEGIN
    ENTER(ctx);
    if (n_checks == 0) return 0;
    size_t m = pc ? pc->n : 0;
    if (pc && pc->device != ctx->device) return fail(BN254_ERR_ARG, "precomp tables live on another device");
    if ((m && !g1_shared) || (v && (!g1_var || !g2_var))) return fail(BN254_ERR_ARG, "null pointer");
    size_t sz[5] = {n_checks * m * sizeof(G1), n_checks * v * sizeof(G1), n_checks * v * sizeof(G2), n_checks * sizeof(Fq12), n_checks};
    void* d[5];
    TRY(ctx_io(ctx, sz, 5, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], g1_shared, sz[0]));
    TRY(h2d(L, d[1], g1_var, sz[1]));
    TRY(h2d(L, d[2], g2_var, sz[2]));
    TRY(product_checks_dev(ctx, pc, m ? d[0] : nullptr, d[1], d[2], v, n_checks, d[3], (uint8_t*)d[4]));
    if (gt_out) TRY(d2h(L, gt_out, d[3], sz[3]));
    if (is_one) TRY(d2h(L, is_one, d[4], n_checks));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
    API_END
}

// ---------------------------------------------------------------------------------------------
// one product sharded over several devices: 384-byte partials gathered by peer copy (no NCCL)
// ---------------------------------------------------------------------------------------------
}  // extern "C"
static double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
// Phase 1 runs one host thread per context: copies in (concurrent across devices also for pageable memory), the split Miller
// path, the tree product; each context records ev_done behind its partial.  Phase 2 enqueues all peer copies on ctxs[0]'s
// stream behind those events; phase 3 multiplies the partials and exponentiates once.  ONE host synchronisation at the end
// (the timed twin adds one per phase).
static int product_sharded(bn254_ctx** ctxs, size_t n_ctx, const void* g1, const void* g2, size_t n, void* gt_out, uint8_t* is_one,
                           double* phase_ms) {
    if (!ctxs || n_ctx == 0) return fail(BN254_ERR_ARG, "no contexts");
    for (size_t d = 0; d < n_ctx; d++) if (!ctxs[d]) return fail(BN254_ERR_ARG, "null context");
    if (n && (!g1 || !g2)) return fail(BN254_ERR_ARG, "null pointer");
    // hold every context for the whole call (fixed order: no deadlock between two sharded calls over the same list)
    std::vector<std::unique_lock<std::recursive_mutex>> locks;
    locks.reserve(n_ctx);
    for (size_t d = 0; d < n_ctx; d++) {
        bool dup = false;
        for (size_t e = 0; e < d; e++) dup = dup || ctxs[e] == ctxs[d];
        if (dup) return fail(BN254_ERR_ARG, "the same context appears twice in the list");
        locks.emplace_back(ctxs[d]->mu);
    }
    struct Shard { int rc = 0; char err[320] = ""; Fq12* res = nullptr; };
    std::vector<Shard> sh(n_ctx);
    double t0 = now_ms();
    Workers workers;
    std::vector<std::thread>& th = workers.th;
    th.reserve(n_ctx);
    for (size_t d = 0; d < n_ctx; d++) {
        size_t lo = n * d / n_ctx, cnt = n * (d + 1) / n_ctx - lo;
        th.emplace_back([=, &sh]() {
            bn254_ctx* c = ctxs[d];
            auto body = [&]() -> int {
                TRY(use(c));
                size_t k = cnt ? cnt : 1;
                size_t sz[4] = {k * sizeof(G1), k * sizeof(G2), k * sizeof(Fq12), ((k + 1) / 2) * sizeof(Fq12)};
                void* dv[4];
                TRY(ctx_io(c, sz, 4, dv));
                Lane L = main_lane(c);
                if (cnt) {
                    TRY(h2d(L, dv[0], (const char*)g1 + lo * sizeof(G1), cnt * sizeof(G1)));
                    TRY(h2d(L, dv[1], (const char*)g2 + lo * sizeof(G2), cnt * sizeof(G2)));
                }
                TRY(miller_product_dev(L, (const G1*)dv[0], (const G2*)dv[1], cnt, (Fq12*)dv[2], (Fq12*)dv[3], (uint8_t*)nullptr, &sh[d].res));
                CU(cudaEventRecord(c->ev_done, c->stream));
                if (phase_ms) CU(cudaStreamSynchronize(c->stream));
                return 0;
            };
            int rc;
            try { rc = body(); } catch (...) { rc = fail(BN254_ERR_NOMEM, "exception in a device worker"); }
            sh[d].rc = rc;
            if (rc != 0) snprintf(sh[d].err, sizeof sh[d].err, "device %d: %s", c->device, g_err);
        });
    }
    workers.join();
    for (size_t d = 0; d < n_ctx; d++) if (sh[d].rc != 0) return fail(sh[d].rc, sh[d].err);
    double t1 = now_ms();
    // phase 2: gather on device 0 (peer copies over NVLink when the devices are peers; staged by the driver otherwise)
    bn254_ctx* c0 = ctxs[0];
    TRY(use(c0));
    size_t sz[4] = {n_ctx * sizeof(Fq12), ((n_ctx + 1) / 2) * sizeof(Fq12), sizeof(Fq12), 16};
    void* a[4];
    TRY(ctx_aux(c0, sz, 4, a));
    Lane L0 = main_lane(c0);
    for (size_t d = 0; d < n_ctx; d++) {
        bn254_ctx* c = ctxs[d];
        if (c != c0) CU(cudaStreamWaitEvent(c0->stream, c->ev_done, 0));
        if (c->device == c0->device) CU(cudaMemcpyAsync((Fq12*)a[0] + d, sh[d].res, sizeof(Fq12), cudaMemcpyDeviceToDevice, c0->stream));
        else CU(cudaMemcpyPeerAsync((Fq12*)a[0] + d, c0->device, sh[d].res, c->device, sizeof(Fq12), c0->stream));
    }
    if (phase_ms) CU(cudaStreamSynchronize(c0->stream));
    double t2 = now_ms();
    // phase 3: <= n_ctx - 1 Fq12 multiplications and ONE final exponentiation on device 0
    Fq12* res = nullptr;
    TRY(tree_product(L0, (Fq12*)a[0], (Fq12*)a[1], n_ctx, &res));
    TRY(final_exp_dev(L0, (const Fq12*)res, (Fq12*)a[2], (uint8_t*)nullptr, (uint8_t*)a[3], 1));
    if (gt_out) TRY(d2h(L0, gt_out, a[2], sizeof(Fq12)));
    if (is_one) TRY(d2h(L0, is_one, a[3], 1));
    CU(cudaStreamSynchronize(c0->stream));
    if (phase_ms) { phase_ms[0] = t1 - t0; phase_ms[1] = t2 - t1; phase_ms[2] = now_ms() - t2; }
    return 0;
}
extern "C" {
int bn254_pairing_product_sharded(bn254_ctx** ctxs, size_t n_ctx, const void* g1, const void* g2, size_t n, void* gt_out,
                                  uint8_t* is_one) {
    API_BEGIN
    return product_sharded(ctxs, n_ctx, g1, g2, n, gt_out, is_one, nullptr);
    API_END
}
int bn254_pairing_product_sharded_timed(bn254_ctx** ctxs, size_t n_ctx, const void* g1, const void* g2, size_t n, void* gt_out,
                                        uint8_t* is_one, double* phase_ms) {
    API_BEGIN
    if (!phase_ms) return fail(BN254_ERR_ARG, "null pointer");
    return product_sharded(ctxs, n_ctx, g1, g2, n, gt_out, is_one, phase_ms);
    API_END
}

// ---------------------------------------------------------------------------------------------
// field-level test hook and roofline microbenchmarks
// ---------------------------------------------------------------------------------------------
int bn254_field_op_many(bn254_ctx* ctx, int op, const void* a, const void* b, void* out, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    static const size_t unit_of[16] = {32, 32, 32, 64, 64, 192, 384, 384, 384, 32, 64, 192, 384, 384, 384, 192};
    if (op < 0 || op > 15 || !a || !out) return fail(BN254_ERR_ARG, "bad argument");
    bool binary = (op == 0 || op == 1 || op == 2 || op == 3 || op == 5 || op == 6 || op == 11 || op == 12 || op == 15);
    if (binary && !b) return fail(BN254_ERR_ARG, "null pointer");
    size_t bytes = unit_of[op] * n;
    size_t sz[3] = {bytes, bytes, bytes};
    void* d[3];
    TRY(ctx_io(ctx, sz, 3, d));
    Lane L = main_lane(ctx);
    TRY(h2d(L, d[0], a, bytes));
    if (binary) TRY(h2d(L, d[1], b, bytes));
    if (op >= 11) {
        LAUNCH_SYNC(L, k_slot_op, grid_for(n), BN_BLOCK, BN_BLOCK * BN_SLOT_STRIDE, op, (const uint4*)d[0], (const uint4*)d[1], (uint4*)d[2], n);
    } else {
        LAUNCH(L, k_field_op, n, op, (const uint4*)d[0], (const uint4*)d[1], (uint4*)d[2], n);
    }
    TRY(d2h(L, out, d[2], bytes));
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
    API_END
}

int bn254_microbench(bn254_ctx* ctx, int kind, int grid, int block, int iters, double* ms, double* ops) {
    API_BEGIN
    ENTER(ctx);
    if (kind < 0 || kind > 2 || grid <= 0 || block <= 0 || block > 256 || iters <= 0 || !ms || !ops) return fail(BN254_ERR_ARG, "bad argument");
    size_t sz[1] = {16};
    void* sink[1];
    TRY(ctx_aux(ctx, sz, 1, sink));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
    for (int rep = 0; rep < 2; rep++) {   // first pass warms up
        CU(cudaEventRecord(e0, ctx->stream));
        if (kind == 0) k_microbench<0><<<grid, block, 0, ctx->stream>>>((uint32_t*)sink[0], iters, 12345u);
        else if (kind == 1) k_microbench<1><<<grid, block, 0, ctx->stream>>>((uint32_t*)sink[0], iters, 12345u);
        else k_microbench<2><<<grid, block, 0, ctx->stream>>>((uint32_t*)sink[0], iters, 12345u);
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
    API_END
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// bn254_final_exp_many_dev_noalloc: the final exponentiation with its scratch in a static device array.
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, tight serving loops, harnesses that forbid it).  The regular entry point keeps the easy part's values
// and Option flags in the context's grow-only workspace (cudaMalloc on growth); this one keeps them here.
//
// Every GPU that loads the library holds the array until exit: BN254_NOALLOC_MAX_N x (384 + 1) bytes = 1.5 MB.
// All calls on a device SHARE it and it keeps the last call's contents.  Nothing in it has to start at zero: the easy-part
// kernel writes value i and flag i for every i < n before the hard-part kernel, ordered behind it on the same stream, reads
// them, and no kernel reads an index >= n.  Calls that overlap on the GPU (two contexts / streams of one device) race on it.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char bn254_noalloc_scratch[(size_t)BN254_NOALLOC_MAX_N * (sizeof(Fq12) + 1)];

// In host code the array's name is a host-side stand-in, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress, taken once per device with that device current (context set-up) and cached here.
static const int NOALLOC_MAX_DEVICES = 64;
static void* g_noalloc_addr[NOALLOC_MAX_DEVICES];
static cudaError_t g_noalloc_err[NOALLOC_MAX_DEVICES];
static bool g_noalloc_done[NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mu;
static void noalloc_setup(int device) {
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return;
    std::lock_guard<std::mutex> g(g_noalloc_mu);
    if (g_noalloc_done[device]) return;
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&ptr, bn254_noalloc_scratch);
    g_noalloc_err[device] = e;
    g_noalloc_addr[device] = e == cudaSuccess ? ptr : nullptr;
    g_noalloc_done[device] = true;
    if (e != cudaSuccess) cudaGetLastError();   // reported by the entry point, not by bn254_ctx_create
}

extern "C" int bn254_final_exp_many_dev_noalloc(bn254_ctx* ctx, const void* d_gt_in, void* d_gt_out, uint8_t* d_some, size_t n) {
    API_BEGIN
    ENTER(ctx);
    if (n == 0) return 0;
    if (n > (size_t)BN254_NOALLOC_MAX_N)
        return fail(BN254_ERR_ARG, "bn254_final_exp_many_dev_noalloc: n exceeds BN254_NOALLOC_MAX_N (the static scratch holds 4096 values); "
                                   "split the batch or use bn254_final_exp_many_dev");
    if (!d_gt_in || !d_gt_out) return fail(BN254_ERR_ARG, "null pointer");
    CHECK_ALIGNED(d_gt_in, d_gt_out);
    if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES) return fail(BN254_ERR_ARG, "device index beyond the no-alloc table");
    void* scratch = g_noalloc_addr[ctx->device];
    if (!scratch) return fail(BN254_ERR_CUDA, "cudaGetSymbolAddress(bn254_noalloc_scratch)", g_noalloc_err[ctx->device]);
    if (ctx->opt.fexp_smem != 2) return fail(BN254_ERR_ARG, "the no-alloc variant runs the default (fexp_smem = 2) kernels only");
    Lane L = main_lane(ctx);
    Fq12* t = (Fq12*)scratch;                                                   // easy-part values
    uint8_t* flags = (uint8_t*)scratch + (size_t)BN254_NOALLOC_MAX_N * sizeof(Fq12);   // Option flags of the easy part
    LAUNCH(L, k_final_exp_easy, n, (const Fq12*)d_gt_in, t, flags, n);
    unsigned bs = sync_block(L, n), g = (unsigned)((n + bs - 1) / bs);
    LAUNCH_SYNC(L, k_final_exp_hard_sync, g, bs, sync_smem(bs), (const Fq12*)t, (const uint8_t*)flags, (Fq12*)d_gt_out, d_some, (uint8_t*)nullptr, n);
    return 0;
    API_END
}

#ifdef BN_SLOT_CHECK
// ---------------------------------------------------------------------------------------------
// Test build only (-DBN_SLOT_CHECK, libbn254_b200_chk.so): read-out of the shared-memory ownership / bounds checker of
// bn_slots.cuh, and a negative control that proves it is live.  Not declared in include/bn254_b200.h.
// ---------------------------------------------------------------------------------------------
namespace bn {
// thread 5 writes the first word of thread 6's slots through the checked accessor: exactly one violation, nothing stored
__global__ void k_slot_check_poke() {
    const sp base = s_base();
    if (threadIdx.x == 5) sts128(base + BN_SLOT_STRIDE, make_uint4(1, 2, 3, 4));
}
}  // namespace bn
extern "C" int bn254_debug_slot_violations(bn254_ctx* ctx, unsigned long long* out4) {
    API_BEGIN
    ENTER(ctx);
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream2));
    CU(cudaMemcpyFromSymbol(out4, bn_slot_violations, 4 * sizeof(unsigned long long)));
    return 0;
    API_END
}
extern "C" int bn254_debug_slot_reset(bn254_ctx* ctx) {
    API_BEGIN
    ENTER(ctx);
    unsigned long long z[4] = {0, 0, 0, 0};
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaMemcpyToSymbol(bn_slot_violations, z, sizeof z));
    return 0;
    API_END
}
extern "C" int bn254_debug_slot_poke(bn254_ctx* ctx) {
    API_BEGIN
    ENTER(ctx);
    Lane L = main_lane(ctx);
    LAUNCH_SYNC(L, k_slot_check_poke, 1, BN_BLOCK, BN_BLOCK * BN_SLOT_STRIDE);
    CU(cudaStreamSynchronize(ctx->stream));
    return 0;
    API_END
}
#endif
