// wkv6_capi.cu -- the C ABI of include/wkv6_b200.h: argument validation, TMA tensor-map encoding,
// kernel selection and launch.  No torch types, no allocation, no synchronisation.
#include "../../include/wkv6_b200.h"
#include "wkv6_scan.cuh"

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>

namespace wkv6 {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};
static std::atomic<int> g_algo{WKV6_ALGO_AUTO};

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static unsigned long long* g_dbg = nullptr;    // debug aid only (wkv6_b200_debug_trace), never on the data path
unsigned long long* debug_trace_buffer() { return g_dbg; }

static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

static int cuda_fail(int e, const char* where) {
    if (e == 0) return 0;
    snprintf(g_err, sizeof(g_err), "%s: %s (%d)", where, cudaGetErrorString((cudaError_t)e), e);
    return e > 0 ? e : -e;
}

typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_fn_t get_encode() {
    static encode_fn_t fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<encode_fn_t>(p);
    });
    return fn;
}

// Encoded tensor maps are pure functions of (pointer, shape, box, swizzle): a small per-thread direct-mapped cache saves
// the 10 (forward) / 22 (backward) cuTensorMapEncodeTiled driver calls of a bidirectional launch whenever a caller comes
// back with the same buffers (the caching allocator of PyTorch hands the same blocks out step after step).
struct MapKey {
    const void* ptr;
    int kind, B, T, C, rows, swz;
    bool operator==(const MapKey& o) const {
        return ptr == o.ptr && kind == o.kind && B == o.B && T == o.T && C == o.C && rows == o.rows && swz == o.swz;
    }
};
struct MapSlot {
    MapKey key;
    CUtensorMap map;
    bool valid;
};
constexpr int kMapSlots = 256;
static MapSlot* map_cache() {
    static thread_local MapSlot slots[kMapSlots];   // zero-initialised: valid = false
    return slots;
}
static unsigned map_hash(const MapKey& k) {
    unsigned long long h = reinterpret_cast<uintptr_t>(k.ptr) >> 4;
    h = h * 0x9E3779B97F4A7C15ull ^ ((unsigned long long)k.T * 131u + (unsigned)k.B * 31u + (unsigned)k.C + (unsigned)k.rows * 7u +
                                     (unsigned)k.kind * 3u + (unsigned)k.swz);
    return (unsigned)(h >> 40) & (kMapSlots - 1);
}

static thread_local bool ctx_bound = false;
static void bind_context_once() {
    // cuTensorMapEncodeTiled wants a current context; threads that never touched the runtime (autograd workers)
    // have none until a runtime call binds the primary context.
    if (!ctx_bound) { cudaFree(nullptr); ctx_bound = true; }
}

// 3-D view [B][T][C] of a contiguous tensor, box = [1][rows][64 channels]
static int make_map(CUtensorMap* m, int dtype, const void* ptr, int B, int T, int C, int rows, bool swizzle128 = false) {
    const MapKey key = {ptr, dtype, B, T, C, rows, swizzle128 ? 1 : 0};
    MapSlot& slot = map_cache()[map_hash(key)];
    if (slot.valid && slot.key == key) { *m = slot.map; return 0; }
    encode_fn_t enc = get_encode();
    if (!enc) return fail(WKV6_E_NODRIVER, "cuTensorMapEncodeTiled is not available from the CUDA driver");
    bind_context_once();
    const size_t es = dtype == WKV6_DTYPE_BF16 ? 2 : 4;
    const cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)T, (cuuint64_t)B};
    const cuuint64_t strides[2] = {(cuuint64_t)C * es, (cuuint64_t)T * C * es};
    const cuuint32_t box[3] = {64u, (cuuint32_t)rows, 1u};
    const cuuint32_t estr[3] = {1u, 1u, 1u};
    const CUresult r = enc(m, dtype == WKV6_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                           3, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(WKV6_E_BADARG, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    slot.key = key; slot.map = *m; slot.valid = true;
    return 0;
}

// chunk-start state checkpoints: [B*H*nchunks*64 rows][64] bf16, box 64 x 64, 128B swizzle
static int make_ckpt_map(CUtensorMap* m, const void* ptr, int B, int T, int H) {
    const MapKey key = {ptr, 100, B, T, H, 64, 1};
    MapSlot& slot = map_cache()[map_hash(key)];
    if (slot.valid && slot.key == key) { *m = slot.map; return 0; }
    encode_fn_t enc = get_encode();
    if (!enc) return fail(WKV6_E_NODRIVER, "cuTensorMapEncodeTiled is not available from the CUDA driver");
    bind_context_once();
    const int nchunks = (T + 63) / 64;
    const cuuint64_t dims[2] = {64u, (cuuint64_t)B * H * nchunks * 64};
    const cuuint64_t strides[1] = {128u};
    const cuuint32_t box[2] = {64u, 64u};
    const cuuint32_t estr[2] = {1u, 1u};
    const CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(WKV6_E_BADARG, "cuTensorMapEncodeTiled(ckpt) failed with CUresult %d", (int)r);
    slot.key = key; slot.map = *m; slot.valid = true;
    return 0;
}

static size_t ckpt_bytes(int B, int T, int H) { return (size_t)B * H * ((T + 63) / 64) * 64 * 64 * 2; }

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int check_common(int dtype, int B, int T, int C, int H) {
    if (dtype != WKV6_DTYPE_BF16 && dtype != WKV6_DTYPE_F32) return fail(WKV6_E_BADDTYPE, "unknown dtype %d", dtype);
    if (B < 0 || T < 0 || C <= 0 || H <= 0) return fail(WKV6_E_BADARG, "bad sizes B=%d T=%d C=%d H=%d", B, T, C, H);
    if (C != H * N) return fail(WKV6_E_BADSHAPE, "C (%d) must equal H (%d) * 64: head size 64 only", C, H);
    return 0;
}

struct FwdSet {
    const void *r, *k, *v, *w, *u;
    const float* s0;
    void* y;
    float* sT;
    void* ckpt;
};
struct BwdSet {
    const void *r, *k, *v, *w, *u;
    const float* s0;
    const void* gy;
    const float* gsT;
    void *gr, *gk, *gv, *gw, *gu;
    float* gs0;
    const void* ckpt;
};

static bool use_chunk(int dtype, int T) {
    const int a = g_algo.load(std::memory_order_relaxed);
    if (dtype != WKV6_DTYPE_BF16) return false;
    if (a == WKV6_ALGO_SCAN) return false;
    (void)T;
    return true;    // AUTO and CHUNK: bf16 forward runs on the tcgen05 chunk kernel (2.3x the exact scan on cfg2)
}

static int fill_fwd(DirArgs& d, int dtype, int B, int T, int C, const FwdSet& s, int reverse, int rows, bool want_ymap,
                    bool skip_y = false) {
    const void* ptrs[5] = {s.r, s.k, s.v, s.w, skip_y ? s.r : s.y};
    for (const void* q : ptrs)
        if (!q || !aligned16(q)) return fail(WKV6_E_BADARG, "forward: NULL or misaligned tensor pointer");
    if (!s.u) return fail(WKV6_E_BADARG, "forward: u is NULL");
    if ((s.s0 && !aligned16(s.s0)) || (s.sT && !aligned16(s.sT))) return fail(WKV6_E_BADARG, "forward: misaligned state");
    int e;
    if ((e = make_map(&d.mr, dtype, s.r, B, T, C, rows, want_ymap))) return e;
    if ((e = make_map(&d.mk, dtype, s.k, B, T, C, rows, want_ymap))) return e;
    if ((e = make_map(&d.mv, dtype, s.v, B, T, C, rows, want_ymap))) return e;
    if ((e = make_map(&d.mw, dtype, s.w, B, T, C, rows, want_ymap))) return e;
    if (want_ymap && s.y && (e = make_map(&d.my, dtype, s.y, B, T, C, rows, true))) return e;
    d.ckpt = nullptr;
    if (want_ymap && s.ckpt) {
        if (!aligned16(s.ckpt)) return fail(WKV6_E_BADARG, "forward: misaligned checkpoint buffer");
        if ((e = make_ckpt_map(&d.mc, s.ckpt, B, T, C / N))) return e;
        d.ckpt = s.ckpt;
    }
    d.u = s.u; d.s0 = s.s0; d.sT = s.sT; d.y = s.y;
    d.reverse = reverse;
    return 0;
}

// ---- wkv6_b200_forward_static_scratch: the segmentation scratch of callers that must not allocate.
// One array per device that loads this library, held until exit.  `g_static_scratch` in HOST code is only a handle of the
// symbol (its host-side shadow), never a device address: the device address comes from cudaGetSymbolAddress, looked up once
// per device with that device current, and is what the kernels are given.
static __device__ __align__(256) unsigned char g_static_scratch[WKV6_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
static std::atomic<void*> g_static_scratch_ptr[kMaxDevices];    // zero-initialised

static int static_scratch_address(void** out) {
    int dev = 0;
    cudaError_t ce = cudaGetDevice(&dev);
    if (ce != cudaSuccess) return cuda_fail((int)ce, "forward_static_scratch: cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) return fail(WKV6_E_BADARG, "forward_static_scratch: device ordinal %d (at most %d)", dev, kMaxDevices);
    if (!ctx_bound) {
        // a thread that never touched the runtime has no current context for cuTensorMapEncodeTiled: cudaSetDevice binds the
        // primary context (and, unlike bind_context_once, is not spelled cudaFree)
        if ((ce = cudaSetDevice(dev)) != cudaSuccess) return cuda_fail((int)ce, "forward_static_scratch: cudaSetDevice");
        ctx_bound = true;
    }
    void* ptr = g_static_scratch_ptr[dev].load(std::memory_order_acquire);
    if (!ptr) {
        ce = cudaGetSymbolAddress(&ptr, g_static_scratch);     // the current device's copy of the array
        if (ce != cudaSuccess) return cuda_fail((int)ce, "forward_static_scratch: cudaGetSymbolAddress");
        if (!ptr || !aligned16(ptr)) return fail(WKV6_E_BADARG, "forward_static_scratch: bad symbol address");
        g_static_scratch_ptr[dev].store(ptr, std::memory_order_release);   // racing threads store the same value
    }
    *out = ptr;
    return 0;
}

static int run_forward(int dtype, int B, int T, int C, int H, const FwdSet* sets, const int* reverse, int nset,
                       void* scratch, size_t scratch_bytes, void* stream) {
    int e = check_common(dtype, B, T, C, H);
    if (e) return e;
    if (nset == 0 || B == 0) return 0;
    Params p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.T = T; p.C = C; p.H = H; p.ndir = nset;
    if (T == 0) {   // empty sequence: S_T = S_0
        for (int i = 0; i < nset; i++)
            if (sets[i].sT) {
                const size_t bytes = (size_t)B * H * N * N * sizeof(float);
                cudaError_t ce = sets[i].s0 ? cudaMemcpyAsync(sets[i].sT, sets[i].s0, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream)
                                            : cudaMemsetAsync(sets[i].sT, 0, bytes, (cudaStream_t)stream);
                if (ce != cudaSuccess) return cuda_fail((int)ce, "forward(T=0)");
            }
        return 0;
    }
    const bool chunk = use_chunk(dtype, T);
    if (scratch && !aligned16(scratch)) return fail(WKV6_E_BADARG, "forward: misaligned scratch buffer");
    for (int i = 0; i < nset; i++)
        if ((e = fill_fwd(p.d[i], dtype, B, T, C, sets[i], reverse[i], chunk ? 64 : TT, chunk))) return e;
    if (chunk) {
        // long sequences with few (batch, head) pairs: time-segmented two-pass forward on the caller's scratch buffer
        e = launch_fwd_chunk_segmented(p, scratch, scratch_bytes, (cudaStream_t)stream);
        if (e < 0) e = launch_fwd_chunk(p, (cudaStream_t)stream);
    } else {
        e = launch_fwd_scan_segmented(dtype, p, scratch, scratch_bytes, (cudaStream_t)stream);
        if (e < 0) e = launch_fwd_scan(dtype, p, (cudaStream_t)stream);
    }
    return cuda_fail(e, "wkv6 forward launch");
}

extern "C" size_t wkv6_b200_backward_workspace_bytes(int, int, int, int, int, int);
extern "C" size_t wkv6_b200_backward_scratch_bytes(int, int, int, int, int, int);

static int run_backward(int dtype, int B, int T, int C, int H, const BwdSet* sets, const int* reverse, int nset,
                        void* ws, size_t ws_bytes, void* stream) {
    int e = check_common(dtype, B, T, C, H);
    if (e) return e;
    if (nset == 0 || B == 0) return 0;
    const size_t es = dtype == WKV6_DTYPE_BF16 ? 2 : 4;
    if (T == 0) {   // no steps: all sequence gradients are empty, gu = 0, gs0 = gsT
        for (int i = 0; i < nset; i++) {
            cudaError_t ce = cudaMemsetAsync(sets[i].gu, 0, (size_t)B * C * es, (cudaStream_t)stream);
            if (ce == cudaSuccess && sets[i].gs0) {
                const size_t bytes = (size_t)B * H * N * N * sizeof(float);
                ce = sets[i].gsT ? cudaMemcpyAsync(sets[i].gs0, sets[i].gsT, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream)
                                 : cudaMemsetAsync(sets[i].gs0, 0, bytes, (cudaStream_t)stream);
            }
            if (ce != cudaSuccess) return cuda_fail((int)ce, "backward(T=0)");
        }
        return 0;
    }
    Params p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.T = T; p.C = C; p.H = H; p.ndir = nset;
    if (use_chunk(dtype, T)) {
        // tcgen05 chunk backward: one reverse sweep over checkpoints of the chunk-start states.  Checkpoints come
        // from the forward (ckpt argument) or are recomputed here into the workspace by a state-only forward.
        const size_t cb = ckpt_bytes(B, T, H);
        bool recompute = false;
        Params pf;
        memset(&pf, 0, sizeof(pf));
        pf.B = B; pf.T = T; pf.C = C; pf.H = H; pf.ndir = 0; pf.skip_y = 1;
        for (int i = 0; i < nset; i++) {
            const BwdSet& s = sets[i];
            DirArgs& d = p.d[i];
            const void* ptrs[10] = {s.r, s.k, s.v, s.w, s.gy, s.gr, s.gk, s.gv, s.gw, s.gu};
            for (const void* q : ptrs)
                if (!q || !aligned16(q)) return fail(WKV6_E_BADARG, "backward: NULL or misaligned tensor pointer");
            if (!s.u) return fail(WKV6_E_BADARG, "backward: u is NULL");
            const void* ck = s.ckpt;
            if (!ck) {
                if (!ws || ws_bytes < cb * nset || !aligned16(ws))
                    return fail(WKV6_E_WORKSPACE, "backward without checkpoints needs %zu workspace bytes", cb * nset);
                void* mine = reinterpret_cast<char*>(ws) + cb * i;
                const FwdSet fs = {s.r, s.k, s.v, s.w, s.u, s.s0, nullptr, nullptr, mine};
                if ((e = fill_fwd(pf.d[pf.ndir], dtype, B, T, C, fs, reverse[i], 64, true, true))) return e;
                pf.ndir++;
                recompute = true;
                ck = mine;
            }
            if ((e = make_map(&d.mr, dtype, s.r, B, T, C, 64, true))) return e;
            if ((e = make_map(&d.mk, dtype, s.k, B, T, C, 64, true))) return e;
            if ((e = make_map(&d.mv, dtype, s.v, B, T, C, 64, true))) return e;
            if ((e = make_map(&d.mw, dtype, s.w, B, T, C, 64, true))) return e;
            if ((e = make_map(&d.mg, dtype, s.gy, B, T, C, 64, true))) return e;
            if ((e = make_map(&d.mgr, dtype, s.gr, B, T, C, 64, true))) return e;
            if ((e = make_map(&d.mgk, dtype, s.gk, B, T, C, 64, true))) return e;
            if ((e = make_map(&d.mgv, dtype, s.gv, B, T, C, 64, true))) return e;
            if ((e = make_map(&d.mgw, dtype, s.gw, B, T, C, 64, true))) return e;
            if ((e = make_ckpt_map(&d.mc, ck, B, T, H))) return e;
            d.ckpt = const_cast<void*>(ck);
            d.u = s.u; d.s0 = s.s0; d.gsT = s.gsT; d.gs0 = s.gs0; d.gu = s.gu;
            d.reverse = reverse[i];
        }
        if (recompute) {
            e = launch_fwd_chunk(pf, (cudaStream_t)stream);
            if (e) return cuda_fail(e, "wkv6 checkpoint recompute launch");
        }
        // few, long sequences: time-segmented sweep on the scratch behind the recomputed checkpoints (if any) in the workspace
        const size_t soff = recompute ? cb * nset : 0;
        e = -1;
        if (ws && aligned16(ws) && ws_bytes > soff)
            e = launch_bwd_chunk_segmented(p, reinterpret_cast<char*>(ws) + soff, ws_bytes - soff, (cudaStream_t)stream);
        // more sequences than SMs: balanced persistent grid with hand-over records in the same scratch
        if (e < 0 && ws && aligned16(ws) && ws_bytes > soff)
            e = launch_bwd_chunk_balanced(p, reinterpret_cast<char*>(ws) + soff, ws_bytes - soff, (cudaStream_t)stream);
        if (e < 0) e = launch_bwd_chunk(p, (cudaStream_t)stream);
        return cuda_fail(e, "wkv6 chunk backward launch");
    }
    const size_t per_dir = (size_t)B * H * N * sizeof(float);
    for (int i = 0; i < nset; i++) {
        const BwdSet& s = sets[i];
        DirArgs& d = p.d[i];
        const void* ptrs[10] = {s.r, s.k, s.v, s.w, s.gy, s.gr, s.gk, s.gv, s.gw, s.gu};
        for (const void* q : ptrs)
            if (!q || !aligned16(q)) return fail(WKV6_E_BADARG, "backward: NULL or misaligned tensor pointer");
        if (!s.u) return fail(WKV6_E_BADARG, "backward: u is NULL");
        if ((e = make_map(&d.mr, dtype, s.r, B, T, C, TT))) return e;
        if ((e = make_map(&d.mk, dtype, s.k, B, T, C, TT))) return e;
        if ((e = make_map(&d.mv, dtype, s.v, B, T, C, TT))) return e;
        if ((e = make_map(&d.mw, dtype, s.w, B, T, C, TT))) return e;
        if ((e = make_map(&d.mg, dtype, s.gy, B, T, C, TT))) return e;
        if ((e = make_map(&d.mq, dtype, s.gr, B, T, C, TT))) return e;
        d.u = s.u; d.s0 = s.s0; d.gsT = s.gsT; d.gs0 = s.gs0;
        d.gr = s.gr; d.gk = s.gk; d.gv = s.gv; d.gw = s.gw; d.gu = s.gu;
        d.reverse = reverse[i];
        d.hT = nullptr;
        d.grlo = nullptr;
        const size_t need = wkv6_b200_backward_workspace_bytes(dtype, B, T, C, H, nset);
        if (s.gsT || dtype == WKV6_DTYPE_BF16) {
            if (!ws || ws_bytes < need || !aligned16(ws))
                return fail(WKV6_E_WORKSPACE, "backward needs %zu workspace bytes (wkv6_b200_backward_workspace_bytes)", need);
            d.hT = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + per_dir * i);
            if (dtype == WKV6_DTYPE_BF16)
                d.grlo = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<char*>(ws) + per_dir * nset + (size_t)B * T * C * 2 * i);
        }
    }
    e = launch_bwd_scan(dtype, p, (cudaStream_t)stream);
    return cuda_fail(e, "wkv6 backward launch");
}

}  // namespace wkv6

using namespace wkv6;

extern "C" {

int wkv6_b200_abi_version(void) { return 4; }
/* Debug: enable != 0 allocates a device buffer that CTA 0 of the chunk kernels fills with clock64() stamps (forward:
 * slots 0..31, backward: 64..95) and, from slot 128 on, every CTA < 1024 of the forward chunk kernel with
 * {globaltimer at start, SM id, globaltimer at end, 0}; `out` (host, 128 entries) receives the first 128 slots;
 * wkv6_b200_debug_trace_ctas copies the per-CTA part (4 * 1024 entries). */
int wkv6_b200_debug_trace(int enable, unsigned long long* out) {
    if (enable && !g_dbg) {
        if (cudaMalloc(&g_dbg, (128 + 4096) * sizeof(unsigned long long)) != cudaSuccess) return -1;
        cudaMemset(g_dbg, 0, (128 + 4096) * sizeof(unsigned long long));
    }
    if (out && g_dbg) cudaMemcpy(out, g_dbg, 128 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    if (!enable && g_dbg) { cudaFree(g_dbg); g_dbg = nullptr; }
    return 0;
}
int wkv6_b200_debug_trace_ctas(unsigned long long* out) {
    if (!out || !g_dbg) return -1;
    return (int)cudaMemcpy(out, g_dbg + 128, 4096 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
}
size_t wkv6_b200_checkpoint_bytes(int B, int T, int C, int H) {
    (void)C;
    return (B <= 0 || T <= 0 || H <= 0) ? 0 : ckpt_bytes(B, T, H);
}
const char* wkv6_b200_last_error(void) { return g_err; }
long long wkv6_b200_launch_count(void) { return g_launches.load(); }
void wkv6_b200_set_algo(int algo) { g_algo.store(algo); }
int wkv6_b200_get_algo(void) { return g_algo.load(); }

size_t wkv6_b200_forward_scratch_bytes(int dtype, int B, int T, int C, int H, int ndirections) {
    if (check_common(dtype, B, T, C, H) || B == 0 || T == 0 || ndirections <= 0) return 0;
    Params p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.T = T; p.C = C; p.H = H; p.ndir = ndirections;
    return segment_scratch_bytes(p, use_chunk(dtype, T) ? 2 : 4);
}

int wkv6_b200_forward(int dtype, int B, int T, int C, int H, const void* r, const void* k, const void* v, const void* w,
                      const void* u, const float* s0, void* y, float* sT, void* ckpt, int reverse, void* scratch,
                      size_t scratch_bytes, void* stream) {
    const FwdSet s = {r, k, v, w, u, s0, y, sT, ckpt};
    const int rv = reverse != 0;
    return run_forward(dtype, B, T, C, H, &s, &rv, 1, scratch, scratch_bytes, stream);
}

int wkv6_b200_forward_static_scratch(int dtype, int B, int T, int C, int H, const void* r, const void* k, const void* v,
                                     const void* w, const void* u, const float* s0, void* y, float* sT, void* ckpt,
                                     int reverse, void* stream) {
    int e = check_common(dtype, B, T, C, H);
    if (e) return e;
    const size_t need = wkv6_b200_forward_scratch_bytes(dtype, B, T, C, H, 1);
    if (need > (size_t)WKV6_STATIC_SCRATCH_BYTES)
        return fail(WKV6_E_BADARG, "forward_static_scratch: this shape segments on %zu bytes of scratch, the static array holds "
                    "%zu (WKV6_STATIC_SCRATCH_BYTES): use wkv6_b200_forward with a caller-owned buffer", need,
                    (size_t)WKV6_STATIC_SCRATCH_BYTES);
    void* scratch = nullptr;
    if (need && B > 0) {
        if ((e = static_scratch_address(&scratch))) return e;
    }
    const FwdSet s = {r, k, v, w, u, s0, y, sT, ckpt};
    const int rv = reverse != 0;
    return run_forward(dtype, B, T, C, H, &s, &rv, 1, scratch, scratch ? (size_t)WKV6_STATIC_SCRATCH_BYTES : 0, stream);
}

size_t wkv6_b200_backward_scratch_bytes(int dtype, int B, int T, int C, int H, int ndirections) {
    if (check_common(dtype, B, T, C, H) || B == 0 || T == 0 || ndirections <= 0 || !use_chunk(dtype, T)) return 0;
    Params p;
    memset(&p, 0, sizeof(p));
    p.B = B; p.T = T; p.C = C; p.H = H; p.ndir = ndirections;
    return bwd_segment_scratch_bytes(p);
}

size_t wkv6_b200_backward_workspace_bytes(int dtype, int B, int T, int C, int H, int ndirections) {
    if (B <= 0 || H <= 0 || ndirections <= 0) return 0;
    size_t n = (size_t)ndirections * B * H * N * sizeof(float);                    /* hT  */
    if (dtype == WKV6_DTYPE_BF16 && T > 0) {
        n += (size_t)ndirections * B * T * C * 2;                                  /* grlo (exact scan)            */
        const size_t c = (size_t)ndirections * ckpt_bytes(B, T, H);                /* checkpoints (chunk backward) */
        if (c > n) n = c;
        n += wkv6_b200_backward_scratch_bytes(dtype, B, T, C, H, ndirections);     /* time-segmented sweep         */
    }
    return n;
}

int wkv6_b200_backward(int dtype, int B, int T, int C, int H, const void* r, const void* k, const void* v, const void* w,
                       const void* u, const float* s0, const void* gy, const float* gsT, const void* ckpt, void* gr,
                       void* gk, void* gv, void* gw, void* gu, float* gs0, int reverse, void* workspace,
                       size_t workspace_bytes, void* stream) {
    const BwdSet s = {r, k, v, w, u, s0, gy, gsT, gr, gk, gv, gw, gu, gs0, ckpt};
    const int rv = reverse != 0;
    return run_backward(dtype, B, T, C, H, &s, &rv, 1, workspace, workspace_bytes, stream);
}

int wkv6_b200_bi_forward(int dtype, int B, int T, int C, int H,
                         const void* r_f, const void* k_f, const void* v_f, const void* w_f, const void* u_f,
                         const float* s0_f, void* y_f, float* sT_f, void* ckpt_f,
                         const void* r_b, const void* k_b, const void* v_b, const void* w_b, const void* u_b,
                         const float* s0_b, void* y_b, float* sT_b, void* ckpt_b, int flags, void* scratch,
                         size_t scratch_bytes, void* stream) {
    FwdSet sets[2];
    int rv[2];
    int n = 0;
    if (flags & WKV6_BI_L2R) { sets[n] = {r_f, k_f, v_f, w_f, u_f, s0_f, y_f, sT_f, ckpt_f}; rv[n++] = 0; }
    if (flags & WKV6_BI_R2L) { sets[n] = {r_b, k_b, v_b, w_b, u_b, s0_b, y_b, sT_b, ckpt_b}; rv[n++] = 1; }
    return run_forward(dtype, B, T, C, H, sets, rv, n, scratch, scratch_bytes, stream);
}

int wkv6_b200_bi_backward(int dtype, int B, int T, int C, int H,
                          const void* r_f, const void* k_f, const void* v_f, const void* w_f, const void* u_f,
                          const float* s0_f, const void* gy_f, const float* gsT_f, const void* ckpt_f,
                          void* gr_f, void* gk_f, void* gv_f, void* gw_f, void* gu_f, float* gs0_f,
                          const void* r_b, const void* k_b, const void* v_b, const void* w_b, const void* u_b,
                          const float* s0_b, const void* gy_b, const float* gsT_b, const void* ckpt_b,
                          void* gr_b, void* gk_b, void* gv_b, void* gw_b, void* gu_b, float* gs0_b,
                          int flags, void* workspace, size_t workspace_bytes, void* stream) {
    BwdSet sets[2];
    int rv[2];
    int n = 0;
    if (flags & WKV6_BI_L2R) { sets[n] = {r_f, k_f, v_f, w_f, u_f, s0_f, gy_f, gsT_f, gr_f, gk_f, gv_f, gw_f, gu_f, gs0_f, ckpt_f}; rv[n++] = 0; }
    if (flags & WKV6_BI_R2L) { sets[n] = {r_b, k_b, v_b, w_b, u_b, s0_b, gy_b, gsT_b, gr_b, gk_b, gv_b, gw_b, gu_b, gs0_b, ckpt_b}; rv[n++] = 1; }
    return run_backward(dtype, B, T, C, H, sets, rv, n, workspace, workspace_bytes, stream);
}

int wkv6_b200_combine(int dtype, size_t n, const void* a, const void* b, float scale, int round_sum, float* out,
                      void* stream) {
    if (dtype != WKV6_DTYPE_BF16 && dtype != WKV6_DTYPE_F32) return fail(WKV6_E_BADDTYPE, "unknown dtype %d", dtype);
    if (n == 0) return 0;
    if (!a || !out || !aligned16(a) || !aligned16(out) || (b && !aligned16(b)) || (n % 8))
        return fail(WKV6_E_BADARG, "combine: NULL/misaligned pointer or n %% 8 != 0");
    return cuda_fail(launch_combine(dtype, n, a, b, scale, round_sum, out, (cudaStream_t)stream), "wkv6 combine launch");
}

int wkv6_b200_tmix_mix(int dtype, int B, int T, int C, const void* x, const void* x_edge, const void* m, const void* maa_x, const void* maa_r,
                       const void* maa_k, const void* maa_v, const void* maa_w, int reverse, void* xxx, void* xr, void* xk, void* xv,
                       void* xw, void* stream) {
    if (dtype != WKV6_DTYPE_BF16 && dtype != WKV6_DTYPE_F32) return fail(WKV6_E_BADDTYPE, "unknown dtype %d", dtype);
    if (B < 0 || T < 0 || C <= 0 || C % 8) return fail(WKV6_E_BADARG, "tmix_mix: bad shape B=%d T=%d C=%d (C %% 8 == 0)", B, T, C);
    if (B == 0 || T == 0) return 0;
    const void* maa[5] = {maa_x, maa_r, maa_k, maa_v, maa_w};
    void* out[4] = {m ? xr : xxx, xk, xv, xw};
    if (!x || !aligned16(x) || (m && !aligned16(m)) || (x_edge && !aligned16(x_edge)))
        return fail(WKV6_E_BADARG, "tmix_mix: NULL/misaligned x, x_edge or m");
    if (!m) {
        if (!maa_x || !xxx || !aligned16(maa_x) || !aligned16(xxx)) return fail(WKV6_E_BADARG, "tmix_mix stage 1: NULL/misaligned maa_x or xxx");
    } else {
        for (int i = 0; i < 4; i++)
            if (!maa[i + 1] || !out[i] || !aligned16(maa[i + 1]) || !aligned16(out[i]))
                return fail(WKV6_E_BADARG, "tmix_mix stage 2: NULL/misaligned maa or output %d", i);
    }
    return cuda_fail(launch_tmix_mix(dtype, B, T, C, x, x_edge, m, maa, reverse, out, (cudaStream_t)stream), "wkv6 tmix_mix launch");
}

int wkv6_b200_ln_pair_cat(int dtype, long long rows, int C, const void* y_f, const void* y_b, const void* w_f, const void* b_f,
                          const void* w_b, const void* b_b, float eps, void* out, void* stream) {
    if (dtype != WKV6_DTYPE_BF16 && dtype != WKV6_DTYPE_F32) return fail(WKV6_E_BADDTYPE, "unknown dtype %d", dtype);
    if (rows < 0 || C <= 0 || C % 8 || C > 2048) return fail(WKV6_E_BADARG, "ln_pair_cat: bad shape rows=%lld C=%d (C %% 8 == 0, C <= 2048)", rows, C);
    if (rows == 0) return 0;
    const void* ptrs[7] = {y_f, y_b, w_f, b_f, w_b, b_b, out};
    for (const void* q : ptrs)
        if (!q || !aligned16(q)) return fail(WKV6_E_BADARG, "ln_pair_cat: NULL or misaligned pointer");
    return cuda_fail(launch_ln_pair_cat(dtype, rows, C, y_f, y_b, w_f, b_f, w_b, b_b, eps, out, (cudaStream_t)stream), "wkv6 ln_pair_cat launch");
}

int wkv6_b200_tmix_mix_backward(int dtype, int B, int T, int C, const void* x, const void* m, const void* maa0, const void* maa1,
                                const void* maa2, const void* maa3, const void* g0, const void* g1, const void* g2,
                                const void* g3, int reverse, void* gx, void* gm, float* gmaa, void* stream) {
    if (dtype != WKV6_DTYPE_BF16 && dtype != WKV6_DTYPE_F32) return fail(WKV6_E_BADDTYPE, "unknown dtype %d", dtype);
    if (B < 0 || T < 0 || C <= 0 || C % 8) return fail(WKV6_E_BADARG, "tmix_mix_backward: bad shape B=%d T=%d C=%d (C %% 8 == 0)", B, T, C);
    if (B == 0 || T == 0) return 0;
    const void* maa[4] = {maa0, maa1, maa2, maa3};
    const void* g[4] = {g0, g1, g2, g3};
    const int np = m ? 4 : 1;
    if (!x || !gx || !gmaa || !aligned16(x) || !aligned16(gx) || (m && !aligned16(m)) || (gm && !aligned16(gm)))
        return fail(WKV6_E_BADARG, "tmix_mix_backward: NULL/misaligned x, gx, gmaa, m or gm");
    for (int i = 0; i < np; i++)
        if (!maa[i] || !g[i] || !aligned16(maa[i]) || !aligned16(g[i]))
            return fail(WKV6_E_BADARG, "tmix_mix_backward: NULL/misaligned maa or gradient %d", i);
    return cuda_fail(launch_tmix_mix_bwd(dtype, B, T, C, x, m, maa, g, reverse, gx, gm, gmaa, (cudaStream_t)stream),
                     "wkv6 tmix_mix_backward launch");
}

int wkv6_b200_convmod_core(int dtype, int B, int Tin, int Tout, int C, int K, int left, const void* z, const float* dw_weight,
                           const float* dw_bias, const float* ln_weight, const float* ln_bias, float eps, void* y, void* stream) {
    if (dtype != WKV6_DTYPE_BF16 && dtype != WKV6_DTYPE_F32) return fail(WKV6_E_BADDTYPE, "unknown dtype %d", dtype);
    if (B < 0 || Tin < 0 || Tout < 0 || left < 0) return fail(WKV6_E_BADARG, "convmod_core: bad sizes B=%d Tin=%d Tout=%d left=%d", B, Tin, Tout, left);
    if (C < 64 || C > 1024 || C % 64) return fail(WKV6_E_BADSHAPE, "convmod_core: C (%d) must be a multiple of 64 in [64, 1024]", C);
    if (K != 15 && K != 31) return fail(WKV6_E_BADSHAPE, "convmod_core: kernel size %d (15 and 31 are built)", K);
    if (B == 0 || Tout == 0) return 0;
    if (!z || !y || !dw_weight || !ln_weight || !ln_bias || !aligned16(z) || !aligned16(y))
        return fail(WKV6_E_BADARG, "convmod_core: NULL or misaligned pointer");
    return cuda_fail(launch_convmod_core(dtype, B, Tin, Tout, C, K, left, z, dw_weight, dw_bias, ln_weight, ln_bias, eps, y,
                                         (cudaStream_t)stream), "convmod_core launch");
}

}  // extern "C"
