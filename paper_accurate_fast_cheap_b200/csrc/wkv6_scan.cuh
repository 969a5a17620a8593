// wkv6_scan.cuh -- parameter blocks shared by the kernels and the C ABI (wkv6_capi.cu).
#pragma once
#include <atomic>
#include "wkv6_common.cuh"

namespace wkv6 {

constexpr int TT = 16;    // time steps per TMA tile of the exact scan kernels
constexpr int NT = 128;   // threads per CTA of the exact scan kernels (one CTA per (direction, b, h))

// One direction ("set") of a call.  Tensor maps are 3-D views [B][T][C] with box [1][TT][64].
struct alignas(64) DirArgs {
    CUtensorMap mr, mk, mv, mw, mg, mq;   // r, k, v, w, gy, gr(as input of pass B)
    CUtensorMap my;                       // y (TMA store; chunk kernel only)
    CUtensorMap mc;                       // chunk-start state checkpoints, 2-D [B*H*nchunks*64][64] bf16 (chunk kernels)
    CUtensorMap mgr, mgk, mgv, mgw;       // gradient outputs (TMA store; chunk backward only)
    void* ckpt;                           // null: the forward does not save checkpoints
    const void* u;                        // [H,64] io dtype
    const float* s0;                      // [B,H,64,64] or null
    float* sT;                            // [B,H,64,64] or null
    void* y;                              // [B,T,C]
    const float* gsT;                     // bwd: gradient into S_T or null
    float* gs0;                           // bwd: gradient w.r.t. S_0 or null
    void *gr, *gk, *gv, *gw;              // [B,T,C]
    void* gu;                             // [B,C]
    float* hT;                            // workspace [B*H*64]: sum_i S_T[i][j]*gsT[i][j]  (pass A -> pass B)
    __nv_bfloat16* grlo;                  // workspace [B,T,C] (bf16 I/O only): gr_fp32 - float(bf16(gr)), so that
                                          // sweep B sees gr to ~16 mantissa bits for the gw suffix sum
    int reverse;                          // scan t = T-1..0
    int pad_;
};

struct Params {
    DirArgs d[2];
    int B, T, C, H;
    int ndir;
    int skip_y;                           // chunk forward: compute states/checkpoints only (backward recompute)
    // time-segmented forward (long sequences, few (batch, head) pairs; wkv6_chunk_lean.cu): 0 = one CTA per sequence
    int nseg, seg_chunks;                 // segments per sequence, 64-step chunks per segment
    int seg_mode;                         // 1: pass 1 (local state + decay product of every segment), 2: pass 2 (outputs)
    float* seg_state;                     // [ndir][B*H][nseg][64 i][64 j]: pass 1 writes dS, the scan turns it into S_start
    float* seg_decay;                     // [ndir][B*H][nseg][64 j]: product of the per-step decays of the segment
    float* seg_gu;                        // time-segmented backward: [ndir][B*H][nseg][64] partial gu of every segment
    int corr_first;                       // one-pass segmented forward: first chunk wkv6_fwd_corr_kernel touches
    int corr_pp, corr_pc;                 // its CTAs: pieces per segment, chunks per piece
    float* seg_prefix;                    // one-pass segmented forward: [ndir][B*H][chunks][64] decay from segment start to chunk start
    // balanced persistent backward (wkv6_chunk_bwd.cu): CTAs per direction (0: off), hand-over records, flag value of the launch
    int bal_ncta;
    unsigned bal_magic_n, bal_magic_h;    // 2^31 / chunks-per-sequence + 1, 2^31 / H + 1 (every chunk-backward launch)
    unsigned char* bal_rec;
    unsigned long long bal_epoch;
    unsigned long long* dbg;              // optional clock64() trace of one CTA (wkv6_b200_debug_trace), else null
};

// launchers (wkv6_scan.cu / wkv6_chunk_lean.cu / wkv6_chunk_bwd.cu); return cudaError_t as int
int launch_fwd_scan(int dtype, const Params& p, cudaStream_t st);
int launch_bwd_scan(int dtype, const Params& p, cudaStream_t st);
int launch_fwd_chunk(const Params& p, cudaStream_t st);   // bf16 only, tcgen05
int launch_tmix_mix(int dtype, int B, int T, int C, const void* x, const void* x_edge, const void* m, const void* const* maa, int reverse,
                    void* const* out, cudaStream_t st);
bool plan_segments(Params& q, void* scratch, size_t scratch_bytes, int min_seg);
size_t segment_scratch_bytes(const Params& p, int min_seg);
int launch_seg_scan(const Params& q, cudaStream_t st);
int launch_fwd_scan_segmented(int dtype, const Params& p, void* scratch, size_t scratch_bytes, cudaStream_t st);
int launch_fwd_chunk_segmented(const Params& p, void* scratch, size_t scratch_bytes, cudaStream_t st);   // -1: not applicable
int launch_bwd_chunk(const Params& p, cudaStream_t st);   // bf16 only, tcgen05; needs DirArgs::ckpt
int launch_bwd_chunk_segmented(const Params& p, void* scratch, size_t scratch_bytes, cudaStream_t st);   // -1: not applicable
int launch_bwd_chunk_balanced(const Params& p, void* scratch, size_t scratch_bytes, cudaStream_t st);    // -1: not applicable
size_t bwd_segment_scratch_bytes(const Params& p);
int launch_tmix_mix_bwd(int dtype, int B, int T, int C, const void* x, const void* m, const void* const* maa, const void* const* g,
                        int reverse, void* gx, void* gm, float* gmaa, cudaStream_t st);
int launch_ln_pair_cat(int dtype, long long rows, int C, const void* y_f, const void* y_b, const void* w_f, const void* b_f,
                       const void* w_b, const void* b_b, float eps, void* out, cudaStream_t st);
int launch_convmod_core(int dtype, int B, int Tin, int Tout, int C, int K, int left, const void* z, const float* dw_w,
                        const float* dw_b, const float* ln_w, const float* ln_b, float eps, void* y, cudaStream_t st);
int launch_combine(int dtype, size_t n, const void* a, const void* b, float scale, int round_sum, float* out, cudaStream_t st);

void count_launch(int n);
unsigned long long* debug_trace_buffer();

// clock64() of CTA 0 at chunk `WKV6_TS_CHUNK`: slots 0..15 thread 0 (MMA issuer), 16..31 thread 32
#define WKV6_TS_CHUNK 4
// needs `const bool wkv6_trace_on = WKV6_TRACE_ON(p, tid);` in scope (one predicate per thread instead of five tests per
// stamp: the stamps were 3.7 % of the executed instructions of the forward kernel)
#define WKV6_TRACE_ON(p, tid) ((p).dbg != nullptr && blockIdx.x == 0 && ((tid) == 0 || (tid) == 32))
// The in-loop phase stamps are compiled in only with -DWKV6_TRACE_PHASES=1 (python -m paper_accurate_fast_cheap_b200.build
// --trace; scripts/trace_phases.py asks for it): 17 predicated branches per chunk were 3 % of the backward kernel's
// instructions.  The per-CTA start / end stamps (outside the chunk loops) are always there.
#ifndef WKV6_TRACE_PHASES
#define WKV6_TRACE_PHASES 0
#endif
#if WKV6_TRACE_PHASES
#define WKV6_TS(base, i)                                                                          \
    do {                                                                                          \
        if (wkv6_trace_on && c_ts == WKV6_TS_CHUNK) p.dbg[(base) + (tid ? 16 : 0) + (i)] = clock64(); \
    } while (0)
#else
#define WKV6_TS(base, i) do { (void)wkv6_trace_on; (void)c_ts; } while (0)
#endif


// Launch with the programmatic-stream-serialization attribute (see pdl_wait() in wkv6_common.cuh).
template <typename Kernel>
inline cudaError_t launch_pdl(Kernel kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const Params& q) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, q);
}

// The dynamic-shared-memory opt-in is a per-DEVICE function attribute: one bit per (kernel slot, device), so that a
// process driving several GPUs opts in on each of them (and pays the driver call once per device, not per launch).
template <typename Kernel>
inline cudaError_t opt_in_smem_once(Kernel kernel, size_t bytes, int slot) {
    static std::atomic<unsigned long long> done[4];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const unsigned long long bit = 1ull << (dev & 63);
    if (done[slot].load(std::memory_order_acquire) & bit) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return e;
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    done[slot].fetch_or(bit, std::memory_order_release);
    return cudaSuccess;
}

}  // namespace wkv6
