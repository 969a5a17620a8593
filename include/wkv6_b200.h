/*
 * wkv6_b200.h -- C ABI of libwkv6_b200.so: the B200 (sm_100a) implementation of the
 * bidirectional RWKV-6 WKV6 scan (forward, reversed, backward pass).
 *
 * This is the drop-in boundary.  Every entry point replaces one host launcher of the
 * reference (paths relative to the reference tree):
 *
 *   wkv6_b200_forward        <- cuda_forward / cuda_forward_fp32 / cuda_forward_fp64
 *                               wenet/rwkv_v6/cuda/wkv6_cuda.cu:265-270, 282-287, 299-304
 *                               (+ initial state: wkv6state_cuda.cu:7-65, 296-300)
 *   wkv6_b200_backward       <- cuda_backward / cuda_backward_fp32 / cuda_backward_fp64
 *                               wenet/rwkv_v6/cuda/wkv6_cuda.cu:272-280, 289-297, 306-314
 *                               (+ gs: wkv6state_cuda.cu:68-293, 302-308)
 *   wkv6_b200_bi_forward     <- the two cuda_forward launches (and the two torch.flip copies
 *   wkv6_b200_bi_backward       around them) issued per layer by
 *                               wenet/rwkv_v6/rwkv_wrapper_bidirectional.py:43-49
 *   wkv6_b200_combine        <- `(x_att_forward + x_att_backward_fliped) / 2` + `.float()`
 *                               wenet/rwkv_v6/rwkv_wrapper_bidirectional.py:48-56
 *
 * Conventions
 *   - All tensor pointers are DEVICE pointers, contiguous, 16-byte aligned.
 *     r,k,v,w,y,gy,gr,gk,gv,gw: [B,T,C] in `dtype`; u: [H,N] in `dtype`; gu: [B,C] in `dtype`
 *     (per-batch partials, the caller sums over B exactly like model.py:151);
 *     N = C / H must be 64.
 *   - States are fp32: s0, sT, gs0, gsT: [B,H,N,N], index [b][h][i_value][j_key]
 *     (layout of wkv6state_cuda.cu:15,24).  Any of them may be NULL
 *     (s0 = NULL means S_0 = 0; gsT = NULL means no gradient flows into S_T).
 *   - `reverse` != 0 runs the recurrence for t = T-1 .. 0 and writes every output at its
 *     natural (un-flipped) position: identical to flip -> scan -> flip over the padded time
 *     axis, without the copies.
 *   - `stream` is a cudaStream_t (NULL = legacy default stream).  Calls are asynchronous.
 *   - The library keeps no per-call state between entry points: every buffer a launch uses (outputs, checkpoints,
 *     backward workspace, forward scratch) is an argument, so concurrent calls on different streams / devices / host
 *     threads are independent.  Process-wide: the launch counter, the bf16 arithmetic selector (wkv6_b200_set_algo)
 *     and the debug trace buffer.
 *   - Return value: 0 on success, a negative WKV6_E_* code for rejected arguments,
 *     a positive cudaError_t for CUDA failures.  wkv6_b200_last_error() gives the text.
 *   - No CPU fallback exists: without a CUDA device every compute entry point fails.
 */
#ifndef WKV6_B200_H_
#define WKV6_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WKV6_DTYPE_BF16 0
#define WKV6_DTYPE_F32  1

#define WKV6_E_BADARG   (-1)   /* NULL pointer, non-positive size, misaligned pointer          */
#define WKV6_E_BADSHAPE (-2)   /* C != H*64                                                   */
#define WKV6_E_BADDTYPE (-3)
#define WKV6_E_NODRIVER (-4)   /* cuTensorMapEncodeTiled not obtainable from the driver       */
#define WKV6_E_WORKSPACE (-5)  /* workspace too small                                         */

/* flags of the bidirectional entry points */
#define WKV6_BI_L2R      1     /* run the left-to-right direction (set "f")                   */
#define WKV6_BI_R2L      2     /* run the right-to-left direction (set "b"), flip-free        */

/* Arithmetic selector for bf16 I/O (f32 I/O always uses the exact scan): AUTO = CHUNK = the tcgen05 chunk kernels (the
 * only bf16 production path); SCAN = the exact fp32-math recurrence on bf16 I/O, a documented numerics option (it is the
 * fp32 kernel instantiated for bf16 tensors, not a second implementation of the chunk path). */
#define WKV6_ALGO_AUTO   0
#define WKV6_ALGO_SCAN   1     /* fp32 recurrent scan on the FMA pipe (exact form)            */
#define WKV6_ALGO_CHUNK  2     /* chunked dense-contraction form on tcgen05/TMEM              */

int  wkv6_b200_abi_version(void);
const char* wkv6_b200_last_error(void);
/* Number of kernels this library has launched since load (the bench's gpu_launches). */
long long wkv6_b200_launch_count(void);
/* Debug aid: clock64() trace of one CTA of the chunk kernels (see csrc/wkv6_capi.cu); not part of the data path. */
int  wkv6_b200_debug_trace(int enable, unsigned long long* out);
int  wkv6_b200_debug_trace_ctas(unsigned long long* out);   /* 4*1024 entries: per-CTA {t_start, smid, t_end, 0} */
void wkv6_b200_set_algo(int algo);
int  wkv6_b200_get_algo(void);

/* Bytes of one direction's checkpoint buffer (bf16 state at the start of every 64-step chunk):
 * B*H*ceil(T/64)*64*64*2.  A forward that is given `ckpt` (bf16 I/O, tcgen05 chunk kernel) fills it; a backward
 * that is given the same buffer skips the state recomputation.  The reference keeps nothing between its forward
 * and backward launches (model.py:131,149), so NULL is always accepted. */
size_t wkv6_b200_checkpoint_bytes(int B, int T, int C, int H);

/* Time segmentation of the forward (optional).  One CTA per (direction, batch, head) leaves most of the 148 SMs idle when
 * a GPU holds few, long sequences (directions * B * H <= 148 for bf16, <= 74 for f32; T >= 2048: one to eight 5-minute
 * utterances, a 1-hour utterance).  Given `scratch` (device memory, caller-owned, 16-byte aligned, used only by this call
 * on `stream`) of at least wkv6_b200_forward_scratch_bytes(...) bytes, such a forward runs in 2..40 time segments per
 * sequence over all SMs (three launches: local states, segment scan, outputs); same results to the dtype's tolerance.
 * The function returns 0 for shapes that would not be segmented; scratch == NULL always means one CTA per sequence.
 * The reference has no counterpart: it runs one thread block per (batch, head) whatever T is (wkv6_cuda.cu:269). */
size_t wkv6_b200_forward_scratch_bytes(int dtype, int B, int T, int C, int H, int ndirections);

int wkv6_b200_forward(int dtype, int B, int T, int C, int H,
                      const void* r, const void* k, const void* v, const void* w, const void* u,
                      const float* s0, void* y, float* sT, void* ckpt, int reverse,
                      void* scratch, size_t scratch_bytes, void* stream);

/* wkv6_b200_forward for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc): the same kernels on the same arguments, with the time-segmentation scratch taken
 * from a fixed array inside the library (a file-scope `static __device__` array of WKV6_STATIC_SCRATCH_BYTES, which
 * every GPU that loads the library holds until exit) instead of a caller-owned buffer.  The call allocates and frees
 * nothing.  A shape whose wkv6_b200_forward_scratch_bytes() exceeds WKV6_STATIC_SCRATCH_BYTES is rejected with
 * WKV6_E_BADARG (wkv6_b200_last_error() names both sizes); a shape that would not be segmented runs one CTA per sequence
 * and does not touch the array.  Nothing in the array has to start at zero (every element a launch reads was written
 * earlier in the same call), so it is not cleared, and it keeps the last call's contents.  ALL calls on a device share
 * the one array: calls that can overlap on the GPU (different streams without an ordering between them, or different
 * host threads) race on it -- keep them on one stream or order them with events.
 * No counterpart in the reference. */
#define WKV6_STATIC_SCRATCH_BYTES (16u << 20)
int wkv6_b200_forward_static_scratch(int dtype, int B, int T, int C, int H,
                                     const void* r, const void* k, const void* v, const void* w, const void* u,
                                     const float* s0, void* y, float* sT, void* ckpt, int reverse, void* stream);

/* Bytes of workspace the backward entry points need when they are given no checkpoints (0 is possible). */
size_t wkv6_b200_backward_workspace_bytes(int dtype, int B, int T, int C, int H, int ndirections);
/* Time segmentation of the bf16 backward (optional), the counterpart of the forward's: with few, long sequences
 * (directions * B * H <= 74, T >= 2048: long-form fine-tuning) and a workspace of at least this many bytes a backward that
 * IS given checkpoints sweeps every sequence in 2..40 time segments over all SMs (four launches: segment-local reverse
 * states, segment scan, gradients, gu reduction); same results to the dtype's tolerance.  Returns 0 for shapes that would
 * not be segmented; wkv6_b200_backward_workspace_bytes() includes these bytes.  The reference cannot run these shapes at
 * all: its backward indexes a ctx_len-sized array (wkv6_cuda.cu:207).
 * The same bytes serve the opposite case, MORE sequences than SMs (directions * B * H > SM count: the training batch):
 * given them, the backward runs a balanced persistent grid -- one CTA per SM, every direction's (sequence, 64-step chunk)
 * units cut into equal ranges, a sequence that is cut between two CTAs hands its reverse state and partial gu over through
 * a 16.5 KB record in this scratch ((CTAs + 1) records, 2.5 MB on a B200) -- instead of one CTA per sequence in several
 * waves.  One launch either way; contents of the scratch on entry are arbitrary, calls that overlap on the GPU must not
 * share it.  Gradients are those of the one-CTA-per-sequence launch bit for bit, except gu (its partial sums regroup). */
size_t wkv6_b200_backward_scratch_bytes(int dtype, int B, int T, int C, int H, int ndirections);

int wkv6_b200_backward(int dtype, int B, int T, int C, int H,
                       const void* r, const void* k, const void* v, const void* w, const void* u,
                       const float* s0, const void* gy, const float* gsT, const void* ckpt,
                       void* gr, void* gk, void* gv, void* gw, void* gu, float* gs0,
                       int reverse, void* workspace, size_t workspace_bytes, void* stream);

/* Both directions in ONE launch per pass.  Set "f" scans t = 0..T-1, set "b" scans
 * t = T-1..0 (flip-free).  `flags` selects which sets run (direction dropout): a set whose
 * bit is clear is neither read nor written. */
int wkv6_b200_bi_forward(int dtype, int B, int T, int C, int H,
                         const void* r_f, const void* k_f, const void* v_f, const void* w_f, const void* u_f,
                         const float* s0_f, void* y_f, float* sT_f, void* ckpt_f,
                         const void* r_b, const void* k_b, const void* v_b, const void* w_b, const void* u_b,
                         const float* s0_b, void* y_b, float* sT_b, void* ckpt_b,
                         int flags, void* scratch, size_t scratch_bytes, void* stream);

int wkv6_b200_bi_backward(int dtype, int B, int T, int C, int H,
                          const void* r_f, const void* k_f, const void* v_f, const void* w_f, const void* u_f,
                          const float* s0_f, const void* gy_f, const float* gsT_f, const void* ckpt_f,
                          void* gr_f, void* gk_f, void* gv_f, void* gw_f, void* gu_f, float* gs0_f,
                          const void* r_b, const void* k_b, const void* v_b, const void* w_b, const void* u_b,
                          const float* s0_b, const void* gy_b, const float* gsT_b, const void* ckpt_b,
                          void* gr_b, void* gk_b, void* gv_b, void* gw_b, void* gu_b, float* gs0_b,
                          int flags, void* workspace, size_t workspace_bytes, void* stream);

/* out[i] = (a[i] + b[i]) * scale as fp32, a/b in `dtype`; b may be NULL (out = a*scale).
 * n = number of elements.  round_sum != 0 (bf16 only): the sum is rounded to bf16 (RNE) before the
 * scale, bit-identical to the reference's `(x_f + x_b) / 2` in bf16 followed by `.float()`. */
int wkv6_b200_combine(int dtype, size_t n, const void* a, const void* b, float scale, int round_sum,
                      float* out, void* stream);

/* Elementwise part of the time-mix prologue: token shift + "maa" mixes of RWKV_Tmix_x060c.forward,
 * wenet/rwkv_v6/src/model.py:270-289, one pass per stage instead of ~14 eager kernels, bit-exact with them.
 *   xx = shift(x) - x,  shift = x[t-1] (reverse = 0) or x[t+1] (reverse = 1: the flip-free right-to-left block); beyond the
 *   edge of the sequence: x_edge[b] ([B][C]: streaming, the last frame of the previous piece) or 0 when x_edge == NULL
 *   stage 1 (m == NULL):  xxx = x + xx * maa_x
 *   stage 2 (m != NULL):  xr = x + xx * (maa_r + m[0]), xk (m[1]), xv (m[2]), xw (m[3]);   m is [4][B][T][C] contiguous
 * x and outputs [B][T][C] contiguous in `dtype`, maa_* [C] in `dtype`; C % 8 == 0; 16-byte aligned pointers. */
int wkv6_b200_tmix_mix(int dtype, int B, int T, int C, const void* x, const void* x_edge, const void* m, const void* maa_x, const void* maa_r,
                       const void* maa_k, const void* maa_v, const void* maa_w, int reverse, void* xxx, void* xr, void* xk,
                       void* xv, void* xw, void* stream);

/* The two LayerNorms of a bidirectional block's epilogue (RWKV_Tmix_x060c.ln_x of each direction, model.py:323), written
 * side by side: out[row] = [ ln_f(y_f[row]) || ln_b(y_b[row]) ], out is [rows][2C].  With it the two output projections,
 * the add and the /2 of rwkv_wrapper_bidirectional.py:44-56 are ONE GEMM out x [W_f ; W_b]^T / 2 (library call).
 * y_*, weights / biases [C] and out in `dtype`; C % 8 == 0, C <= 2048; statistics in fp32. */
int wkv6_b200_ln_pair_cat(int dtype, long long rows, int C, const void* y_f, const void* y_b, const void* w_f, const void* b_f,
                          const void* w_b, const void* b_b, float eps, void* out, void* stream);

/* Backward of one stage of wkv6_b200_tmix_mix (training: the fused prologue with autograd on).  With xx = shift(x) - x:
 *   stage 1 (m == NULL): g0 = gradient of xxx, maa0 = maa_x;   stage 2 (m != NULL): g0..g3 = gradients of xr, xk, xv, xw,
 *   maa0..maa3 = maa_r, maa_k, maa_v, maa_w, m as in the forward.
 *   gx   [B][T][C]    = gradient w.r.t. x through this stage (sum over the stage's outputs, including the shifted use of x)
 *   gm   [4][B][T][C] = gradient w.r.t. m (stage 2; may be NULL)
 *   gmaa [1 or 4][C]  = gradient w.r.t. the maa vectors, fp32, ACCUMULATED with atomics: zero-fill before the call
 * x_edge of the forward is treated as a constant.  fp32 arithmetic, one rounding per stored value (the reference's eager
 * autograd rounds after every elementwise op). */
int wkv6_b200_tmix_mix_backward(int dtype, int B, int T, int C, const void* x, const void* m, const void* maa0, const void* maa1,
                                const void* maa2, const void* maa3, const void* g0, const void* g1, const void* g2,
                                const void* g3, int reverse, void* gx, void* gm, float* gmaa, void* stream);

/* The block that follows the time-mix in every encoder layer: the bandwidth-bound middle of the Conformer
 * ConvolutionModule, wenet/transformer/convolution.py:123-137 -- GLU, depthwise Conv1d (+ bias), LayerNorm over the
 * channels, swish -- in one pass over the [B][T][C] layout (no transposes), between the two pointwise convolutions
 * (GEMMs, :122 and :138, which stay library calls):
 *   g[t][c]   = z[t][c] * sigmoid(z[t][C + c])                          z: [B][Tin][2C] in `dtype` = pointwise_conv1(x)
 *   d[t][c]   = dw_bias[c] + sum_k dw_weight[c][k] * g[t + k - left][c]  (g = 0 outside [0, Tin))
 *   y[t][c]   = swish(ln_weight[c] * (d[t][c] - mean_t) / sqrt(var_t + eps) + ln_bias[c]),   t in [0, Tout)
 * Non-causal (reference padding (K-1)/2): Tin = Tout = T, left = (K-1)/2.  Causal with the lorder = K-1 frames of left
 * context (zeros or the cache) already in front of x (:111-117): Tin = T + K - 1, Tout = T, left = 0.
 * dw_weight [C][K], dw_bias [C] (may be NULL), ln_weight / ln_bias [C]: fp32.  K is 15 or 31; C % 64 == 0, 64 <= C <= 1024;
 * fp32 arithmetic for either I/O dtype. */
int wkv6_b200_convmod_core(int dtype, int B, int Tin, int Tout, int C, int K, int left, const void* z,
                           const float* dw_weight, const float* dw_bias, const float* ln_weight, const float* ln_bias,
                           float eps, void* y, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* WKV6_B200_H_ */
