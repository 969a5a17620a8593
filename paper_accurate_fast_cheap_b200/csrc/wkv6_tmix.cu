// wkv6_tmix.cu -- the elementwise part of the time-mix prologue that feeds the WKV6 scan (first step):
// token shift and the five "maa" mixes of RWKV_Tmix_x060c.forward (reference: wenet/rwkv_v6/src/model.py:270-289),
//
//     xx  = shift(x) - x                    shift = x[t-1] (left-to-right block) or x[t+1] (flip-free right-to-left block);
//                                           outside the sequence: zero, or the caller's edge frame (streaming: the last
//                                           frame of the previous piece)
//     stage 1:  xxx = x + xx * maa_x                                            (input of the LoRA GEMM)
//     stage 2:  x_c = x + xx * (maa_c + m_c)   for c in {r, k, v, w}            (inputs of the r / k / v / decay GEMMs)
//
// which the reference runs as ~14 separate elementwise kernels (pad, sub, 5 x {add, mul, add}).  Here each stage is ONE
// pass: stage 1 reads x once (the shifted row comes from L2) and writes xxx; stage 2 reads x and the four m_c and writes
// the four x_c: 2 and 9 tensor passes instead of ~8 and ~36.  HBM-bound by construction; every access is a 16-byte
// vector, consecutive threads touch consecutive vectors of a row.
//
// Bit-exact with the reference's eager ops: every intermediate the reference materialises is rounded to the I/O type
// here too (bf16: packed sub / add / mul, one rounding per operation like ATen's bf16 elementwise kernels; fp32: plain
// IEEE ops, no FMA contraction).
#include "wkv6_scan.cuh"

namespace wkv6 {

namespace {

template <typename T> struct Vec;
template <> struct Vec<__nv_bfloat16> { static constexpr int W = 8; };   // elements per 16-byte vector
template <> struct Vec<float> { static constexpr int W = 4; };

struct TmixArgs {
    const void* x;
    const void* x_edge;   // [B][C] frame beyond the edge (x[-1], or x[T] when reverse), or null: zeros
    const void* m;        // [4][B*T*C] (r, k, v, w) or null: stage 1
    const void* maa[5];   // x, r, k, v, w: [C]
    void* out[4];         // stage 1: out[0] = xxx; stage 2: x_r, x_k, x_v, x_w
    long long rows;       // B*T
    int T, C, reverse;
};

// One 16-byte vector of the elementwise chain.  bf16: packed bf16x2 instructions (sub / add / mul round to nearest even
// once per operation, which is what the eager ops produce: ATen computes each op in fp32 and rounds the result to bf16, and
// for bf16 operands the fp32 intermediate is exact or far from a bf16 tie, so the two roundings agree); fp32: plain IEEE
// operations without FMA contraction.
template <typename T> struct Ops;
template <> struct Ops<__nv_bfloat16> {
    using V = uint4;
    static __device__ __forceinline__ uint32_t sub2(uint32_t a, uint32_t b) { uint32_t d; asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
    static __device__ __forceinline__ uint32_t add2(uint32_t a, uint32_t b) { uint32_t d; asm("add.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
    static __device__ __forceinline__ uint32_t mul2(uint32_t a, uint32_t b) { uint32_t d; asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
    static __device__ __forceinline__ V zero() { return make_uint4(0, 0, 0, 0); }
    static __device__ __forceinline__ V sub(V a, V b) { return make_uint4(sub2(a.x, b.x), sub2(a.y, b.y), sub2(a.z, b.z), sub2(a.w, b.w)); }
    static __device__ __forceinline__ V add(V a, V b) { return make_uint4(add2(a.x, b.x), add2(a.y, b.y), add2(a.z, b.z), add2(a.w, b.w)); }
    static __device__ __forceinline__ V mul(V a, V b) { return make_uint4(mul2(a.x, b.x), mul2(a.y, b.y), mul2(a.z, b.z), mul2(a.w, b.w)); }
};
template <> struct Ops<float> {
    using V = float4;
    static __device__ __forceinline__ V zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
    static __device__ __forceinline__ V sub(V a, V b) { return make_float4(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z), __fsub_rn(a.w, b.w)); }
    static __device__ __forceinline__ V add(V a, V b) { return make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), __fadd_rn(a.w, b.w)); }
    static __device__ __forceinline__ V mul(V a, V b) { return make_float4(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y), __fmul_rn(a.z, b.z), __fmul_rn(a.w, b.w)); }
};

// Thread block = (vx, vy): vx threads along the vectors of a row, vy rows; grid-stride over row groups.  One 32-bit
// modulo per thread and row (t = row % T), no other index arithmetic.
template <typename T, bool STAGE2>
__global__ void __launch_bounds__(256) wkv6_tmix_mix_kernel(const TmixArgs a) {
    using O = Ops<T>;
    using V = typename O::V;
    constexpr int W = Vec<T>::W;
    const int vpr = a.C / W;                                   // vectors per row
    const T* x = static_cast<const T*>(a.x);
    const size_t plane = (size_t)a.rows * a.C;
    for (long long row = (long long)blockIdx.x * blockDim.y + threadIdx.y; row < a.rows; row += (long long)gridDim.x * blockDim.y) {
        const int t = (int)((unsigned long long)row % (unsigned)a.T);
        const bool edge = a.reverse ? (t == a.T - 1) : (t == 0);
        for (int vi = threadIdx.x; vi < vpr; vi += blockDim.x) {
            const int c = vi * W;
            const size_t off = (size_t)row * a.C + c;
            const V xv = *reinterpret_cast<const V*>(x + off);
            V xs;
            if (!edge) xs = *reinterpret_cast<const V*>(a.reverse ? x + off + a.C : x + off - a.C);
            else if (a.x_edge) xs = *reinterpret_cast<const V*>(static_cast<const T*>(a.x_edge) + (size_t)(row / a.T) * a.C + c);
            else xs = O::zero();
            if (!STAGE2) {
                const V mx = *reinterpret_cast<const V*>(static_cast<const T*>(a.maa[0]) + c);
                *reinterpret_cast<V*>(static_cast<T*>(a.out[0]) + off) = O::add(xv, O::mul(O::sub(xs, xv), mx));
            } else {
                V qm[4];   // the four m_c vectors are read once: streaming loads, all in flight before the first use
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint4 q = __ldcs(reinterpret_cast<const uint4*>(static_cast<const T*>(a.m) + k * plane + off));
                    qm[k] = *reinterpret_cast<const V*>(&q);
                }
                const V xx = O::sub(xs, xv);
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const V mk = *reinterpret_cast<const V*>(static_cast<const T*>(a.maa[k + 1]) + c);
                    *reinterpret_cast<V*>(static_cast<T*>(a.out[k]) + off) = O::add(xv, O::mul(xx, O::add(mk, qm[k])));
                }
            }
        }
    }
}

template <typename T, bool STAGE2>
int launch_t(const TmixArgs& a, cudaStream_t st) {
    if (a.rows == 0) return 0;
    const int vpr = a.C / Vec<T>::W;
    int vx = 32;
    while (vx < vpr && vx < 256) vx <<= 1;
    const int vy = 256 / vx;
    long long blocks = (a.rows + vy - 1) / vy;
    if (blocks > 148 * 32) blocks = 148 * 32;   // up to 148 SMs x 8 resident CTAs x 4 waves, grid-stride beyond that
    wkv6_tmix_mix_kernel<T, STAGE2><<<dim3((unsigned)blocks), dim3(vx, vy), 0, st>>>(a);
    count_launch(1);
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// Backward of one stage (training: the fused prologue with autograd on).  With xx = shift(x) - x and, per output c,
// out_c = x + xx * coef_c, coef_c = maa_c (+ m_c in stage 2):
//     gm_c[t]   = g_c[t] * xx[t]                       (stage 2: gradient of the LoRA output m, input of the bmm backward)
//     gmaa_c   += sum over rows of g_c * xx            (fp32, atomically accumulated: the caller zero-fills)
//     gb[t]     = sum_c g_c[t] * coef_c[t]             (gradient of xx)
//     gx[t]     = sum_c g_c[t] - gb[t] + gb[t + 1]     (left-to-right: x[t] also is the shifted input of row t + 1;
//                                                       right-to-left: gb[t - 1]; zero beyond the sequence)
// A thread owns one 16-byte vector column and walks RB consecutive rows of one sequence in the direction in which the
// neighbour's gb becomes available one step later, so x, g_c and m_c are read once and gx is written once; the row just
// beyond the run contributes its gb only.  fp32 arithmetic, one rounding per stored value.
struct TmixBwdArgs {
    const void* x;
    const void* m;        // [4][rows][C] or null (stage 1)
    const void* maa[4];   // stage 1: maa_x in slot 0
    const void* g[4];     // gradients of the stage's outputs (stage 1: g_xxx in slot 0)
    void* gx;             // [B][T][C]
    void* gm;             // [4][rows][C] or null
    float* gmaa;          // [NP][C] fp32, accumulated
    int B, T, C, reverse;
};

// 16-byte vector <-> W floats
template <typename T> struct VecF;
template <> struct VecF<__nv_bfloat16> {
    static constexpr int W = 8;
    static __device__ __forceinline__ void unpack(const uint4& q, float* f) {
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int i = 0; i < 4; i++) { f[2 * i] = __uint_as_float(w[i] << 16); f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
    }
    static __device__ __forceinline__ uint4 pack(const float* f) {
        return make_uint4(pack2(f[0], f[1]), pack2(f[2], f[3]), pack2(f[4], f[5]), pack2(f[6], f[7]));
    }
    static __device__ __forceinline__ uint32_t pack2(float a, float b) {
        __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t*>(&h);
    }
};
template <> struct VecF<float> {
    static constexpr int W = 4;
    static __device__ __forceinline__ void unpack(const uint4& q, float* f) {
        f[0] = __uint_as_float(q.x); f[1] = __uint_as_float(q.y); f[2] = __uint_as_float(q.z); f[3] = __uint_as_float(q.w);
    }
    static __device__ __forceinline__ uint4 pack(const float* f) {
        return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
    }
};

constexpr int kBwdRows = 64;    // rows of one sequence per CTA

template <typename T, int NP>
__global__ void __launch_bounds__(256) wkv6_tmix_mix_bwd_kernel(const TmixBwdArgs a) {
    using VF = VecF<T>;
    constexpr int W = VF::W;
    extern __shared__ float sm_f[];
    const int C = a.C, Tn = a.T, vpr = C / W;
    float* colsum = sm_f;                                   // [NP][C] column sums of g_c * xx
    float* maa_s = sm_f + NP * C;                           // [NP][C] maa_c as fp32
    const int vx = blockDim.x, vy = blockDim.y, RB = kBwdRows / vy;
    const int b = blockIdx.y;
#pragma unroll
    for (int k = 0; k < NP; k++)
        for (int i = threadIdx.y * vx + threadIdx.x; i < C; i += vx * vy) {
            colsum[k * C + i] = 0.f;
            maa_s[k * C + i] = io_traits<T>::to_float(static_cast<const T*>(a.maa[k])[i]);
        }
    __syncthreads();
    const size_t plane = (size_t)a.B * Tn * C, boff = (size_t)b * Tn * C;
    const T* x = static_cast<const T*>(a.x) + boff;
    T* gx = static_cast<T*>(a.gx) + boff;
    const int run_lo = blockIdx.x * kBwdRows + threadIdx.y * RB;          // rows [run_lo, run_hi) of sequence b
    const int run_hi = min(run_lo + RB, Tn);
    const int n = run_hi - run_lo;
    const int step = a.reverse ? -1 : 1;          // walking direction: the consumer of x[t] as a shifted input is row t + step
    const bool has_m = a.m != nullptr;
    for (int vi = threadIdx.x; vi < vpr && n > 0; vi += vx) {
        const int c = vi * W;
        float acc[NP][W];
#pragma unroll
        for (int k = 0; k < NP; k++)
#pragma unroll
            for (int e = 0; e < W; e++) acc[k][e] = 0.f;
        float pend[W];
#pragma unroll
        for (int e = 0; e < W; e++) pend[e] = 0.f;
        // rows of the run in walking order, then the row beyond it (only its gb is needed, to finalise the last row)
        for (int i = 0; i <= n; i++) {
            const int t = a.reverse ? run_hi - 1 - i : run_lo + i;
            const bool inside = t >= 0 && t < Tn, own = i < n;
            const size_t off = (size_t)t * C + c;
            uint4 gq[NP], mq[NP], xq = make_uint4(0, 0, 0, 0), sq = make_uint4(0, 0, 0, 0);
            if (inside) {
#pragma unroll
                for (int k = 0; k < NP; k++) {
                    gq[k] = __ldcs(reinterpret_cast<const uint4*>(static_cast<const T*>(a.g[k]) + boff + off));
                    if (has_m) mq[k] = __ldcs(reinterpret_cast<const uint4*>(static_cast<const T*>(a.m) + k * plane + boff + off));
                }
                if (own) {
                    xq = *reinterpret_cast<const uint4*>(x + off);
                    const int ts = t - step;                               // xx[t] = x[t - step] - x[t], zero beyond the sequence
                    if (ts >= 0 && ts < Tn) sq = *reinterpret_cast<const uint4*>(x + (size_t)ts * C + c);
                }
            }
            float xx[W], gb[W], ga[W];
            {
                float xv[W], xs[W];
                VF::unpack(xq, xv);
                VF::unpack(sq, xs);
#pragma unroll
                for (int e = 0; e < W; e++) { xx[e] = xs[e] - xv[e]; gb[e] = 0.f; ga[e] = 0.f; }
            }
            if (inside) {
#pragma unroll
                for (int k = 0; k < NP; k++) {
                    float gv[W], mv[W], gmv[W];
                    VF::unpack(gq[k], gv);
                    if (has_m) VF::unpack(mq[k], mv);
#pragma unroll
                    for (int e = 0; e < W; e++) {
                        const float cf = has_m ? maa_s[k * C + c + e] + mv[e] : maa_s[k * C + c + e];
                        gb[e] = fmaf(gv[e], cf, gb[e]);
                        ga[e] += gv[e];
                        gmv[e] = gv[e] * xx[e];
                        acc[k][e] += gmv[e];
                    }
                    if (own && a.gm) *reinterpret_cast<uint4*>(static_cast<T*>(a.gm) + k * plane + boff + off) = VF::pack(gmv);
                }
            }
            if (i > 0) {                                                    // finalise the previous row of the walk
#pragma unroll
                for (int e = 0; e < W; e++) pend[e] += gb[e];
                *reinterpret_cast<uint4*>(gx + (size_t)(t - step) * C + c) = VF::pack(pend);
            }
#pragma unroll
            for (int e = 0; e < W; e++) pend[e] = ga[e] - gb[e];
        }
#pragma unroll
        for (int k = 0; k < NP; k++)
#pragma unroll
            for (int e = 0; e < W; e++) atomicAdd(&colsum[k * C + c + e], acc[k][e]);
    }
    __syncthreads();
    for (int i = threadIdx.y * vx + threadIdx.x; i < NP * C; i += vx * vy) atomicAdd(a.gmaa + i, colsum[i]);
}

template <typename T, int NP>
int launch_bwd_t(const TmixBwdArgs& a, cudaStream_t st) {
    if (a.B == 0 || a.T == 0) return 0;
    const int vpr = a.C / VecF<T>::W;
    int vx = 32;
    while (vx < vpr && vx < 256) vx <<= 1;
    const int vy = 256 / vx;                                 // 1 .. 8 runs of 64 .. 8 rows
    const dim3 grid((a.T + kBwdRows - 1) / kBwdRows, a.B);
    wkv6_tmix_mix_bwd_kernel<T, NP><<<grid, dim3(vx, vy), (size_t)2 * NP * a.C * sizeof(float), st>>>(a);
    count_launch(1);
    return (int)cudaGetLastError();
}

}  // namespace

int launch_tmix_mix_bwd(int dtype, int B, int T, int C, const void* x, const void* m, const void* const* maa, const void* const* g,
                        int reverse, void* gx, void* gm, float* gmaa, cudaStream_t st) {
    TmixBwdArgs a;
    a.x = x; a.m = m; a.gx = gx; a.gm = gm; a.gmaa = gmaa;
    const int np = m ? 4 : 1;
    for (int i = 0; i < 4; i++) { a.maa[i] = i < np ? maa[i] : nullptr; a.g[i] = i < np ? g[i] : nullptr; }
    a.B = B; a.T = T; a.C = C; a.reverse = reverse;
    if (dtype == 0) return m ? launch_bwd_t<__nv_bfloat16, 4>(a, st) : launch_bwd_t<__nv_bfloat16, 1>(a, st);
    return m ? launch_bwd_t<float, 4>(a, st) : launch_bwd_t<float, 1>(a, st);
}

// ---------------------------------------------------------------------------------------------------------------
// Epilogue of the bidirectional block (last part): out = (W_f ln_f(y_f) + W_b ln_b(y_b)) / 2 is ONE GEMM
// [ln_f(y_f) || ln_b(y_b)] x [W_f ; W_b]^T / 2 once the two LayerNorms (model.py:323, one per direction) write their
// results side by side.  This kernel is that pair of LayerNorms: one warp per (frame, direction), statistics in fp32 from
// registers (two passes over the registers, not over memory), output rounded to the I/O type like torch's LayerNorm.
namespace {
template <typename T>
__global__ void __launch_bounds__(256) wkv6_ln_pair_cat_kernel(long long rows, int C, const T* __restrict__ y_f, const T* __restrict__ y_b,
                                                               const T* __restrict__ w_f, const T* __restrict__ b_f,
                                                               const T* __restrict__ w_b, const T* __restrict__ b_b, float eps,
                                                               T* __restrict__ out) {
    using VF = VecF<T>;
    constexpr int W = VF::W, MAXV = 2048 / (32 * W) > 0 ? 2048 / (32 * W) : 1;   // vectors per lane for C <= 2048
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const int vpr = C / W;
    for (long long item = (long long)blockIdx.x * wpb + warp; item < 2 * rows; item += (long long)gridDim.x * wpb) {
        const long long row = item >> 1;
        const int dir = (int)(item & 1);
        const T* y = (dir ? y_b : y_f) + row * C;
        const T* wv = dir ? w_b : w_f;
        const T* bv = dir ? b_b : b_f;
        float v[MAXV][W];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < MAXV; i++) {
            const int vi = lane + 32 * i;
            if (vi < vpr) {
                VF::unpack(__ldcs(reinterpret_cast<const uint4*>(y + vi * W)), v[i]);
#pragma unroll
                for (int e = 0; e < W; e++) s += v[i][e];
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const float mean = s / (float)C;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < MAXV; i++)
            if (lane + 32 * i < vpr) {
#pragma unroll
                for (int e = 0; e < W; e++) { const float d = v[i][e] - mean; q = fmaf(d, d, q); }
            }
#pragma unroll
        for (int o = 16; o; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
        const float rstd = rsqrtf(q / (float)C + eps);
        T* o_ = out + row * 2 * C + (size_t)dir * C;
#pragma unroll
        for (int i = 0; i < MAXV; i++) {
            const int vi = lane + 32 * i;
            if (vi < vpr) {
                float wf[W], bf[W], r[W];
                VF::unpack(*reinterpret_cast<const uint4*>(wv + vi * W), wf);
                VF::unpack(*reinterpret_cast<const uint4*>(bv + vi * W), bf);
#pragma unroll
                for (int e = 0; e < W; e++) r[e] = fmaf((v[i][e] - mean) * rstd, wf[e], bf[e]);
                *reinterpret_cast<uint4*>(o_ + vi * W) = VF::pack(r);
            }
        }
    }
}
}  // namespace

int launch_ln_pair_cat(int dtype, long long rows, int C, const void* y_f, const void* y_b, const void* w_f, const void* b_f,
                       const void* w_b, const void* b_b, float eps, void* out, cudaStream_t st) {
    if (rows == 0) return 0;
    long long blocks = (2 * rows + 7) / 8;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (dtype == 0)
        wkv6_ln_pair_cat_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(
            rows, C, static_cast<const __nv_bfloat16*>(y_f), static_cast<const __nv_bfloat16*>(y_b), static_cast<const __nv_bfloat16*>(w_f),
            static_cast<const __nv_bfloat16*>(b_f), static_cast<const __nv_bfloat16*>(w_b), static_cast<const __nv_bfloat16*>(b_b), eps,
            static_cast<__nv_bfloat16*>(out));
    else
        wkv6_ln_pair_cat_kernel<float><<<(unsigned)blocks, 256, 0, st>>>(
            rows, C, static_cast<const float*>(y_f), static_cast<const float*>(y_b), static_cast<const float*>(w_f),
            static_cast<const float*>(b_f), static_cast<const float*>(w_b), static_cast<const float*>(b_b), eps, static_cast<float*>(out));
    count_launch(1);
    return (int)cudaGetLastError();
}

int launch_tmix_mix(int dtype, int B, int T, int C, const void* x, const void* x_edge, const void* m, const void* const* maa, int reverse,
                    void* const* out, cudaStream_t st) {
    TmixArgs a;
    a.x = x;
    a.x_edge = x_edge;
    a.m = m;
    for (int i = 0; i < 5; i++) a.maa[i] = maa[i];
    for (int i = 0; i < 4; i++) a.out[i] = out[i];
    a.rows = (long long)B * T;
    a.T = T;
    a.C = C;
    a.reverse = reverse;
    if (dtype == 0) return m ? launch_t<__nv_bfloat16, true>(a, st) : launch_t<__nv_bfloat16, false>(a, st);
    return m ? launch_t<float, true>(a, st) : launch_t<float, false>(a, st);
}

}  // namespace wkv6
