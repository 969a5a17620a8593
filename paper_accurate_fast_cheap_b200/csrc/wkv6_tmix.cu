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

}  // namespace

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
