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
// here too (bf16: sub, add, mul, add each round once, as ATen's bf16 elementwise kernels do; fp32: plain IEEE ops, no
// FMA contraction).
#include "wkv6_scan.cuh"
#include "wkv6_umma.cuh"

namespace wkv6 {

namespace {

template <typename T> struct Vec;
template <> struct Vec<__nv_bfloat16> {
    static constexpr int W = 8;   // elements per 16-byte vector
    static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&f)[8]) { unpack(*reinterpret_cast<const uint4*>(p), f); }
    static __device__ __forceinline__ void unpack(const uint4 q, float (&f)[8]) {
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            f[2 * i] = __uint_as_float(w[i] << 16);
            f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&f)[8]) {
        uint4 q;
        q.x = pack_bf16(f[0], f[1]); q.y = pack_bf16(f[2], f[3]); q.z = pack_bf16(f[4], f[5]); q.w = pack_bf16(f[6], f[7]);
        *reinterpret_cast<uint4*>(p) = q;
    }
    static __device__ __forceinline__ float rnd(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
};
template <> struct Vec<float> {
    static constexpr int W = 4;
    static __device__ __forceinline__ void load(const float* p, float (&f)[4]) { unpack(*reinterpret_cast<const uint4*>(p), f); }
    static __device__ __forceinline__ void unpack(const uint4 q, float (&f)[4]) {
        f[0] = __uint_as_float(q.x); f[1] = __uint_as_float(q.y); f[2] = __uint_as_float(q.z); f[3] = __uint_as_float(q.w);
    }
    static __device__ __forceinline__ void store(float* p, const float (&f)[4]) {
        *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
    }
    static __device__ __forceinline__ float rnd(float v) { return v; }
};

struct TmixArgs {
    const void* x;
    const void* x_edge;   // [B][C] frame beyond the edge (x[-1], or x[T] when reverse), or null: zeros
    const void* m;        // [4][B*T*C] (r, k, v, w) or null: stage 1
    const void* maa[5];   // x, r, k, v, w: [C]
    void* out[4];         // stage 1: out[0] = xxx; stage 2: x_r, x_k, x_v, x_w
    long long rows;       // B*T
    int T, C, reverse;
};

template <typename T, bool STAGE2>
__global__ void __launch_bounds__(256) wkv6_tmix_mix_kernel(const TmixArgs a) {
    constexpr int W = Vec<T>::W;
    const int vpr = a.C / W;                                   // vectors per row
    const unsigned nvec = (unsigned)(a.rows * vpr);            // < 2^31, checked by the launcher: 32-bit index arithmetic
    const T* x = static_cast<const T*>(a.x);
    for (unsigned v = blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += gridDim.x * blockDim.x) {
        const unsigned row = v / (unsigned)vpr;
        const int c = (int)(v - row * vpr) * W;
        const int t = (int)(row % (unsigned)a.T);
        const size_t off = (size_t)row * a.C + c;
        float xv[W], xs[W], xx[W];
        Vec<T>::load(x + off, xv);
        const bool edge = a.reverse ? (t == a.T - 1) : (t == 0);
        if (edge) {
            if (a.x_edge) {
                Vec<T>::load(static_cast<const T*>(a.x_edge) + (size_t)(row / (unsigned)a.T) * a.C + c, xs);
            } else {
#pragma unroll
                for (int i = 0; i < W; i++) xs[i] = 0.f;
            }
        } else {
            Vec<T>::load(a.reverse ? x + off + a.C : x + off - a.C, xs);
        }
#pragma unroll
        for (int i = 0; i < W; i++) xx[i] = Vec<T>::rnd(__fsub_rn(xs[i], xv[i]));
        if (!STAGE2) {
            float mx[W], o[W];
            Vec<T>::load(static_cast<const T*>(a.maa[0]) + c, mx);
#pragma unroll
            for (int i = 0; i < W; i++) o[i] = Vec<T>::rnd(__fadd_rn(xv[i], Vec<T>::rnd(__fmul_rn(xx[i], mx[i]))));
            Vec<T>::store(static_cast<T*>(a.out[0]) + off, o);
        } else {
            const size_t plane = (size_t)a.rows * a.C;
            uint4 qm[4];   // the four m_c vectors are read once: streaming loads, all in flight before the first use
#pragma unroll
            for (int k = 0; k < 4; k++) qm[k] = __ldcs(reinterpret_cast<const uint4*>(static_cast<const T*>(a.m) + k * plane + off));
#pragma unroll
            for (int k = 0; k < 4; k++) {
                float mk[W], mm[W], o[W];
                Vec<T>::load(static_cast<const T*>(a.maa[k + 1]) + c, mk);
                Vec<T>::unpack(qm[k], mm);
#pragma unroll
                for (int i = 0; i < W; i++) {
                    const float s = Vec<T>::rnd(__fadd_rn(mk[i], mm[i]));
                    o[i] = Vec<T>::rnd(__fadd_rn(xv[i], Vec<T>::rnd(__fmul_rn(xx[i], s))));
                }
                Vec<T>::store(static_cast<T*>(a.out[k]) + off, o);
            }
        }
    }
}

template <typename T, bool STAGE2>
int launch_t(const TmixArgs& a, cudaStream_t st) {
    const long long nvec = a.rows * (a.C / Vec<T>::W);
    if (nvec == 0) return 0;
    if (nvec >= (1ll << 31)) return (int)cudaErrorInvalidValue;
    // up to 148 SMs x 8 resident CTAs x 4 waves of 256 threads, grid-stride beyond that
    long long blocks = (nvec + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    wkv6_tmix_mix_kernel<T, STAGE2><<<dim3((unsigned)blocks), dim3(256), 0, st>>>(a);
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
