// wkv6_umma.cuh -- tcgen05 / TMEM helpers shared by the chunked forward and backward kernels (sm_100a).
// Descriptor conventions (K-major / MN-major tiles of [64][64] bf16 with 128-byte rows and the 128B swizzle, the
// kind::f16 instruction descriptor, the M=64 TMEM lane layout and the 16x256b fragment map) were verified on a
// B200 by tests/probes/umma_probe.cu.
#pragma once
#include "wkv6_scan.cuh"

namespace wkv6 {

constexpr int CL = 64;                              // chunk length
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kClampL2 = 5.0f * 1.4426950408889634f;   // per-step decay floor exp(-5) = 6.7e-3

__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;   // SWIZZLE_128B
    return d;
}
__device__ __forceinline__ uint64_t desc_k(uint32_t saddr) { return umma_desc(saddr, 16, 1024); }     // K-major, +32 B per 16 k
__device__ __forceinline__ uint64_t desc_mn(uint32_t saddr) { return umma_desc(saddr, 8192, 1024); }  // MN-major, +2048 B per 16 k

// kind::f16 instruction descriptor: bf16 x bf16 -> fp32
__host__ __device__ constexpr uint32_t umma_idesc(int a_mn, int b_mn, int M, int Nn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(Nn >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void umma(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// one lane of a converged warp (the same one every time): MMA issue without a divergent-predicate waterfall loop
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xffffffff;\nselp.u32 %0, 1, 0, P;\n}\n" : "=r"(pred));
    return pred != 0;
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// 16 TMEM lanes x 64 columns -> 32 registers per thread: reg 4*rep + 2*hh + e  <->  row (lane/4) + 8*hh,
// column 8*rep + 2*(lane%4) + e
__device__ __forceinline__ void tmem_ld_16x64(uint32_t taddr, float (&v)[32]) {
    uint32_t u[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,"
        "%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]), "=r"(u[8]), "=r"(u[9]),
          "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15]), "=r"(u[16]), "=r"(u[17]), "=r"(u[18]),
          "=r"(u[19]), "=r"(u[20]), "=r"(u[21]), "=r"(u[22]), "=r"(u[23]), "=r"(u[24]), "=r"(u[25]), "=r"(u[26]), "=r"(u[27]),
          "=r"(u[28]), "=r"(u[29]), "=r"(u[30]), "=r"(u[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; i++) v[i] = __uint_as_float(u[i]);
}

// byte offset of the 16-byte chunk `cg` of row `row` in a [rows][64] bf16 tile with the 128B swizzle
__device__ __forceinline__ int sw_chunk(int row, int cg) { return row * 128 + ((cg ^ (row & 7)) << 4); }

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

__device__ __forceinline__ void tmem_ld_16x32(uint32_t taddr, float (&v)[16]) {   // 16 lanes x 32 columns
    uint32_t u[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]), "=r"(u[8]), "=r"(u[9]),
          "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = __uint_as_float(u[i]);
}

__device__ __forceinline__ void ld4bf(const uint8_t* p, float (&f)[4]) {   // 4 consecutive bf16 -> fp32
    const uint2 q = *reinterpret_cast<const uint2*>(p);
    f[0] = __uint_as_float(q.x << 16); f[1] = __uint_as_float(q.x & 0xffff0000u);
    f[2] = __uint_as_float(q.y << 16); f[3] = __uint_as_float(q.y & 0xffff0000u);
}
__device__ __forceinline__ void up4bf(const uint2 q, float (&f)[4]) {
    f[0] = __uint_as_float(q.x << 16); f[1] = __uint_as_float(q.x & 0xffff0000u);
    f[2] = __uint_as_float(q.y << 16); f[3] = __uint_as_float(q.y & 0xffff0000u);
}
__device__ __forceinline__ void st4bf(uint8_t* p, const float (&f)[4]) {
    *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]));
}
// x = hi + lo with hi = bf16(x), lo = bf16(x - hi): two bf16 tiles carry ~16 mantissa bits through the MMA.  Needed
// where two differently factorised sums must cancel (the gw prefix sum), see DESIGN.md.
__device__ __forceinline__ void st4bf_hilo(uint8_t* hi, uint8_t* lo, const float (&f)[4]) {
    const uint32_t h0 = pack_bf16(f[0], f[1]), h1 = pack_bf16(f[2], f[3]);
    *reinterpret_cast<uint2*>(hi) = make_uint2(h0, h1);
    const float l0 = f[0] - __uint_as_float(h0 << 16), l1 = f[1] - __uint_as_float(h0 & 0xffff0000u);
    const float l2 = f[2] - __uint_as_float(h1 << 16), l3 = f[3] - __uint_as_float(h1 & 0xffff0000u);
    *reinterpret_cast<uint2*>(lo) = make_uint2(pack_bf16(l0, l1), pack_bf16(l2, l3));
}


// 16 TMEM lanes x 16 columns -> 8 registers per thread (same register map as tmem_ld_16x32 with rep = 0, 1)
__device__ __forceinline__ void tmem_ld_16x16(uint32_t taddr, float (&v)[8]) {
    uint32_t u[8];
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = __uint_as_float(u[i]);
}

// two-channel forms of the helpers above (one 32-bit word = 2 consecutive bf16)
__device__ __forceinline__ void up2bf(const uint32_t q, float (&f)[2]) {
    f[0] = __uint_as_float(q << 16);
    f[1] = __uint_as_float(q & 0xffff0000u);
}
__device__ __forceinline__ void st2bf(uint8_t* p, const float (&f)[2]) { *reinterpret_cast<uint32_t*>(p) = pack_bf16(f[0], f[1]); }
__device__ __forceinline__ void st2bf_hilo(uint8_t* hi, uint8_t* lo, const float (&f)[2]) {
    const uint32_t h0 = pack_bf16(f[0], f[1]);
    *reinterpret_cast<uint32_t*>(hi) = h0;
    *reinterpret_cast<uint32_t*>(lo) = pack_bf16(f[0] - __uint_as_float(h0 << 16), f[1] - __uint_as_float(h0 & 0xffff0000u));
}

}  // namespace wkv6
