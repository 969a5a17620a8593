// wkv6_fwd_tiles.cuh -- pieces shared by the forward chunk kernels (wkv6_chunk_lean.cu, wkv6_chunk_ws.cu): packed
// bf16x2 arithmetic, stmatrix / 64-bit shared stores, the compact Ktil variant layout and the cross-block tile pass.
#pragma once
#include "wkv6_umma.cuh"

namespace wkv6 {

__device__ __forceinline__ uint32_t bmul2(uint32_t a, uint32_t b) {
    uint32_t d;
    asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t bfma2(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ void stsm_x4(uint32_t saddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("stmatrix.sync.aligned.m8n8.x4.shared.b16 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(a), "r"(b), "r"(c), "r"(d)
                 : "memory");
}
// transposed store: memory row r of matrix k (address from lane 8k + r) receives COLUMN r of the fragment matrix
__device__ __forceinline__ void stsm_x4_trans(uint32_t saddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("stmatrix.sync.aligned.m8n8.x4.trans.shared.b16 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(a), "r"(b), "r"(c), "r"(d)
                 : "memory");
}
__device__ __forceinline__ void sts64(uint32_t saddr, uint32_t a, uint32_t b) {
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(saddr), "r"(a), "r"(b) : "memory");
}

constexpr float kLog2Log2e = 0.5287663729448977f;   // log2(log2(e)):  2^(w log2e + this) = e^w log2e

// Ktil variant bases inside LeanSmem::kt (virtual row 0 of variant I), per direction
__device__ __forceinline__ constexpr int ktbase(bool REV, int I) {
    return REV ? (I == 0 ? 20 * 1024 : I == 1 ? 18 * 1024 : I == 2 ? 14 * 1024 : I == 3 ? 0 : 8 * 1024)
               : (I == 0 ? 0 : I == 1 ? 2 * 1024 : I == 2 ? 6 * 1024 : I == 3 ? 12 * 1024 : 20 * 1024);
}

// phase 2 of one thread: cross-block factors -> Rst and the Ktil variants I+1..4 of this thread's 4 rows x 4 channels.
// state_only (pass 1 of a time-segmented launch: no y): only the variant relative to the chunk end, the operand of dS.
template <int I>
__device__ __forceinline__ void lean_phase2(const bool REV, uint32_t rst_u32, uint32_t kt_u32, const float* bt, int c0, const int (&offs)[4],
                                            const uint32_t (&ro2)[4][2], const uint32_t (&ko2)[4][2], const bool state_only) {
    float tot[4][4];
#pragma unroll
    for (int m = 0; m < 4; m++) {
        const float4 t = *reinterpret_cast<const float4*>(bt + m * N + c0);
        tot[m][0] = t.x; tot[m][1] = t.y; tot[m][2] = t.z; tot[m][3] = t.w;
    }
    if (state_only) {
        float f[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            float s = 0.f;
#pragma unroll
            for (int m = I; m < 4; m++) s += tot[m][e];
            f[e] = ex2(-s);                     // 2^(A_L - a_I)
        }
        const uint32_t f0 = pack_bf16(f[0], f[1]), f1 = pack_bf16(f[2], f[3]);
        const uint32_t base = kt_u32 + ktbase(REV, 4);
#pragma unroll
        for (int r = 0; r < 4; r++) sts64(base + offs[r], bmul2(ko2[r][0], f0), bmul2(ko2[r][1], f1));
        return;
    }
    if (I == 0) {
#pragma unroll
        for (int r = 0; r < 4; r++) sts64(rst_u32 + offs[r], ro2[r][0], ro2[r][1]);
    } else {
        float f[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            float s = 0.f;
#pragma unroll
            for (int m = 0; m < I; m++) s += tot[m][e];
            f[e] = ex2(-s);                     // 2^(a_I)
        }
        const uint32_t f0 = pack_bf16(f[0], f[1]), f1 = pack_bf16(f[2], f[3]);
#pragma unroll
        for (int r = 0; r < 4; r++) sts64(rst_u32 + offs[r], bmul2(ro2[r][0], f0), bmul2(ro2[r][1], f1));
    }
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int m = I; m < 4; m++) {
        float f[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            acc[e] += tot[m][e];
            f[e] = ex2(-acc[e]);                // 2^(a_{m+1} - a_I)
        }
        const uint32_t f0 = pack_bf16(f[0], f[1]), f1 = pack_bf16(f[2], f[3]);
        const uint32_t base = kt_u32 + ktbase(REV, m + 1);
#pragma unroll
        for (int r = 0; r < 4; r++) sts64(base + offs[r], bmul2(ko2[r][0], f0), bmul2(ko2[r][1], f1));
    }
}

}  // namespace wkv6
