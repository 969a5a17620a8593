// umma_probe.cu -- standalone hardware probe (test infrastructure): checks, on a real B200, the smem-descriptor,
// instruction-descriptor and TMEM-load conventions the chunk kernel relies on:
//   T1: D = A * B^T, A [64 m][64 k] and B [64 n][64 k] stored row-major with 128B swizzle  (K-major operands)
//   T2: D = A^T * B, A [64 k][64 m] and B [64 k][64 n] stored row-major with 128B swizzle  (MN-major operands)
// D (M=64, N=64, fp32) is read back with tcgen05.ld 16x256b.x8 and compared on the host.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_probe umma_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;   // SWIZZLE_128B
    return d;
}

// kind::f16 instruction descriptor: bf16 x bf16 -> fp32, M=64, N=64
__host__ __device__ constexpr uint32_t make_idesc(int a_mn_major, int b_mn_major, int M, int Nn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
           ((uint32_t)(Nn >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void mma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred P1;\nLAB_WAIT:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra DONE;\nbra LAB_WAIT;\nDONE:\n}\n" ::"r"(
            smem_u32(bar)),
        "r"(parity)
        : "memory");
}

// element (row, col) of a [64][64] bf16 tile with 128-byte rows and the 128B swizzle
__host__ __device__ inline int sw128_index(int row, int col) { return row * 64 + (((col >> 3) ^ (row & 7)) << 3) + (col & 7); }

struct Smem {
    alignas(1024) __nv_bfloat16 A[64 * 64];
    alignas(1024) __nv_bfloat16 B[64 * 64];
    alignas(8) uint64_t bar;
    uint32_t tmem_base;
};

// mode 0: T1 (K-major), mode 1: T2 (MN-major).  out: [128 threads][32 regs] raw 16x256b.x8 dump
__global__ void __launch_bounds__(128) probe_kernel(const __nv_bfloat16* gA, const __nv_bfloat16* gB, float* out, int mode) {
    extern __shared__ __align__(1024) uint8_t raw[];
    Smem& sm = *reinterpret_cast<Smem*>(raw);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 64 * 64; i += 128) {
        const int r = i >> 6, c = i & 63;
        sm.A[sw128_index(r, c)] = gA[i];
        sm.B[sw128_index(r, c)] = gB[i];
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&sm.bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&sm.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic smem writes -> visible to the MMA (async proxy)
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = sm.tmem_base;
    if (tid == 0) {
        const uint32_t idesc = make_idesc(mode, mode, 64, 64);
        for (int ks = 0; ks < 4; ks++) {
            uint64_t ad, bd;
            if (mode == 0) {   // K-major: 16 k = 32 bytes further along the swizzled row
                ad = make_desc(smem_u32(sm.A) + ks * 32, 16, 1024);
                bd = make_desc(smem_u32(sm.B) + ks * 32, 16, 1024);
            } else {           // MN-major: 16 k = 16 rows = 2048 bytes further; 8-row groups 1024 B apart
                ad = make_desc(smem_u32(sm.A) + ks * 2048, 8192, 1024);
                bd = make_desc(smem_u32(sm.B) + ks * 2048, 8192, 1024);
            }
            mma_f16(tbase, ad, bd, idesc, ks > 0);
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&sm.bar)) : "memory");
    }
    mbar_wait(&sm.bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t v[32];
    const uint32_t taddr = tbase + ((uint32_t)(warp * 32) << 16);
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,"
        "%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int i = 0; i < 32; i++) out[tid * 32 + i] = __uint_as_float(v[i]);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tbase) : "memory");
}

int main() {
    __nv_bfloat16 hA[4096], hB[4096];
    float fA[4096], fB[4096];
    srand(7);
    for (int i = 0; i < 4096; i++) {
        fA[i] = (float)(rand() % 9 - 4);
        fB[i] = (float)(rand() % 7 - 3);
        hA[i] = __float2bfloat16(fA[i]);
        hB[i] = __float2bfloat16(fB[i]);
    }
    __nv_bfloat16 *dA, *dB;
    float* dO;
    cudaMalloc(&dA, 8192); cudaMalloc(&dB, 8192); cudaMalloc(&dO, 128 * 32 * 4);
    cudaMemcpy(dA, hA, 8192, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB, 8192, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(Smem) + 1024);
    int bad_total = 0;
    for (int mode = 0; mode < 2; mode++) {
        cudaMemset(dO, 0, 128 * 32 * 4);
        probe_kernel<<<1, 128, sizeof(Smem) + 1024>>>(dA, dB, dO, mode);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("mode %d: CUDA error %s\n", mode, cudaGetErrorString(e)); return 1; }
        static float ho[128 * 32];
        cudaMemcpy(ho, dO, sizeof(ho), cudaMemcpyDeviceToHost);
        // hypothesis: warp w, lane l, reg = 4*rep + q: row = 16*w + (l/4) + 8*(q/2), col = 8*rep + 2*(l%4) + (q%2)
        int bad = 0;
        for (int t = 0; t < 128; t++)
            for (int g = 0; g < 32; g++) {
                const int w = t / 32, l = t % 32, rep = g / 4, q = g % 4;
                const int row = 16 * w + l / 4 + 8 * (q / 2), col = 8 * rep + 2 * (l % 4) + (q % 2);
                float ref = 0;
                for (int k = 0; k < 64; k++)
                    ref += mode == 0 ? fA[row * 64 + k] * fB[col * 64 + k] : fA[k * 64 + row] * fB[k * 64 + col];
                if (ho[t * 32 + g] != ref) {
                    if (bad < 6) printf("mode %d mismatch t=%d reg=%d (row %d col %d): got %g want %g\n", mode, t, g, row, col, ho[t * 32 + g], ref);
                    bad++;
                }
            }
        printf("mode %d (%s): %d / 4096 mismatches\n", mode, mode == 0 ? "K-major SW128" : "MN-major SW128", bad);
        bad_total += bad;
    }
    printf(bad_total == 0 ? "PROBE OK\n" : "PROBE FAILED\n");
    return bad_total != 0;
}
