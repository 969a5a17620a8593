// umma_timing.cu -- standalone hardware probe (test infrastructure): how long do tcgen05.mma kind::f16 instructions of
// the shapes the chunk kernels use take on a B200?  One CTA, one issuing thread; for each shape it issues n MMAs back
// to back (k-steps of one [64|128][64] x [N][64] product), commits to an mbarrier and waits:
//     issue  = cycles the issuing thread spent on the n tcgen05.mma instructions
//     done   = cycles from the first issue to the mbarrier completion
// The slope of `done` over n is the tensor-pipe cost per instruction, the intercept the pipeline latency.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o umma_timing umma_timing.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__host__ __device__ constexpr uint32_t make_idesc(int a_mn, int b_mn, int M, int Nn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(Nn >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(d_tmem),
                 "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile("{\n.reg .pred P1;\nLAB_WAIT:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra DONE;\nbra LAB_WAIT;\nDONE:\n}\n" ::"r"(
                     smem_u32(bar)),
                 "r"(parity)
                 : "memory");
}

struct Smem {
    alignas(1024) uint8_t A[32 * 1024];
    alignas(1024) uint8_t B[32 * 1024];
    alignas(8) uint64_t bar;
    uint32_t tmem_base;
};

struct Shape { int M, N, a_mn, b_mn, n, ndst; };   // ndst: number of distinct accumulators the n MMAs rotate over

__global__ void __launch_bounds__(128) timing_kernel(const Shape* shapes, int nshapes, long long* out) {
    extern __shared__ __align__(1024) uint8_t raw[];
    Smem& sm = *reinterpret_cast<Smem*>(raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u));
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 64 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(sm.A)[i] = make_uint4(0, 0, 0, 0);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&sm.bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&sm.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tbase = sm.tmem_base;
    if (__shfl_sync(0xffffffffu, warp, 0) == 0) {   // whole warp 0, uniform operands, one elected lane issues (as the kernels do)
        uint32_t parity = 0;
        const uint32_t a_u32 = smem_u32(sm.A), b_u32 = smem_u32(sm.B);
        const uint32_t tb = __shfl_sync(0xffffffffu, tbase, 0);
        for (int s = 0; s < nshapes; s++) {
            const Shape sh = shapes[s];
            const uint32_t idesc = make_idesc(sh.a_mn, sh.b_mn, sh.M, sh.N);
            uint64_t da[4], db[4];
            for (int ks = 0; ks < 4; ks++) {
                da[ks] = sh.a_mn ? make_desc(a_u32 + ks * 2048, 8192, 1024) : make_desc(a_u32 + ks * 32, 16, 1024);
                db[ks] = sh.b_mn ? make_desc(b_u32 + ks * 2048, 8192, 1024) : make_desc(b_u32 + ks * 32, 16, 1024);
            }
            for (int rep = 0; rep < 3; rep++) {   // last repetition is reported (warm)
                uint32_t el;
                asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xffffffff;\nselp.u32 %0, 1, 0, P;\n}\n" : "=r"(el));
                if (el) {
                    const long long t0 = clock64();
                    if (sh.n == 1) mma_f16(tb, da[0], db[0], idesc, 0);
                    for (int g = 0; g < sh.n / 4; g++) {
                        const uint32_t dst = tb + 64 * (g % sh.ndst);
                        mma_f16(dst, da[0], db[0], idesc, 0);
                        mma_f16(dst, da[1], db[1], idesc, 1);
                        mma_f16(dst, da[2], db[2], idesc, 1);
                        mma_f16(dst, da[3], db[3], idesc, 1);
                    }
                    const long long t1 = clock64();
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&sm.bar)) : "memory");
                    mbar_wait(&sm.bar, parity);
                    const long long t2 = clock64();
                    out[2 * s] = t1 - t0;
                    out[2 * s + 1] = t2 - t0;
                }
                parity ^= 1;
                __syncwarp();
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}

int main() {
    const int counts[] = {1, 4, 8, 16, 32};
    const int base[][5] = {   // M, N, a_mn, b_mn, ndst
        {64, 16, 0, 0, 1}, {64, 64, 0, 1, 1}, {64, 64, 1, 1, 1}, {64, 64, 0, 0, 1}, {64, 64, 1, 1, 4}, {128, 64, 1, 1, 1},
        {128, 64, 0, 1, 1}, {128, 16, 0, 0, 1}, {128, 128, 1, 1, 1}, {64, 128, 1, 1, 1}, {64, 256, 0, 1, 1}, {128, 256, 0, 1, 1}};
    const int nb = sizeof(base) / sizeof(base[0]);
    Shape h[64];
    int ns = 0;
    for (int b = 0; b < nb; b++)
        for (int c = 0; c < 5; c++) h[ns++] = Shape{base[b][0], base[b][1], base[b][2], base[b][3], counts[c], base[b][4]};
    Shape* d;
    long long* dout;
    cudaMalloc(&d, sizeof(h));
    cudaMalloc(&dout, 2 * 64 * sizeof(long long));
    cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice);
    const size_t smem = sizeof(Smem) + 1024;
    cudaFuncSetAttribute(timing_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    timing_kernel<<<1, 128, smem>>>(d, ns, dout);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    long long o[128];
    cudaMemcpy(o, dout, sizeof(o), cudaMemcpyDeviceToHost);
    printf("   M    N  A  B  dst |  n: issue/done cycles\n");
    for (int b = 0; b < nb; b++) {
        printf("%4d %4d %s %s %3d  |", base[b][0], base[b][1], base[b][2] ? "MN" : "K ", base[b][3] ? "MN" : "K ", base[b][4]);
        for (int c = 0; c < 5; c++) printf("  %2d: %5lld/%5lld", counts[c], o[2 * (b * 5 + c)], o[2 * (b * 5 + c) + 1]);
        printf("\n");
    }
    return 0;
}
