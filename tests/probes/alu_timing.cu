// alu_timing.cu -- standalone hardware probe (test infrastructure): issue rate per SM sub-partition of the instructions
// the tile-building phases are made of (MUFU.EX2, FFMA, F2FP.BF16 pack, HMUL2.BF16, SHFL, FMNMX), on one SM of a B200.
// One CTA of `warps` warps; every thread runs 8 independent chains of the instruction for ITER iterations.
// Prints cycles per warp-instruction per sub-partition (SMSP): 1.0 = full rate (32 lanes/clk/SMSP), 8.0 = 4 lanes/clk.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o alu_timing alu_timing.cu
#include <cstdint>
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

constexpr int ITER = 512;

template <int OP>
__global__ void k(float* out, long long* cyc, float seed) {
    float x[8];
    uint32_t u[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { x[i] = seed + 0.001f * (threadIdx.x + i); u[i] = __float_as_uint(x[i]); }
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 1) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(x[i]));
            if (OP == 2) asm volatile("cvt.rn.bf16x2.f32 %0, %1, %1;" : "=r"(u[i]) : "f"(x[i]), "r"(u[i]));
            if (OP == 3) asm volatile("mul.rn.bf16x2 %0, %0, %0;" : "+r"(u[i]));
            if (OP == 4) asm volatile("shfl.sync.bfly.b32 %0, %0, 1, 0x1f, 0xffffffff;" : "+r"(u[i]));
            if (OP == 5) asm volatile("min.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(seed));
            if (OP == 6) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 7) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(u[i]));
        }
    }
    const long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += x[i] + __uint_as_float(u[i]);
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

template <int OP>
void run(const char* name, float* out, long long* cyc) {
    printf("%-22s", name);
    for (int warps : {4, 8, 16, 32}) {
        k<OP><<<1, warps * 32>>>(out, cyc, 0.5f);
        cudaDeviceSynchronize();
        k<OP><<<1, warps * 32>>>(out, cyc, 0.5f);
        cudaDeviceSynchronize();
        long long c;
        cudaMemcpy(&c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
        const double per = (double)c / ((double)ITER * 8 * (warps / 4));   // warp-instructions per SMSP
        printf("  %2dw: %6.2f", warps, per);
    }
    printf("\n");
}

int main() {
    float* out;
    long long* cyc;
    cudaMalloc(&out, 4096);
    cudaMalloc(&cyc, 8);
    printf("cycles per warp-instruction per SMSP (4 / 8 / 16 / 32 warps in one CTA)\n");
    run<0>("ex2.approx.ftz.f32", out, cyc);
    run<6>("rcp.approx.ftz.f32", out, cyc);
    run<7>("ex2.approx.ftz.bf16x2", out, cyc);
    run<1>("fma.rn.f32", out, cyc);
    run<2>("cvt.rn.bf16x2.f32", out, cyc);
    run<3>("mul.rn.bf16x2", out, cyc);
    run<4>("shfl.bfly", out, cyc);
    run<5>("min.f32", out, cyc);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
