/*
 * ref_exports.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * extern "C" doors onto the UNMODIFIED reference launchers, so that tests/ and the golden-vector
 * generator can run the reference's own CUDA kernels on the B200 box through ctypes.  The kernels
 * themselves are compiled from where they lie (/root/reference/wenet/rwkv_v6/cuda/wkv6_cuda.cu,
 * launchers at :265-314) by oracle/build_ref.py; nothing of the reference is copied into this repo.
 * All pointers are device pointers; launches go to the legacy default stream like the reference's.
 */
#include <cuda_runtime.h>
#include <stdint.h>

namespace c10 { struct BFloat16; }
namespace at { using c10::BFloat16; }
typedef at::BFloat16 bf16;

void cuda_forward(int B, int T, int C, int H, bf16* r, bf16* k, bf16* v, bf16* w, bf16* u, bf16* y);
void cuda_backward(int B, int T, int C, int H, bf16* r, bf16* k, bf16* v, bf16* w, bf16* u, bf16* gy, bf16* gr,
                   bf16* gk, bf16* gv, bf16* gw, bf16* gu);
void cuda_forward_fp32(int B, int T, int C, int H, float* r, float* k, float* v, float* w, float* u, float* y);
void cuda_backward_fp32(int B, int T, int C, int H, float* r, float* k, float* v, float* w, float* u, float* gy,
                        float* gr, float* gk, float* gv, float* gw, float* gu);

extern "C" {

int wkv6_ref_ctx_len(void) { return WKV6_REF_T; }

int wkv6_ref_forward(int dtype, int B, int T, int C, int H, void* r, void* k, void* v, void* w, void* u, void* y) {
    if (dtype == 0)
        cuda_forward(B, T, C, H, (bf16*)r, (bf16*)k, (bf16*)v, (bf16*)w, (bf16*)u, (bf16*)y);
    else
        cuda_forward_fp32(B, T, C, H, (float*)r, (float*)k, (float*)v, (float*)w, (float*)u, (float*)y);
    return (int)cudaGetLastError();
}

int wkv6_ref_backward(int dtype, int B, int T, int C, int H, void* r, void* k, void* v, void* w, void* u, void* gy,
                      void* gr, void* gk, void* gv, void* gw, void* gu) {
    if (T > WKV6_REF_T) return -1;   /* kernel_backward_201 indexes a [_T_-2] array (wkv6_cuda.cu:207) */
    if (dtype == 0)
        cuda_backward(B, T, C, H, (bf16*)r, (bf16*)k, (bf16*)v, (bf16*)w, (bf16*)u, (bf16*)gy, (bf16*)gr, (bf16*)gk,
                      (bf16*)gv, (bf16*)gw, (bf16*)gu);
    else
        cuda_backward_fp32(B, T, C, H, (float*)r, (float*)k, (float*)v, (float*)w, (float*)u, (float*)gy, (float*)gr,
                           (float*)gk, (float*)gv, (float*)gw, (float*)gu);
    return (int)cudaGetLastError();
}

}  /* extern "C" */
