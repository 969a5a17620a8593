// convmod.cu -- the bandwidth-bound middle of the Conformer ConvolutionModule as ONE kernel:
//
//     z = pointwise_conv1(x)            [B, Tin, 2C]   (a GEMM; stays a library call)
//     g = glu(z)                        a * sigmoid(b),  a = z[..., :C], b = z[..., C:]
//     d = depthwise_conv(g) + bias      d[t][c] = sum_k w[c][k] g[t + k - left][c], zero beyond [0, Tin)
//     y = swish(layer_norm_C(d))        LayerNorm over the channels of a frame, then x * sigmoid(x)
//     out = pointwise_conv2(y)          (a GEMM; stays a library call)
//
// reference: wenet/transformer/convolution.py:89-144 (forward), :40-87 (shapes of the parameters), with
// kernel_size 31, norm 'layer_norm', activation swish, non-causal in every RWKV config (conf/rwkv/*.yaml:14-18).
// The reference transposes to [B, C, T] for the Conv1d calls and back for the LayerNorm; here everything stays in the
// [B, T, C] layout of the encoder (channels contiguous), so g, d and the LayerNorm input never exist in HBM: the kernel
// reads z once (plus a K-1 row halo per time tile, served by L2) and writes y once -- 3*C*s algorithmic bytes per frame
// against ~11*C*s for the eager chain (glu out, conv out, 2 transposes, LayerNorm out, swish out).
//
// One CTA = one batch row x one tile of TT output frames x all C channels (the LayerNorm needs them all), C/2 threads:
//   phase 1: the (TT + K - 1) x C tile of g, computed from coalesced 16-byte loads of z, goes to shared memory (fp32);
//   phase 2: thread p owns channels 2p, 2p+1 with their K taps in registers (packed f32x2 FMAs); frames are processed in
//            groups of 8, every input row is read from shared memory once per group (K + 7 LDS.64 for 8 x K x 2 FMAs);
//            LayerNorm statistics of the 8 frames: transposing shuffle reduction (16 shuffles), one exchange through
//            shared memory across the warps; y leaves as 8-byte stores (a warp writes 256 contiguous bytes of a frame).
// fp32 math throughout; I/O fp32 or bf16.
#include "wkv6_scan.cuh"

namespace wkv6 {

namespace {

struct ConvArgs {
    const void* z;        // [B][Tin][2C]
    const float* dw_w;    // [C][K]  (Conv1d weight [C,1,K], fp32)
    const float* dw_b;    // [C] or null
    const float* ln_w;    // [C]
    const float* ln_b;    // [C]
    void* y;              // [B][Tout][C]
    int B, Tin, Tout, C, left, TT;
    float eps;
};

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
        "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c)));
    return *reinterpret_cast<float2*>(&d);
}
// 1 / (1 + 2^(-x log2 e)): two MUFU operations (ex2.approx, rcp.approx: ~1 ulp each) and two FMA-pipe ones.  __frcp_rn
// would add a Newton step and a slow-path branch per element (it was a third of the kernel's instructions).
__device__ __forceinline__ float sigmoidf_fast(float x) {
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * -1.4426950408889634f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f + e));
    return r;
}

template <typename IO> __device__ __forceinline__ void load4(const IO* p, float (&f)[4]);
template <> __device__ __forceinline__ void load4<float>(const float* p, float (&f)[4]) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p));
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
}
template <> __device__ __forceinline__ void load4<__nv_bfloat16>(const __nv_bfloat16* p, float (&f)[4]) {
    const uint2 q = __ldg(reinterpret_cast<const uint2*>(p));
    f[0] = __uint_as_float(q.x << 16); f[1] = __uint_as_float(q.x & 0xffff0000u);
    f[2] = __uint_as_float(q.y << 16); f[3] = __uint_as_float(q.y & 0xffff0000u);
}
template <typename IO> __device__ __forceinline__ void store2(IO* p, float a, float b);
template <> __device__ __forceinline__ void store2<float>(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }
template <> __device__ __forceinline__ void store2<__nv_bfloat16>(__nv_bfloat16* p, float a, float b) {
    *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b);
}

// Dynamic-shared-memory opt-in of one kernel, once per device.  (opt_in_smem_once of wkv6_scan.cuh keys its once-flags by the
// kernel's TYPE, which all eight instantiations in this file share: the flags live here, one set per call site.)
template <typename Kern>
static cudaError_t opt_in_smem_impl(Kern kernel, std::atomic<unsigned long long>& done) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    const unsigned long long bit = 1ull << (dev & 63);
    if (done.load(std::memory_order_acquire) & bit) return cudaSuccess;
    e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
    if (e != cudaSuccess) return e;
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    done.fetch_or(bit, std::memory_order_release);
    return cudaSuccess;
}
#define opt_in_smem(...) ([&]() { static std::atomic<unsigned long long> done_{0}; return opt_in_smem_impl(__VA_ARGS__, done_); }())

constexpr int kGroup = 8;      // frames per LayerNorm exchange
constexpr int kMaxWarps = 16;  // C <= 1024

template <typename IO, int K>
__global__ void __launch_bounds__(512) convmod_core_kernel(const ConvArgs a) {
    extern __shared__ __align__(16) float smem[];
    const int C = a.C, TT = a.TT, rows = TT + K - 1;
    float* tile = smem;                                   // [rows][C]
    float* red = smem + (size_t)rows * C;                 // [nwarps][16] partial sums, then [16] totals at red_tot
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    float* red_tot = red + kMaxWarps * 16;
    const int b = blockIdx.y, t0 = blockIdx.x * TT;
    const IO* zb = static_cast<const IO*>(a.z) + (size_t)b * a.Tin * 2 * C;

    // ---- this thread's two channels: K taps each.  The [C][K] weight block is staged through the (still unused) tile with
    // coalesced 16-byte loads; read straight from global memory a warp's load of one tap touches 32 different 128-byte
    // lines (stride K floats), 2.3x the sector traffic of the whole z tile.
    const int c0 = 2 * tid;
    float2 w[K];
    {
        const int nvec = (C * K) >> 2;                    // C % 64 == 0: the block is a whole number of float4
        for (int i = tid; i < nvec; i += blockDim.x)
            reinterpret_cast<float4*>(tile)[i] = __ldg(reinterpret_cast<const float4*>(a.dw_w) + i);
        __syncthreads();
#pragma unroll
        for (int k = 0; k < K; k++) w[k] = make_float2(tile[c0 * K + k], tile[(c0 + 1) * K + k]);   // odd stride: conflict-free
        __syncthreads();
    }

    // ---- phase 1: g = a * sigmoid(b) of input frames [t0 - left, t0 - left + rows), zero outside [0, Tin).
    // blockDim = C/2 threads cover two rows of C/4 four-channel vectors per pass: a thread keeps its vector column and walks
    // the rows; four rows per iteration with all eight loads issued before the first use (the loop is latency-bound otherwise:
    // one dependent global load per iteration and 27 iterations per tile).
    {
        const int vpr = C >> 2;
        const int rsel = tid >= vpr ? 1 : 0, c = (tid - rsel * vpr) << 2;
#pragma unroll 1
        for (int r0 = rsel; r0 < rows; r0 += 8) {
            float av[4][4], bv[4][4];
            bool ok[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int r = r0 + 2 * u, tin = t0 - a.left + r;
                ok[u] = r < rows && tin >= 0 && tin < a.Tin;
                if (ok[u]) {
                    load4<IO>(zb + (size_t)tin * 2 * C + c, av[u]);
                    load4<IO>(zb + (size_t)tin * 2 * C + C + c, bv[u]);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int r = r0 + 2 * u;
                if (r < rows) {
                    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (ok[u])
                        g = make_float4(av[u][0] * sigmoidf_fast(bv[u][0]), av[u][1] * sigmoidf_fast(bv[u][1]),
                                        av[u][2] * sigmoidf_fast(bv[u][2]), av[u][3] * sigmoidf_fast(bv[u][3]));
                    *reinterpret_cast<float4*>(tile + (size_t)r * C + c) = g;
                }
            }
        }
    }

    // ---- bias, LayerNorm affine
    const float2 bias = a.dw_b ? make_float2(__ldg(a.dw_b + c0), __ldg(a.dw_b + c0 + 1)) : make_float2(0.f, 0.f);
    const float2 lw = make_float2(__ldg(a.ln_w + c0), __ldg(a.ln_w + c0 + 1));
    const float2 lb = make_float2(__ldg(a.ln_b + c0), __ldg(a.ln_b + c0 + 1));
    const float invC = 1.f / (float)C;
    IO* yb = static_cast<IO*>(a.y) + (size_t)b * a.Tout * C;
    __syncthreads();

    // ---- phase 2: groups of 8 output frames
    for (int g0 = 0; g0 < TT && t0 + g0 < a.Tout; g0 += kGroup) {
        float2 acc[kGroup];
#pragma unroll
        for (int o = 0; o < kGroup; o++) acc[o] = bias;
#pragma unroll
        for (int j = 0; j < kGroup + K - 1; j++) {
            const float2 x = *reinterpret_cast<const float2*>(tile + (size_t)(g0 + j) * C + c0);
#pragma unroll
            for (int o = 0; o < kGroup; o++) {
                const int k = j - o;
                if (k >= 0 && k < K) acc[o] = ffma2(x, w[k], acc[o]);
            }
        }
        // LayerNorm statistics of the 8 frames over all channels: v[o] = sum, v[8 + o] = sum of squares.
        // Transposing butterfly: after the xor-16/8/4/2 steps lane l holds the warp total of value (l >> 1) & 15 ... see idx
        float v[16];
#pragma unroll
        for (int o = 0; o < kGroup; o++) {
            v[o] = acc[o].x + acc[o].y;
            v[8 + o] = fmaf(acc[o].x, acc[o].x, acc[o].y * acc[o].y);
        }
#pragma unroll
        for (int i = 0; i < 8; i++) {      // xor 16: lanes < 16 keep v[0..7], lanes >= 16 keep v[8..15]
            const bool hi = (lane & 16) != 0;
            const float keep = hi ? v[8 + i] : v[i], send = hi ? v[i] : v[8 + i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
        }
#pragma unroll
        for (int i = 0; i < 4; i++) {      // xor 8
            const bool hi = (lane & 8) != 0;
            const float keep = hi ? v[4 + i] : v[i], send = hi ? v[i] : v[4 + i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
        }
#pragma unroll
        for (int i = 0; i < 2; i++) {      // xor 4
            const bool hi = (lane & 4) != 0;
            const float keep = hi ? v[2 + i] : v[i], send = hi ? v[i] : v[2 + i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
        }
        {                                  // xor 2, then xor 1 (plain sum: two lanes end with the same value)
            const bool hi = (lane & 2) != 0;
            const float keep = hi ? v[1] : v[0], send = hi ? v[0] : v[1];
            v[0] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
            v[0] += __shfl_xor_sync(0xffffffffu, v[0], 1);
        }
        // lane bits (16, 8, 4, 2) select the value: index = 8*b16 + 4*b8 + 2*b4 + b2
        const int idx = ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
        if ((lane & 1) == 0) red[warp * 16 + idx] = v[0];
        __syncthreads();
        if (tid < 16) {
            float s = 0.f;
            for (int wq = 0; wq < nwarps; wq++) s += red[wq * 16 + tid];
            red_tot[tid] = s;
        }
        __syncthreads();
#pragma unroll
        for (int o = 0; o < kGroup; o++) {
            const int t = t0 + g0 + o;
            const float mean = red_tot[o] * invC;
            const float var = fmaxf(red_tot[8 + o] * invC - mean * mean, 0.f);
            const float rstd = rsqrtf(var + a.eps);
            const float y0 = fmaf((acc[o].x - mean) * rstd, lw.x, lb.x);
            const float y1 = fmaf((acc[o].y - mean) * rstd, lw.y, lb.y);
            if (t < a.Tout) store2<IO>(yb + (size_t)t * C + c0, y0 * sigmoidf_fast(y0), y1 * sigmoidf_fast(y1));
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Register-ring variant (C <= 512): the sliding window of the depthwise convolution lives
// in REGISTERS.  A CTA owns one batch row and a span of output frames, thread p owns channels 2p, 2p+1 with their K taps
// (float2 each) and a ring of K partial outputs (float2 each): when g[tin] = glu(z[tin]) arrives it is added to the K outputs
// it touches (K packed FMAs on K independent accumulators), after which the oldest of them is complete.  The ring index of
// an output is (frame - tap) mod K, a compile-time constant in a loop unrolled over K frames.  So glu(z) is computed once per
// frame and span (K - 1 warm-up frames per span instead of a (TT + K - 1)-row tile per TT frames) and never stored; z streams
// through a small shared-memory ring with cp.async (a thread only ever reads what it copied itself: no barrier on the load
// path, kRingPF frames in flight per thread).  Completed frames are parked in shared memory, eight at a time (double-buffered);
// then every warp takes one of the eight frames -- all C channels, so the LayerNorm statistics are a plain warp reduction --
// normalises, applies swish and writes the frame with 16-byte stores: one CTA barrier per eight frames.
constexpr int kEmit = 8;          // frames per LayerNorm batch = frames per cp.async group
constexpr int kZRows = 32;        // raw z rows in the shared-memory ring: the current 8 + 16 in flight + 8 of slack

__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(void* dst, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
template <typename IO> __device__ __forceinline__ void cp_async_pair(IO* dst, const IO* src);   // two consecutive channels
template <> __device__ __forceinline__ void cp_async_pair<float>(float* dst, const float* src) { cp_async8(dst, src); }
template <> __device__ __forceinline__ void cp_async_pair<__nv_bfloat16>(__nv_bfloat16* dst, const __nv_bfloat16* src) { cp_async4(dst, src); }
template <typename IO> __device__ __forceinline__ float2 lds_pair(const IO* p);
template <> __device__ __forceinline__ float2 lds_pair<float>(const float* p) { return *reinterpret_cast<const float2*>(p); }
template <> __device__ __forceinline__ float2 lds_pair<__nv_bfloat16>(const __nv_bfloat16* p) {
    const uint32_t q = *reinterpret_cast<const uint32_t*>(p);
    return make_float2(__uint_as_float(q << 16), __uint_as_float(q & 0xffff0000u));
}
template <typename IO> __device__ __forceinline__ void store4(IO* p, const float* f);
template <> __device__ __forceinline__ void store4<float>(float* p, const float* f) { *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]); }
template <> __device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, const float* f) {
    __nv_bfloat162 a = __floats2bfloat162_rn(f[0], f[1]), b = __floats2bfloat162_rn(f[2], f[3]);
    *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
}

// The ring of partial outputs has R = 32 (K = 31) or 16 (K = 15) slots and the frame loop is unrolled over R frames = R / 8
// sub-blocks of 8: slot (frame - tap) mod R, the row of the z ring, the parking position of a finished frame and the point
// where a batch of eight is complete (always after the 6th frame of a sub-block, as K - 1 = 6 mod 8 for both kernel sizes)
// are all compile-time.  One cp.async group per sub-block, two sub-blocks ahead; inside a sub-block there is no volatile
// instruction, so the compiler overlaps glu(z) of a frame with the FMAs of the one before.
// CT: compile-time channel count (512 and 256 are the shipped encoder widths: every shared / global offset becomes an
// immediate), 0: run-time.
template <typename IO, int K, int CT>
__global__ void __launch_bounds__(256, 1) convmod_ring_kernel(const ConvArgs a) {
    constexpr int R = K > 15 ? 32 : 16;
    static_assert((K - 1) % 8 == 6 && K <= R, "emit alignment");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int C = CT ? CT : a.C, span = a.TT;
    IO* zring = reinterpret_cast<IO*>(smem_raw);                                        // [kZRows][2C] raw z rows
    float* stage = reinterpret_cast<float*>(smem_raw + (size_t)kZRows * 2 * C * sizeof(IO));      // [2][kEmit][C] finished conv frames
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const int b = blockIdx.y, ts = blockIdx.x * span, te = min(ts + span, a.Tout);
    if (ts >= te) return;
    const int c0 = 2 * tid;
    const IO* zb = static_cast<const IO*>(a.z) + (size_t)b * a.Tin * 2 * C;
    IO* yb = static_cast<IO*>(a.y) + (size_t)b * a.Tout * C;

    // ---- taps of this thread's two channels, staged through shared memory with coalesced loads (see convmod_core_kernel)
    float2 w[K];
    {
        float* wst = reinterpret_cast<float*>(smem_raw);
        const int nvec = (C * K) >> 2;
        for (int i = tid; i < nvec; i += blockDim.x) reinterpret_cast<float4*>(wst)[i] = __ldg(reinterpret_cast<const float4*>(a.dw_w) + i);
        __syncthreads();
#pragma unroll
        for (int k = 0; k < K; k++) w[k] = make_float2(wst[c0 * K + k], wst[(c0 + 1) * K + k]);
        __syncthreads();
    }
    const float2 bias = a.dw_b ? make_float2(__ldg(a.dw_b + c0), __ldg(a.dw_b + c0 + 1)) : make_float2(0.f, 0.f);
    const float invC = 1.f / (float)C;

    // input frames tin0 .. tin0 + n_in - 1 feed the outputs ts .. te - 1 (output t reads inputs t - left .. t - left + K - 1);
    // input ordinal i, output ordinal e: input i carries tap k of output e = i - k; output e is complete after input e + K - 1
    const int tin0 = ts - a.left, n_in = te - ts + K - 1;
    const int i_lo = max(0, -tin0), i_hi = min(n_in, a.Tin - tin0);   // input ordinals inside the sequence: [i_lo, i_hi)
    const int e_hi = te - ts;                                          // output ordinals of this span: [0, e_hi)
    IO* const zmine = zring + c0;
    const IO* const zsrc = zb + (ptrdiff_t)tin0 * 2 * C + c0;          // this thread's pair in input ordinal 0 (may lie before z)
    auto issue8 = [&](int i0, int row0) {      // cp.async of input frames i0 .. i0 + 7 into ring rows row0 ..; one group, always committed
#pragma unroll
        for (int j = 0; j < kEmit; j++) {
            const int i = i0 + j;
            if (i >= i_lo && i < i_hi) {
                IO* dst = zmine + (size_t)(row0 + j) * 2 * C;
                const IO* src = zsrc + (ptrdiff_t)i * 2 * C;
                cp_async_pair<IO>(dst, src);
                cp_async_pair<IO>(dst + C, src + C);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    issue8(0, 0);
    issue8(8, 8);

    float2 acc[R];
#pragma unroll
    for (int k = 0; k < R; k++) acc[k] = bias;

    // LayerNorm + swish of the parked frames of batch `batch` (outputs ts + 8 * batch ...): one warp per frame
    auto finish_batch = [&](int batch) {
        const float* st = stage + (size_t)(batch & 1) * kEmit * C;
        for (int fr = warp; fr < kEmit; fr += nwarps) {
            const int t = ts + batch * kEmit + fr;
            if (t >= te) break;
            const float* row = st + (size_t)fr * C;
            float v[4][4];                                     // C <= 512: up to 4 float4 per lane
            float sp[4] = {0.f, 0.f, 0.f, 0.f};                // independent partial sums: no 16-deep dependent chains
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int c = (lane + 32 * i) * 4;
                if (c < C) {
                    const float4 q = *reinterpret_cast<const float4*>(row + c);
                    v[i][0] = q.x; v[i][1] = q.y; v[i][2] = q.z; v[i][3] = q.w;
                    sp[i] = (q.x + q.y) + (q.z + q.w);
                }
            }
            float s = (sp[0] + sp[1]) + (sp[2] + sp[3]);
#pragma unroll
            for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            const float mean = s * invC;
            float qp[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 4; i++)
                if ((lane + 32 * i) * 4 < C) {
#pragma unroll
                    for (int e = 0; e < 4; e++) { const float d = v[i][e] - mean; qp[e] = fmaf(d, d, qp[e]); }
                }
            float q2 = (qp[0] + qp[1]) + (qp[2] + qp[3]);
#pragma unroll
            for (int o = 16; o; o >>= 1) q2 += __shfl_xor_sync(0xffffffffu, q2, o);
            const float rstd = rsqrtf(q2 * invC + a.eps);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int c = (lane + 32 * i) * 4;
                if (c < C) {
                    const float4 lw = __ldg(reinterpret_cast<const float4*>(a.ln_w + c)), lb = __ldg(reinterpret_cast<const float4*>(a.ln_b + c));
                    const float lwv[4] = {lw.x, lw.y, lw.z, lw.w}, lbv[4] = {lb.x, lb.y, lb.z, lb.w};
                    float r[4];
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        const float y = fmaf((v[i][e] - mean) * rstd, lwv[e], lbv[e]);
                        r[e] = y * sigmoidf_fast(y);
                    }
                    store4<IO>(yb + (size_t)t * C + c, r);
                }
            }
        }
    };

    for (int base = 0; base < n_in; base += R) {
#pragma unroll
        for (int sb = 0; sb < R / kEmit; sb++) {
            const int i0 = base + kEmit * sb;
            if (i0 >= n_in) break;                                   // CTA-uniform
            // z ring row of input i = i mod kZRows; base is a multiple of R, and R divides kZRows: rows are compile-time
            // up to the parity of base / R when R = 16
            const int rowbase = (R == 32) ? 0 : (base & 16);
            issue8(i0 + 16, (rowbase + kEmit * sb + 16) & (kZRows - 1));
            asm volatile("cp.async.wait_group 2;" ::: "memory");     // this thread's copies of frames i0 .. i0 + 7 have landed
#pragma unroll
            for (int j = 0; j < kEmit; j++) {
                constexpr int dummy_ = 0; (void)dummy_;
                const int f = kEmit * sb + j;                        // frame within the unrolled period (compile-time)
                const int i = base + f;
                float2 g = make_float2(0.f, 0.f);
                if (i >= i_lo && i < i_hi) {
                    const IO* zr = zmine + (size_t)((rowbase + f) & (kZRows - 1)) * 2 * C;
                    const float2 av = lds_pair<IO>(zr), bv = lds_pair<IO>(zr + C);
                    g = make_float2(av.x * sigmoidf_fast(bv.x), av.y * sigmoidf_fast(bv.y));
                }
#pragma unroll
                for (int k = 0; k < K; k++) acc[(f - k + R) % R] = ffma2(g, w[k], acc[(f - k + R) % R]);
                // the output that just received its last tap: ordinal e = i - (K - 1), slot (f - (K - 1)) mod R
                const int e = i - (K - 1);
                constexpr int kPark = 0; (void)kPark;
                const int sd = (f - (K - 1) + 2 * R) % R;
                // parking position: (e >> 3) & 1 and e & 7 with e = base + f - (K - 1), base a multiple of 16
                const int e_lo = (f - (K - 1) + 64) & 15;            // e mod 16 (compile-time)
                if (e >= 0 && e < e_hi) *reinterpret_cast<float2*>(stage + (size_t)e_lo * C + c0) = acc[sd];
                acc[sd] = bias;
                if (j == 5) {                                        // (K - 1) mod 8 == 6: output ordinal e = 7 mod 8 right here
                    if (e >= 0 && e < e_hi) {                        // CTA-uniform
                        __syncthreads();
                        finish_batch(e >> 3);
                    }
                }
            }
        }
    }
    if ((te - ts) & 7) {                                             // the last span of a row may end inside a batch
        __syncthreads();
        finish_batch((te - ts) >> 3);
    }
}

template <typename IO, int K>
int launch_k(const ConvArgs& a, cudaStream_t st) {
    const size_t smem = ((size_t)(a.TT + K - 1) * a.C + kMaxWarps * 16 + 16) * sizeof(float);
    cudaError_t e = opt_in_smem(convmod_core_kernel<IO, K>);   // + max-shared carve-out: two tiles per SM
    if (e != cudaSuccess) return (int)e;
    const dim3 grid((a.Tout + a.TT - 1) / a.TT, a.B);
    convmod_core_kernel<IO, K><<<grid, a.C / 2, smem, st>>>(a);
    count_launch(1);
    return (int)cudaGetLastError();
}

}  // namespace

// Tile length (multiple of 8 frames, <= 64): minimises a simple cost model -- waves x resident CTAs per SM x (2 TT + K - 1)
// rows of work per CTA (phase 1 touches TT + K - 1 rows, phase 2 TT frames), with a bonus when two or more CTAs share
// an SM (one's loads overlap the other's FMAs).  cfg2 (B=16, T=750, C=512, K=31): TT = 24, 54-row tile = 108 KB, 2 CTAs/SM.
int convmod_pick_tile(int B, int Tout, int C, int K) {
    int best = 0;
    double best_cost = 0.0;
    for (int tt = 8; tt <= 64; tt += 8) {
        const size_t smem = ((size_t)(tt + K - 1) * C + kMaxWarps * 16 + 16) * sizeof(float) + 1024;   // + per-CTA reserve
        if (smem > 226 * 1024) break;
        int occ = (int)(227 * 1024 / smem);
        if (occ > 2048 / (C / 2)) occ = 2048 / (C / 2);
        if (occ > 4) occ = 4;
        const long long ncta = (long long)B * ((Tout + tt - 1) / tt);
        const long long waves = (ncta + 148LL * occ - 1) / (148LL * occ);
        const double cost = (double)waves * occ * (2 * tt + K - 1) * (occ >= 2 ? 0.6 : 1.0);
        if (best == 0 || cost < best_cost) { best = tt; best_cost = cost; }
    }
    return best;
}

// Span of output frames per CTA of the register-ring kernel: one wave of CTAs over the 148 SMs when the problem is small
// (every span pays K - 1 warm-up frames), ~128-frame spans when it is large; multiples of 8 (the emit batch).
static int convmod_pick_span(int B, int Tout) {
    int nspan = (Tout + 127) / 128;
    if ((long long)B * nspan < 148) nspan = 148 / B > 0 ? 148 / B : 1;
    int span = ((Tout + nspan - 1) / nspan + 7) / 8 * 8;
    return span < 16 ? 16 : span;
}

template <typename IO, int K, int CT>
static int launch_ring_ct(ConvArgs a, cudaStream_t st) {
    a.TT = convmod_pick_span(a.B, a.Tout);
    size_t smem = (size_t)kZRows * 2 * a.C * sizeof(IO) + (size_t)2 * kEmit * a.C * sizeof(float);
    const size_t wbytes = (size_t)a.C * K * sizeof(float);           // the taps are staged through the same memory first
    if (smem < wbytes) smem = wbytes;
    cudaError_t e = opt_in_smem(convmod_ring_kernel<IO, K, CT>);
    if (e != cudaSuccess) return (int)e;
    const dim3 grid((a.Tout + a.TT - 1) / a.TT, a.B);
    convmod_ring_kernel<IO, K, CT><<<grid, a.C / 2, smem, st>>>(a);
    count_launch(1);
    return (int)cudaGetLastError();
}

template <typename IO, int K>
static int launch_ring(const ConvArgs& a, cudaStream_t st) {
    if (a.C == 512) return launch_ring_ct<IO, K, 512>(a, st);
    if (a.C == 256) return launch_ring_ct<IO, K, 256>(a, st);
    return launch_ring_ct<IO, K, 0>(a, st);
}

int launch_convmod_core(int dtype, int B, int Tin, int Tout, int C, int K, int left, const void* z, const float* dw_w,
                        const float* dw_b, const float* ln_w, const float* ln_b, float eps, void* y, cudaStream_t st) {
    ConvArgs a;
    a.z = z; a.dw_w = dw_w; a.dw_b = dw_b; a.ln_w = ln_w; a.ln_b = ln_b; a.y = y;
    a.B = B; a.Tin = Tin; a.Tout = Tout; a.C = C; a.left = left; a.eps = eps;
    a.TT = 0;
    // the register ring (4 R + ~40 registers per thread, a 32-row ring of raw z in shared memory) covers C <= 512; wider
    // layers take the shared-memory tile kernel
    const bool ring = C <= 512;
    if (ring) {
        if (dtype == 0) return K == 31 ? launch_ring<__nv_bfloat16, 31>(a, st) : launch_ring<__nv_bfloat16, 15>(a, st);
        return K == 31 ? launch_ring<float, 31>(a, st) : launch_ring<float, 15>(a, st);
    }
    a.TT = convmod_pick_tile(B, Tout, C, K);
    if (a.TT == 0) return (int)cudaErrorInvalidValue;
    if (dtype == 0) return K == 31 ? launch_k<__nv_bfloat16, 31>(a, st) : launch_k<__nv_bfloat16, 15>(a, st);
    return K == 31 ? launch_k<float, 31>(a, st) : launch_k<float, 15>(a, st);
}

}  // namespace wkv6
