// wkv6_scan.cu -- exact (recurrent-form) WKV6 kernels for sm_100a: fp32 math on the FMA pipe.
//
// What it computes (reference: wenet/rwkv_v6/cuda/wkv6_cuda.cu:9-63 forward, :66-263 backward;
// initial state / gs: wkv6state_cuda.cu:7-65, 68-293; math restated in SURVEY.md App. A):
//     y_t[i]   = sum_j r_t[j] * (u[j] k_t[j] v_t[i] + S_t[i][j])
//     S_{t+1}  = S_t * diag(exp(-exp(w_t))) + v_t k_t^T
// and all gradients.  How it differs from the reference kernels (this is a redesign, not a port):
//   * one CTA of 128 threads per (direction, batch, head); each thread owns a 4(i) x 8(j) tile of the
//     64x64 state in registers as float2 pairs and updates it with packed FFMA2/FMUL2;
//   * r/k/v/w/gy arrive as [16 x 64] TMA tiles (cp.async.bulk.tensor + mbarrier ring) instead of
//     per-step scalar global loads; exp(-exp(w)) is evaluated once per element;
//   * reductions over i / j are register reduce-scatters over warp shuffles -- one __syncthreads per
//     16 steps (forward) instead of two per step;
//   * the reversed direction walks the tiles backwards and writes outputs in place (no flip copies);
//   * backward is two sweeps (A: forward-time state S -> gr, gu;  B: reverse-time state C -> gk, gv, gw)
//     and gw uses the suffix-sum identity  H_t = H_{t+1} + r_{t+1} grS_{t+1} - k_t gkC_t,
//     gw_t = -exp(w_t) H_t  -- no O(T) per-thread array, hence no T <= ctx_len limit
//     (the reference's kernel_backward_201 is UB for T > _T_, wkv6_cuda.cu:207).
#include "wkv6_scan.cuh"

namespace wkv6 {

// ---------------------------------------------------------------------------------------
// thread tile: lane bits 0-2 -> j group, lane bits 3-4 + warp -> i group
// ---------------------------------------------------------------------------------------
struct Tile {
    int lane, warp, jg, ig, i0, jA, jB;
    __device__ explicit Tile(int tid) {
        lane = tid & 31;
        warp = tid >> 5;
        jg = lane & 7;
        ig = warp * 4 + (lane >> 3);
        i0 = ig * 4;
        jA = jg * 4;        // j set of a thread: {jA..jA+3} U {jB..jB+3}: 128-bit LDS without bank conflicts
        jB = 32 + jg * 4;
    }
};

// x[a], a = 0..3, summed over the 8 lanes sharing lane>>3.  Returns the sum for a = 2*b0 + b1
// (b0 = lane&1, b1 = (lane>>1)&1); the value is replicated in lanes differing in bit 2.
__device__ __forceinline__ float reduce4_over_j(const float (&x)[4], int lane) {
    const bool b0 = lane & 1, b1 = lane & 2;
    float k0 = b0 ? x[2] : x[0], k1 = b0 ? x[3] : x[1];
    const float s0 = b0 ? x[0] : x[2], s1 = b0 ? x[1] : x[3];
    k0 += __shfl_xor_sync(0xffffffffu, s0, 1);
    k1 += __shfl_xor_sync(0xffffffffu, s1, 1);
    float kp = b1 ? k1 : k0;
    const float sd = b1 ? k0 : k1;
    kp += __shfl_xor_sync(0xffffffffu, sd, 2);
    kp += __shfl_xor_sync(0xffffffffu, kp, 4);
    return kp;
}
__device__ __forceinline__ int reduce4_index(int lane) { return 2 * (lane & 1) + ((lane >> 1) & 1); }

// x[e], e = 0..7, summed over the 4 lanes sharing lane&7 (lane bits 3,4).  Returns the sums for
// e = 4*b3 + 2*b4 + {0,1}.
__device__ __forceinline__ float2 reduce8_over_i(const float (&x)[8], int lane) {
    const bool b3 = lane & 8, b4 = lane & 16;
    float k[4], s[4];
#pragma unroll
    for (int e = 0; e < 4; e++) {
        k[e] = b3 ? x[4 + e] : x[e];
        s[e] = b3 ? x[e] : x[4 + e];
    }
#pragma unroll
    for (int e = 0; e < 4; e++) k[e] += __shfl_xor_sync(0xffffffffu, s[e], 8);
    float2 kp, sd;
    kp.x = b4 ? k[2] : k[0];
    kp.y = b4 ? k[3] : k[1];
    sd.x = b4 ? k[0] : k[2];
    sd.y = b4 ? k[1] : k[3];
    kp.x += __shfl_xor_sync(0xffffffffu, sd.x, 16);
    kp.y += __shfl_xor_sync(0xffffffffu, sd.y, 16);
    return kp;
}
// channel j (0..63) of the first of the two values returned by reduce8_over_i
__device__ __forceinline__ int reduce8_index(const Tile& t) {
    return ((t.lane & 8) ? 32 : 0) + t.jA + ((t.lane & 16) ? 2 : 0);
}

__device__ __forceinline__ void split4(const float4 a, const float4 b, float2 (&o)[4]) {
    o[0] = make_float2(a.x, a.y);
    o[1] = make_float2(a.z, a.w);
    o[2] = make_float2(b.x, b.y);
    o[3] = make_float2(b.z, b.w);
}

__device__ __forceinline__ void load_state(float2 (&S)[4][4], const float* s, const Tile& t) {
#pragma unroll
    for (int a = 0; a < 4; a++) {
        if (s) {
            const float4 x = *reinterpret_cast<const float4*>(s + (t.i0 + a) * N + t.jA);
            const float4 z = *reinterpret_cast<const float4*>(s + (t.i0 + a) * N + t.jB);
            split4(x, z, S[a]);
        } else {
#pragma unroll
            for (int q = 0; q < 4; q++) S[a][q] = make_float2(0.f, 0.f);
        }
    }
}
__device__ __forceinline__ void store_state(const float2 (&S)[4][4], float* s, const Tile& t) {
#pragma unroll
    for (int a = 0; a < 4; a++) {
        *reinterpret_cast<float4*>(s + (t.i0 + a) * N + t.jA) = make_float4(S[a][0].x, S[a][0].y, S[a][1].x, S[a][1].y);
        *reinterpret_cast<float4*>(s + (t.i0 + a) * N + t.jB) = make_float4(S[a][2].x, S[a][2].y, S[a][3].x, S[a][3].y);
    }
}

// =======================================================================================
// forward
// =======================================================================================
constexpr int PSTR = 68;   // padded row of `part`: conflict-free 128-bit stores from 8 j-groups

template <typename IO>
struct FwdSmem {
    static constexpr int NS = io_traits<IO>::kStages;
    alignas(128) IO raw[NS][4][TT * N];   // TMA landing zone: r, k, v, w
    alignas(16) float f[2][4][TT * N];    // fp32: r, k, v, d = exp(-exp(w))
    alignas(16) float part[TT * 8 * PSTR]; // per-step partial y of the 8 j-groups, rows padded to 68 floats
    float bonus[2][TT];                   // sum_j r u k per step
    alignas(8) uint64_t full[NS];
};

template <typename IO>
__global__ void __launch_bounds__(NT, 2) wkv6_fwd_scan_kernel(const __grid_constant__ Params p) {
    using tr = io_traits<IO>;
    using SM = FwdSmem<IO>;
    constexpr int NS = SM::NS;
    extern __shared__ __align__(1024) uint8_t smem_bytes[];
    SM& sm = *reinterpret_cast<SM*>(smem_bytes);

    const int tid = threadIdx.x;
    const Tile tl(tid);
    const int BH = p.B * p.H;
    const int nseg = p.nseg > 1 ? p.nseg : 1;
    const int seq = blockIdx.x / nseg, seg = blockIdx.x - seq * nseg;
    const int ds = seq / BH;
    const int bh = seq - ds * BH;
    const int b = bh / p.H, h = bh - b * p.H;
    const DirArgs& D = p.d[ds];
    const bool rev = D.reverse != 0;
    const int T = p.T, C = p.C;
    const int ntot = (T + TT - 1) / TT;

    // The CTA walks the tiles [tbeg, tbeg + ntiles) of its sequence in processing order.  Time-segmented launch (long
    // sequences, few of them; see launch_fwd_scan_segmented): pass 1 starts every segment from S = 0, writes no y and
    // leaves the segment's final state and the product of its per-step decays in the scratch; pass 2 starts from the
    // segment-start state the scan kernel put there.
    int tbeg = 0, ntiles = ntot;
    const float* s0p = D.s0 ? D.s0 + (size_t)bh * N * N : nullptr;
    float* sTp = D.sT ? D.sT + (size_t)bh * N * N : nullptr;
    float* decay_out = nullptr;
    bool write_y = true;
    if (p.nseg > 1) {
        const int per = p.seg_chunks * (64 / TT);
        tbeg = seg * per;
        ntiles = min(per, ntot - tbeg);
        float* slot = p.seg_state + ((size_t)seq * nseg + seg) * N * N;
        if (p.seg_mode == 1) {
            s0p = nullptr; sTp = slot; write_y = false;
            decay_out = p.seg_decay + ((size_t)seq * nseg + seg) * N;
        } else {
            s0p = slot;
            if (seg != nseg - 1) sTp = nullptr;
        }
    }

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < NS; s++) mbar_init(&sm.full[s], 1);
        fence_barrier_init();
        tma_prefetch_desc(&D.mr); tma_prefetch_desc(&D.mk); tma_prefetch_desc(&D.mv); tma_prefetch_desc(&D.mw);
    }
    __syncthreads();

    auto issue = [&](int n) {   // n: local tile index
        const int q = rev ? ntot - 1 - (tbeg + n) : tbeg + n;
        const int slot = n % NS;
        mbar_arrive_expect_tx(&sm.full[slot], 4u * TT * N * sizeof(IO));
        tma_load_3d(sm.raw[slot][0], &D.mr, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][1], &D.mk, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][2], &D.mv, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][3], &D.mw, &sm.full[slot], h * N, q * TT, b);
    };
    if (tid == 0)
        for (int n = 0; n < NS && n < ntiles; n++) issue(n);

    // per-thread slice of a tile in the conversion passes: row prow, channels pc0..pc0+7
    const int prow = tid >> 3, pc0 = (tid & 7) * 8;
    float u8[8];
    tr::load8(reinterpret_cast<const IO*>(D.u) + h * N + pc0, u8);

    float2 S[4][4];
    load_state(S, s0p, tl);
    float2 dprod[4];   // product of the per-step decays of this thread's 8 key channels (segmented pass 1)
#pragma unroll
    for (int qq = 0; qq < 4; qq++) dprod[qq] = make_float2(1.f, 1.f);

    auto preprocess = [&](int n) {
        const int slot = n % NS, buf = n & 1;
        mbar_wait(&sm.full[slot], (n / NS) & 1);
        float r8[8], k8[8], v8[8], w8[8];
        tr::load8(&sm.raw[slot][0][prow * N + pc0], r8);
        tr::load8(&sm.raw[slot][1][prow * N + pc0], k8);
        tr::load8(&sm.raw[slot][2][prow * N + pc0], v8);
        tr::load8(&sm.raw[slot][3][prow * N + pc0], w8);
        float part = 0.f;
#pragma unroll
        for (int e = 0; e < 8; e++) {
            part = fmaf(r8[e] * u8[e], k8[e], part);
            w8[e] = expf(-expf(w8[e]));
        }
        st8f(&sm.f[buf][0][prow * N + pc0], r8);
        st8f(&sm.f[buf][1][prow * N + pc0], k8);
        st8f(&sm.f[buf][2][prow * N + pc0], v8);
        st8f(&sm.f[buf][3][prow * N + pc0], w8);
        part = sum_lanes8(part);
        if ((tid & 7) == 0) sm.bonus[buf][prow] = part;
    };

    IO* const ybase = reinterpret_cast<IO*>(D.y) + (size_t)b * T * C + h * N + pc0;

    // Two barriers per 16-step tile.  Inside the step loop there is no cross-lane traffic at all: every thread
    // parks its four partial sums (over its 8 j) in `part`, and the epilogue adds the 8 j-groups.  With one warp
    // per scheduler per CTA, shuffle latency in the loop was the bound (215 us -> see profiles/).
    auto step = [&](int buf, int s) {
        const float* fr = sm.f[buf][0] + s * N;
        const float* fk = sm.f[buf][1] + s * N;
        const float* fv = sm.f[buf][2] + s * N;
        const float* fd = sm.f[buf][3] + s * N;
        float2 r2[4], k2[4], d2[4];
        split4(*reinterpret_cast<const float4*>(fr + tl.jA), *reinterpret_cast<const float4*>(fr + tl.jB), r2);
        split4(*reinterpret_cast<const float4*>(fk + tl.jA), *reinterpret_cast<const float4*>(fk + tl.jB), k2);
        split4(*reinterpret_cast<const float4*>(fd + tl.jA), *reinterpret_cast<const float4*>(fd + tl.jB), d2);
        const float4 v4 = *reinterpret_cast<const float4*>(fv + tl.i0);
        const float va[4] = {v4.x, v4.y, v4.z, v4.w};
        float ya[4];
#pragma unroll
        for (int a = 0; a < 4; a++) {
            const float2 vv = make_float2(va[a], va[a]);
            float2 acc = make_float2(0.f, 0.f);
#pragma unroll
            for (int qq = 0; qq < 4; qq++) {
                acc = __ffma2_rn(r2[qq], S[a][qq], acc);
                const float2 x = __fmul2_rn(k2[qq], vv);
                S[a][qq] = __ffma2_rn(S[a][qq], d2[qq], x);
            }
            ya[a] = acc.x + acc.y;
        }
        if (decay_out) {
#pragma unroll
            for (int qq = 0; qq < 4; qq++) dprod[qq] = __fmul2_rn(dprod[qq], d2[qq]);
        }
        *reinterpret_cast<float4*>(&sm.part[(s * 8 + tl.jg) * PSTR + tl.i0]) = make_float4(ya[0], ya[1], ya[2], ya[3]);
    };

    if (ntiles > 0) preprocess(0);
    for (int n = 0; n < ntiles; n++) {
        const int buf = n & 1;
        const int q = rev ? ntot - 1 - (tbeg + n) : tbeg + n;
        const int nvalid = min(TT, T - q * TT);
        __syncthreads();   // f[buf] complete; previous epilogue done with `part`
        if (tid == 0 && n + NS < ntiles) issue(n + NS);
        if (nvalid == TT) {
            if (rev) {
#pragma unroll 4
                for (int m = TT - 1; m >= 0; m--) step(buf, m);
            } else {
#pragma unroll 4
                for (int m = 0; m < TT; m++) step(buf, m);
            }
        } else {
            for (int m = 0; m < nvalid; m++) step(buf, rev ? nvalid - 1 - m : m);
        }
        __syncthreads();
        if (write_y && prow < nvalid) {
            float o8[8], a8[8];
            ld8f(&sm.f[buf][2][prow * N + pc0], o8);
            const float bn = sm.bonus[buf][prow];
#pragma unroll
            for (int e = 0; e < 8; e++) o8[e] *= bn;
#pragma unroll
            for (int g = 0; g < 8; g++) {
                ld8f(&sm.part[(prow * 8 + g) * PSTR + pc0], a8);
#pragma unroll
                for (int e = 0; e < 8; e++) o8[e] += a8[e];
            }
            tr::store8(ybase + (size_t)(q * TT + prow) * C, o8);
        }
        if (n + 1 < ntiles) preprocess(n + 1);
    }
    if (sTp) store_state(S, sTp, tl);
    if (decay_out && tl.i0 == 0) {
        *reinterpret_cast<float4*>(decay_out + tl.jA) = make_float4(dprod[0].x, dprod[0].y, dprod[1].x, dprod[1].y);
        *reinterpret_cast<float4*>(decay_out + tl.jB) = make_float4(dprod[2].x, dprod[2].y, dprod[3].x, dprod[3].y);
    }
}

// =======================================================================================
// backward, sweep A: forward-time state S -> gr, gu (and hT when a final-state gradient exists)
// =======================================================================================
template <typename IO>
struct BwdASmem {
    static constexpr int NS = io_traits<IO>::kStages;
    alignas(128) IO raw[NS][5][TT * N];   // r, k, v, w, gy
    alignas(16) float f[2][4][TT * N];    // k, v, d, gy
    alignas(16) float part[4][TT * N];    // per-warp partial sums over i of gy_i S_ij
    float vg[2][TT];                      // sum_i v_i gy_i per step
    alignas(8) uint64_t full[NS];
};

template <typename IO>
__global__ void __launch_bounds__(NT, 2) wkv6_bwd_a_kernel(const __grid_constant__ Params p) {
    using tr = io_traits<IO>;
    using SM = BwdASmem<IO>;
    constexpr int NS = SM::NS;
    extern __shared__ __align__(1024) uint8_t smem_bytes[];
    SM& sm = *reinterpret_cast<SM*>(smem_bytes);

    const int tid = threadIdx.x;
    const Tile tl(tid);
    const int BH = p.B * p.H;
    const int ds = blockIdx.x / BH;
    const int bh = blockIdx.x - ds * BH;
    const int b = bh / p.H, h = bh - b * p.H;
    const DirArgs& D = p.d[ds];
    const bool rev = D.reverse != 0;
    const int T = p.T, C = p.C;
    const int ntiles = (T + TT - 1) / TT;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < NS; s++) mbar_init(&sm.full[s], 1);
        fence_barrier_init();
    }
    __syncthreads();

    auto issue = [&](int n) {
        const int q = rev ? ntiles - 1 - n : n;
        const int slot = n % NS;
        mbar_arrive_expect_tx(&sm.full[slot], 5u * TT * N * sizeof(IO));
        tma_load_3d(sm.raw[slot][0], &D.mr, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][1], &D.mk, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][2], &D.mv, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][3], &D.mw, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][4], &D.mg, &sm.full[slot], h * N, q * TT, b);
    };
    if (tid == 0)
        for (int n = 0; n < NS && n < ntiles; n++) issue(n);

    const int prow = tid >> 3, pc0 = (tid & 7) * 8;
    float u8[8], gu8[8];
    tr::load8(reinterpret_cast<const IO*>(D.u) + h * N + pc0, u8);
#pragma unroll
    for (int e = 0; e < 8; e++) gu8[e] = 0.f;

    float2 S[4][4];
    load_state(S, D.s0 ? D.s0 + (size_t)bh * N * N : nullptr, tl);

    auto preprocess = [&](int n) {
        const int slot = n % NS, buf = n & 1;
        mbar_wait(&sm.full[slot], (n / NS) & 1);
        float r8[8], k8[8], v8[8], w8[8], g8[8];
        tr::load8(&sm.raw[slot][0][prow * N + pc0], r8);
        tr::load8(&sm.raw[slot][1][prow * N + pc0], k8);
        tr::load8(&sm.raw[slot][2][prow * N + pc0], v8);
        tr::load8(&sm.raw[slot][3][prow * N + pc0], w8);
        tr::load8(&sm.raw[slot][4][prow * N + pc0], g8);
        float vg = 0.f;
#pragma unroll
        for (int e = 0; e < 8; e++) {
            vg = fmaf(v8[e], g8[e], vg);
            w8[e] = expf(-expf(w8[e]));
        }
        vg = sum_lanes8(vg);
#pragma unroll
        for (int e = 0; e < 8; e++) gu8[e] = fmaf(r8[e] * k8[e], vg, gu8[e]);   // OOB rows are all-zero
        st8f(&sm.f[buf][0][prow * N + pc0], k8);
        st8f(&sm.f[buf][1][prow * N + pc0], v8);
        st8f(&sm.f[buf][2][prow * N + pc0], w8);
        st8f(&sm.f[buf][3][prow * N + pc0], g8);
        if ((tid & 7) == 0) sm.vg[buf][prow] = vg;
    };

    IO* const grbase = reinterpret_cast<IO*>(D.gr) + (size_t)b * T * C + h * N + pc0;
    const int jsel = reduce8_index(tl);

    if (ntiles > 0) preprocess(0);
    for (int n = 0; n < ntiles; n++) {
        const int buf = n & 1;
        const int q = rev ? ntiles - 1 - n : n;
        const int nvalid = min(TT, T - q * TT);
        __syncthreads();   // f[buf] complete; previous epilogue done with `part`
        if (tid == 0 && n + NS < ntiles) issue(n + NS);

        for (int m = 0; m < nvalid; m++) {
            const int s = rev ? nvalid - 1 - m : m;
            const float* fk = sm.f[buf][0] + s * N;
            const float* fv = sm.f[buf][1] + s * N;
            const float* fd = sm.f[buf][2] + s * N;
            const float* fg = sm.f[buf][3] + s * N;
            float2 k2[4], d2[4];
            split4(*reinterpret_cast<const float4*>(fk + tl.jA), *reinterpret_cast<const float4*>(fk + tl.jB), k2);
            split4(*reinterpret_cast<const float4*>(fd + tl.jA), *reinterpret_cast<const float4*>(fd + tl.jB), d2);
            const float4 v4 = *reinterpret_cast<const float4*>(fv + tl.i0);
            const float4 g4 = *reinterpret_cast<const float4*>(fg + tl.i0);
            const float va[4] = {v4.x, v4.y, v4.z, v4.w};
            const float ga[4] = {g4.x, g4.y, g4.z, g4.w};
            float2 acc[4];
#pragma unroll
            for (int qq = 0; qq < 4; qq++) acc[qq] = make_float2(0.f, 0.f);
#pragma unroll
            for (int a = 0; a < 4; a++) {
                const float2 vv = make_float2(va[a], va[a]);
                const float2 gg = make_float2(ga[a], ga[a]);
#pragma unroll
                for (int qq = 0; qq < 4; qq++) {
                    acc[qq] = __ffma2_rn(gg, S[a][qq], acc[qq]);
                    const float2 x = __fmul2_rn(k2[qq], vv);
                    S[a][qq] = __ffma2_rn(S[a][qq], d2[qq], x);
                }
            }
            const float x8[8] = {acc[0].x, acc[0].y, acc[1].x, acc[1].y, acc[2].x, acc[2].y, acc[3].x, acc[3].y};
            const float2 kp = reduce8_over_i(x8, tl.lane);
            *reinterpret_cast<float2*>(&sm.part[tl.warp][s * N + jsel]) = kp;
        }
        __syncthreads();
        if (prow < nvalid) {
            float o8[8], a8[8], k8[8];
            ld8f(&sm.part[0][prow * N + pc0], o8);
#pragma unroll
            for (int wv = 1; wv < 4; wv++) {
                ld8f(&sm.part[wv][prow * N + pc0], a8);
#pragma unroll
                for (int e = 0; e < 8; e++) o8[e] += a8[e];
            }
            ld8f(&sm.f[buf][0][prow * N + pc0], k8);
            const float vg = sm.vg[buf][prow];
#pragma unroll
            for (int e = 0; e < 8; e++) o8[e] = fmaf(u8[e] * k8[e], vg, o8[e]);
            tr::store8(grbase + (size_t)(q * TT + prow) * C, o8);
            if constexpr (sizeof(IO) == 2) {
                if (D.grlo) {   // low-order part of gr for sweep B (bf16 gr alone costs gw 2^-9 per summed term)
                    float lo8[8];
#pragma unroll
                    for (int e = 0; e < 8; e++) lo8[e] = o8[e] - tr::to_float(tr::from_float(o8[e]));
                    tr::store8(D.grlo + (size_t)b * T * C + h * N + pc0 + (size_t)(q * TT + prow) * C, lo8);
                }
            }
        }
        if (n + 1 < ntiles) preprocess(n + 1);
    }

    // gu[b][h*64 + j] = sum_t r k (v.gy): reduce the 16 row-groups through `part`
    __syncthreads();
    st8f(&sm.part[0][prow * N + pc0], gu8);
    __syncthreads();
    if (tid < N) {
        float g = 0.f;
#pragma unroll
        for (int rr = 0; rr < TT; rr++) g += sm.part[0][rr * N + tid];
        reinterpret_cast<IO*>(D.gu)[(size_t)b * C + h * N + tid] = tr::from_float(g);
    }
    if (D.gsT && D.hT) {   // hT[j] = sum_i S_T[i][j] gsT[i][j]
        const float* gs = D.gsT + (size_t)bh * N * N;
        float x8[8];
#pragma unroll
        for (int e = 0; e < 8; e++) x8[e] = 0.f;
#pragma unroll
        for (int a = 0; a < 4; a++) {
            const float4 ga = *reinterpret_cast<const float4*>(gs + (tl.i0 + a) * N + tl.jA);
            const float4 gb = *reinterpret_cast<const float4*>(gs + (tl.i0 + a) * N + tl.jB);
            x8[0] = fmaf(S[a][0].x, ga.x, x8[0]); x8[1] = fmaf(S[a][0].y, ga.y, x8[1]);
            x8[2] = fmaf(S[a][1].x, ga.z, x8[2]); x8[3] = fmaf(S[a][1].y, ga.w, x8[3]);
            x8[4] = fmaf(S[a][2].x, gb.x, x8[4]); x8[5] = fmaf(S[a][2].y, gb.y, x8[5]);
            x8[6] = fmaf(S[a][3].x, gb.z, x8[6]); x8[7] = fmaf(S[a][3].y, gb.w, x8[7]);
        }
        const float2 kp = reduce8_over_i(x8, tl.lane);
        __syncthreads();
        *reinterpret_cast<float2*>(&sm.part[tl.warp][jsel]) = kp;
        __syncthreads();
        if (tid < N)
            D.hT[(size_t)bh * N + tid] =
                sm.part[0][tid] + sm.part[1][tid] + sm.part[2][tid] + sm.part[3][tid];
    }
}

// =======================================================================================
// backward, sweep B: reverse-time state C -> gk, gv, gw, gs0
// =======================================================================================
template <typename IO>
struct BwdBSmem {
    static constexpr int NS = io_traits<IO>::kStages;
    alignas(128) IO raw[NS][6][TT * N];   // r, k, v, w, gy, gr
    alignas(16) float f[2][7][TT * N];    // r, k, v, d, gy, a = r*grS, ew = exp(w)
    alignas(16) float part[4][TT * N];    // per-warp partial sums over i of v_i C_ij
    alignas(16) float outv[TT * N];       // sum_j k_j C_ij
    float vg[2][TT];
    float bonus[2][TT];
    alignas(8) uint64_t full[NS];
};

template <typename IO>
__global__ void __launch_bounds__(NT, sizeof(IO) == 2 ? 2 : 1) wkv6_bwd_b_kernel(const __grid_constant__ Params p) {
    using tr = io_traits<IO>;
    using SM = BwdBSmem<IO>;
    constexpr int NS = SM::NS;
    extern __shared__ __align__(1024) uint8_t smem_bytes[];
    SM& sm = *reinterpret_cast<SM*>(smem_bytes);

    const int tid = threadIdx.x;
    const Tile tl(tid);
    const int BH = p.B * p.H;
    const int ds = blockIdx.x / BH;
    const int bh = blockIdx.x - ds * BH;
    const int b = bh / p.H, h = bh - b * p.H;
    const DirArgs& D = p.d[ds];
    const bool brev = D.reverse == 0;   // this sweep runs against the direction's own time order
    const int T = p.T, C = p.C;
    const int ntiles = (T + TT - 1) / TT;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < NS; s++) mbar_init(&sm.full[s], 1);
        fence_barrier_init();
    }
    __syncthreads();

    auto issue = [&](int n) {
        const int q = brev ? ntiles - 1 - n : n;
        const int slot = n % NS;
        mbar_arrive_expect_tx(&sm.full[slot], 6u * TT * N * sizeof(IO));
        tma_load_3d(sm.raw[slot][0], &D.mr, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][1], &D.mk, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][2], &D.mv, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][3], &D.mw, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][4], &D.mg, &sm.full[slot], h * N, q * TT, b);
        tma_load_3d(sm.raw[slot][5], &D.mq, &sm.full[slot], h * N, q * TT, b);
    };
    if (tid == 0)
        for (int n = 0; n < NS && n < ntiles; n++) issue(n);

    const int prow = tid >> 3, pc0 = (tid & 7) * 8;
    float u8[8];
    tr::load8(reinterpret_cast<const IO*>(D.u) + h * N + pc0, u8);

    float2 Cs[4][4];
    load_state(Cs, D.gsT ? D.gsT + (size_t)bh * N * N : nullptr, tl);

    auto preprocess = [&](int n) {
        const int slot = n % NS, buf = n & 1;
        mbar_wait(&sm.full[slot], (n / NS) & 1);
        float r8[8], k8[8], v8[8], w8[8], g8[8], q8[8];
        tr::load8(&sm.raw[slot][0][prow * N + pc0], r8);
        tr::load8(&sm.raw[slot][1][prow * N + pc0], k8);
        tr::load8(&sm.raw[slot][2][prow * N + pc0], v8);
        tr::load8(&sm.raw[slot][3][prow * N + pc0], w8);
        tr::load8(&sm.raw[slot][4][prow * N + pc0], g8);
        tr::load8(&sm.raw[slot][5][prow * N + pc0], q8);
        if constexpr (sizeof(IO) == 2) {
            const int qq = brev ? ntiles - 1 - n : n;
            if (D.grlo && qq * TT + prow < T) {
                float lo8[8];
                tr::load8(D.grlo + (size_t)b * T * C + h * N + pc0 + (size_t)(qq * TT + prow) * C, lo8);
#pragma unroll
                for (int e = 0; e < 8; e++) q8[e] += lo8[e];
            }
        }
        float vg = 0.f, bn = 0.f, ew8[8];
#pragma unroll
        for (int e = 0; e < 8; e++) {
            vg = fmaf(v8[e], g8[e], vg);
            bn = fmaf(r8[e] * u8[e], k8[e], bn);
            ew8[e] = expf(w8[e]);
            w8[e] = expf(-ew8[e]);
        }
        vg = sum_lanes8(vg);
        bn = sum_lanes8(bn);
#pragma unroll
        for (int e = 0; e < 8; e++) q8[e] = r8[e] * fmaf(-u8[e] * k8[e], vg, q8[e]);   // a = r * (gr - u k (v.gy))
        st8f(&sm.f[buf][0][prow * N + pc0], r8);
        st8f(&sm.f[buf][1][prow * N + pc0], k8);
        st8f(&sm.f[buf][2][prow * N + pc0], v8);
        st8f(&sm.f[buf][3][prow * N + pc0], w8);
        st8f(&sm.f[buf][4][prow * N + pc0], g8);
        st8f(&sm.f[buf][5][prow * N + pc0], q8);
        st8f(&sm.f[buf][6][prow * N + pc0], ew8);
        if ((tid & 7) == 0) {
            sm.vg[buf][prow] = vg;
            sm.bonus[buf][prow] = bn;
        }
    };

    const size_t gbase = (size_t)b * T * C + h * N;
    IO* const gkbase = reinterpret_cast<IO*>(D.gk) + gbase + pc0;
    IO* const gvbase = reinterpret_cast<IO*>(D.gv) + gbase + pc0;
    const int col = tid >> 1, half = tid & 1;
    IO* const gwbase = reinterpret_cast<IO*>(D.gw) + gbase + col;
    const int jsel = reduce8_index(tl);
    const int asel = reduce4_index(tl.lane);
    // X = H_{m+1} + a_{m+1} in the direction's own step order (H beyond the last step: hT or 0)
    float X = (D.gsT && D.hT) ? D.hT[(size_t)bh * N + col] : 0.f;

    if (ntiles > 0) preprocess(0);
    for (int n = 0; n < ntiles; n++) {
        const int buf = n & 1;
        const int q = brev ? ntiles - 1 - n : n;
        const int nvalid = min(TT, T - q * TT);
        __syncthreads();
        if (tid == 0 && n + NS < ntiles) issue(n + NS);

        for (int m = 0; m < nvalid; m++) {
            const int s = brev ? nvalid - 1 - m : m;
            const float* fr = sm.f[buf][0] + s * N;
            const float* fk = sm.f[buf][1] + s * N;
            const float* fv = sm.f[buf][2] + s * N;
            const float* fd = sm.f[buf][3] + s * N;
            const float* fg = sm.f[buf][4] + s * N;
            float2 r2[4], k2[4], d2[4];
            split4(*reinterpret_cast<const float4*>(fr + tl.jA), *reinterpret_cast<const float4*>(fr + tl.jB), r2);
            split4(*reinterpret_cast<const float4*>(fk + tl.jA), *reinterpret_cast<const float4*>(fk + tl.jB), k2);
            split4(*reinterpret_cast<const float4*>(fd + tl.jA), *reinterpret_cast<const float4*>(fd + tl.jB), d2);
            const float4 v4 = *reinterpret_cast<const float4*>(fv + tl.i0);
            const float4 g4 = *reinterpret_cast<const float4*>(fg + tl.i0);
            const float va[4] = {v4.x, v4.y, v4.z, v4.w};
            const float ga[4] = {g4.x, g4.y, g4.z, g4.w};
            float2 accK[4];
            float yv[4];
#pragma unroll
            for (int qq = 0; qq < 4; qq++) accK[qq] = make_float2(0.f, 0.f);
#pragma unroll
            for (int a = 0; a < 4; a++) {
                const float2 vv = make_float2(va[a], va[a]);
                const float2 gg = make_float2(ga[a], ga[a]);
                float2 accv = make_float2(0.f, 0.f);
#pragma unroll
                for (int qq = 0; qq < 4; qq++) {
                    accK[qq] = __ffma2_rn(vv, Cs[a][qq], accK[qq]);
                    accv = __ffma2_rn(k2[qq], Cs[a][qq], accv);
                    const float2 x = __fmul2_rn(r2[qq], gg);
                    Cs[a][qq] = __ffma2_rn(Cs[a][qq], d2[qq], x);
                }
                yv[a] = accv.x + accv.y;
            }
            const float gvc = reduce4_over_j(yv, tl.lane);
            if (!(tl.lane & 4)) sm.outv[s * N + tl.i0 + asel] = gvc;
            const float x8[8] = {accK[0].x, accK[0].y, accK[1].x, accK[1].y, accK[2].x, accK[2].y, accK[3].x, accK[3].y};
            const float2 kp = reduce8_over_i(x8, tl.lane);
            *reinterpret_cast<float2*>(&sm.part[tl.warp][s * N + jsel]) = kp;
        }
        __syncthreads();

        // ---- epilogue, phase 1: gk, gv (row x 8 channels per thread, 128-bit stores)
        if (prow < nvalid) {
            float o8[8], a8[8], x8[8];
            ld8f(&sm.part[0][prow * N + pc0], o8);
#pragma unroll
            for (int wv = 1; wv < 4; wv++) {
                ld8f(&sm.part[wv][prow * N + pc0], a8);
#pragma unroll
                for (int e = 0; e < 8; e++) o8[e] += a8[e];
            }
            ld8f(&sm.f[buf][0][prow * N + pc0], x8);   // r
            const float vg = sm.vg[buf][prow];
#pragma unroll
            for (int e = 0; e < 8; e++) o8[e] = fmaf(u8[e] * x8[e], vg, o8[e]);
            tr::store8(gkbase + (size_t)(q * TT + prow) * C, o8);
            ld8f(&sm.outv[prow * N + pc0], o8);
            ld8f(&sm.f[buf][4][prow * N + pc0], x8);   // gy
            const float bn = sm.bonus[buf][prow];
#pragma unroll
            for (int e = 0; e < 8; e++) o8[e] = fmaf(bn, x8[e], o8[e]);
            tr::store8(gvbase + (size_t)(q * TT + prow) * C, o8);
        }
        // ---- epilogue, phase 2: gw by the suffix-sum identity (channel `col`, 8 steps per thread)
        {
            float av[8], kg[8], ew[8];
            float delta = 0.f;
#pragma unroll
            for (int mm = 0; mm < 8; mm++) {
                const int m = half * 8 + mm;
                av[mm] = 0.f; kg[mm] = 0.f; ew[mm] = 0.f;
                if (m < nvalid) {
                    const int s = brev ? nvalid - 1 - m : m;
                    const int o = s * N + col;
                    const float gkc = (sm.part[0][o] + sm.part[1][o]) + (sm.part[2][o] + sm.part[3][o]);
                    kg[mm] = sm.f[buf][1][o] * gkc;
                    av[mm] = sm.f[buf][5][o];
                    ew[mm] = sm.f[buf][6][o];
                    delta += av[mm] - kg[mm];
                }
            }
            const float other = __shfl_xor_sync(0xffffffffu, delta, 1);
            float Xin = half ? X + other : X;
#pragma unroll
            for (int mm = 0; mm < 8; mm++) {
                const int m = half * 8 + mm;
                if (m < nvalid) {
                    const int s = brev ? nvalid - 1 - m : m;
                    const float Hm = Xin - kg[mm];
                    // the direction's first step sees S_0: without an initial state gw is exactly 0 there
                    // (the reference hard-codes it, wkv6_cuda.cu:237); the suffix sum only leaves rounding residue
                    const int t = q * TT + s;
                    const bool first = !D.s0 && (D.reverse ? t == T - 1 : t == 0);
                    gwbase[(size_t)t * C] = tr::from_float(first ? 0.f : -ew[mm] * Hm);
                    Xin = Hm + av[mm];
                }
            }
            X += delta + other;
        }
        if (n + 1 < ntiles) preprocess(n + 1);
    }
    if (D.gs0) store_state(Cs, D.gs0 + (size_t)bh * N * N, tl);
}

// =======================================================================================
// combine:  out = (a + b) * scale  (fp32 out)   -- rwkv_wrapper_bidirectional.py:48-56
// =======================================================================================
template <typename IO>
__global__ void __launch_bounds__(256) wkv6_combine_kernel(size_t n8, const IO* __restrict__ a, const IO* __restrict__ b,
                                                           float scale, int round_sum, float* __restrict__ out) {
    using tr = io_traits<IO>;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        float x[8], y[8];
        tr::load8(a + i * 8, x);
        if (b) {
            tr::load8(b + i * 8, y);
#pragma unroll
            for (int e = 0; e < 8; e++) x[e] += y[e];
            if (round_sum) {
#pragma unroll
                for (int e = 0; e < 8; e++) x[e] = tr::to_float(tr::from_float(x[e]));
            }
        }
#pragma unroll
        for (int e = 0; e < 8; e++) x[e] *= scale;
        st8f(out + i * 8, x);
    }
}

// =======================================================================================
// launchers
// =======================================================================================
template <typename K>
static cudaError_t set_smem(K kernel, size_t bytes) {
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

template <typename IO>
static int launch_fwd_t(const Params& p, cudaStream_t st) {
    const size_t smem = sizeof(FwdSmem<IO>);
    cudaError_t e = set_smem(wkv6_fwd_scan_kernel<IO>, smem);
    if (e != cudaSuccess) return (int)e;
    wkv6_fwd_scan_kernel<IO><<<dim3(p.ndir * p.B * p.H * (p.nseg > 1 ? p.nseg : 1)), dim3(NT), smem, st>>>(p);
    count_launch(1);
    return (int)cudaGetLastError();
}

int launch_fwd_scan(int dtype, const Params& p, cudaStream_t st) {
    return dtype == 0 ? launch_fwd_t<__nv_bfloat16>(p, st) : launch_fwd_t<float>(p, st);
}

// Time-segmented exact scan (same scheme as launch_fwd_chunk_segmented, same scratch layout and segment rule): pass 1,
// segment scan, pass 2.  Returns -1 when the launch does not qualify.
int launch_fwd_scan_segmented(int dtype, const Params& p, void* scratch, size_t scratch_bytes, cudaStream_t st) {
    Params q = p;
    if (!plan_segments(q, scratch, scratch_bytes, 4)) return -1;
    q.seg_mode = 1;
    int e = launch_fwd_scan(dtype, q, st);
    if (e) return e;
    if ((e = launch_seg_scan(q, st))) return e;
    q.seg_mode = 2;
    return launch_fwd_scan(dtype, q, st);
}

template <typename IO>
static int launch_bwd_t(const Params& p, cudaStream_t st) {
    const size_t sa = sizeof(BwdASmem<IO>), sb = sizeof(BwdBSmem<IO>);
    cudaError_t e = set_smem(wkv6_bwd_a_kernel<IO>, sa);
    if (e != cudaSuccess) return (int)e;
    e = set_smem(wkv6_bwd_b_kernel<IO>, sb);
    if (e != cudaSuccess) return (int)e;
    const dim3 grid(p.ndir * p.B * p.H);
    wkv6_bwd_a_kernel<IO><<<grid, dim3(NT), sa, st>>>(p);
    e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    wkv6_bwd_b_kernel<IO><<<grid, dim3(NT), sb, st>>>(p);
    count_launch(2);
    return (int)cudaGetLastError();
}

int launch_bwd_scan(int dtype, const Params& p, cudaStream_t st) {
    return dtype == 0 ? launch_bwd_t<__nv_bfloat16>(p, st) : launch_bwd_t<float>(p, st);
}

int launch_combine(int dtype, size_t n, const void* a, const void* b, float scale, int round_sum, float* out, cudaStream_t st) {
    const size_t n8 = n / 8;
    if (n8 == 0) return 0;
    const int blocks = (int)((n8 + 255) / 256 < 148 * 8 ? (n8 + 255) / 256 : 148 * 8);
    if (dtype == 0)
        wkv6_combine_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(n8, (const __nv_bfloat16*)a, (const __nv_bfloat16*)b, scale, round_sum, out);
    else
        wkv6_combine_kernel<float><<<blocks, 256, 0, st>>>(n8, (const float*)a, (const float*)b, scale, 0, out);
    count_launch(1);
    return (int)cudaGetLastError();
}

}  // namespace wkv6
