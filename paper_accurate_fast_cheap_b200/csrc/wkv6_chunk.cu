// wkv6_chunk.cu -- chunked (dense-contraction) WKV6 forward for bf16 I/O on tcgen05 / TMEM / TMA (sm_100a).
//
// Same function as wkv6_fwd_scan_kernel (reference: wenet/rwkv_v6/cuda/wkv6_cuda.cu:9-63), restated per chunk of
// L = 64 steps so that the O(L*N*N) work runs on the 5th-gen tensor cores:
//     l_t = -exp(w_t) (log2 units, clamped at -5*log2e), A_t = exclusive cumsum, A+_t = inclusive cumsum
//     P[p][q]  = sum_j r_p[j] k_q[j] 2^(A_p[j] - A+_q[j])   (q before p),  P[p][p] = sum_j r_p u k_p
//     Y        = P V  +  (r_p 2^A_p) S^T                     S' = S diag(2^A_L) + V^T (k_q 2^(A_L - A+_q))
// Overflow-free factorisation: the chunk is cut into four 16-step sub-blocks; every factor is taken relative to
// the START of the row's sub-block I:  Rtil_p = r_p 2^(A_p - A_BI) (<= |r|),  Ktil(I)_q = k_q 2^(A_BI - A+_q)
// (<= |k| for earlier sub-blocks, <= 2^(16*7.2) inside the diagonal sub-block: finite in bf16/fp32).
//
// One CTA (256 threads) per (direction, batch, head); two CTAs per SM overlap each other's TMA / MMA latency.
// This file holds the first version of the kernel (kept behind WKV6_FWD_CHUNK_V1 for A/B runs); the default is the
// software-pipelined, packed-bf16 version in wkv6_chunk_lean.cu.
//   G1 (x4): P^T[:, block I] = Ktil(I) [64 q x 64 j, K-major A]  x  Rtil[block I] [16 p x 64 j, K-major B]   M64 N16 K64
//   G3     : Y   = Rst [64 p x 64 j, K-major A] x S' [64 j x 64 i, MN-major B]                              M64 N64 K64
//   G2     : Y  += P^T tile [64 q x 64 p, MN-major A] x V [64 q x 64 i, MN-major B, straight from TMA]
//   G4     : dS^T[j][i] = Ktil(4) [64 q x 64 j, MN-major A] x V [MN-major B]
// All tiles are [64][64] bf16 with 128-byte rows and the 128B swizzle, so the same bytes serve as K-major or
// MN-major operands (conventions verified on hardware by tests/probes/umma_probe.cu).  Accumulators live in TMEM
// (P^T, Y, dS^T: 3 x 64 columns) and are read with tcgen05.ld.16x256b; the fp32 state S^T[j][i] stays in
// registers in that fragment layout for the whole sequence.
#include "wkv6_umma.cuh"
#include <cstdlib>

namespace wkv6 {

// -------------------------------------------------------------------------------------------------------------------
// 256-thread version: same algorithm and tiles, 8 warps per CTA and two CTAs per SM (<= 128 registers per thread), i.e.
// four warps per scheduler instead of two -- the 128-thread kernel above is latency-bound at ~30 % issue utilisation.
// Row-major phases: warp = (sub-block I, channel half), thread = 4 rows x 4 channels.  TMEM fragments: warp =
// (lane quarter, column half), 16 registers per 64-column accumulator.
// -------------------------------------------------------------------------------------------------------------------
struct Chunk256Smem {
    alignas(1024) uint8_t raw_r[CL * 128];
    alignas(1024) uint8_t raw_k[CL * 128];
    alignas(1024) uint8_t raw_w[CL * 128];
    alignas(1024) uint8_t v[2][CL * 128];
    alignas(1024) uint8_t roff[CL * 128];
    alignas(1024) uint8_t rst[CL * 128];
    alignas(1024) uint8_t kt[28 * 1024];
    alignas(1024) uint8_t pt[CL * 128];      // masked P^T [q][p]; reused as the y staging tile [p][i]
    alignas(1024) uint8_t sp[N * 128];       // bf16 state operand S^T [j][i]; also the source of the checkpoint store
    float blocktot[4][N];
    float bonus[2][CL];
    alignas(8) uint64_t bar_rkw, bar_v[2], bar_p, bar_s, bar_y;
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(256, 2) wkv6_fwd_chunk_kernel(const __grid_constant__ Params p) {
    extern __shared__ __align__(1024) uint8_t smem_bytes[];
    Chunk256Smem& sm = *reinterpret_cast<Chunk256Smem*>(smem_bytes + ((1024u - (smem_u32(smem_bytes) & 1023u)) & 1023u));

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool wkv6_trace_on = WKV6_TRACE_ON(p, tid);
    const int BH = p.B * p.H;
    const int ds = blockIdx.x / BH;
    const int bh = blockIdx.x - ds * BH;
    const int b = bh / p.H, h = bh - b * p.H;
    const DirArgs& D = p.d[ds];
    const bool rev = D.reverse != 0;
    const int T = p.T;
    const int nchunks = (T + CL - 1) / CL;

    // Ktil variant tiles, compact: virtual base (row 0) of variant I inside sm.kt
    const int vb[5] = {rev ? 20 * 1024 : 0, rev ? 18 * 1024 : 2 * 1024, rev ? 14 * 1024 : 6 * 1024, rev ? 0 : 12 * 1024,
                       rev ? 8 * 1024 : 20 * 1024};

    if (tid == 0) {
        mbar_init(&sm.bar_rkw, 1);
        mbar_init(&sm.bar_v[0], 1);
        mbar_init(&sm.bar_v[1], 1);
        mbar_init(&sm.bar_p, 4);      // the four G1 groups are issued by four warps' lane 0 in parallel
        mbar_init(&sm.bar_s, 1);
        mbar_init(&sm.bar_y, 1);
        fence_barrier_init();
        tma_prefetch_desc(&D.mr); tma_prefetch_desc(&D.mk); tma_prefetch_desc(&D.mv); tma_prefetch_desc(&D.mw);
        if (!p.skip_y) tma_prefetch_desc(&D.my);
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&sm.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = tid; i < 28 * 1024 / 16; i += 256) reinterpret_cast<uint4*>(sm.kt)[i] = make_uint4(0, 0, 0, 0);

    // TMEM fragment coordinates (16x256b.x4): rows frow, frow+8; columns fcol + 8*rep + {0,1}
    const int qtr = warp & 3, half = warp >> 2;
    const int frow = 16 * qtr + (lane >> 2);
    const int fcol = 32 * half + 2 * (lane & 3);
    const uint32_t lane_field = (uint32_t)(qtr * 32) << 16;
    // row-major coordinates: sub-block I (direction order), channels c0..c0+3, rows 4*rg..4*rg+3 of the sub-block
    const int I = warp & 3, ch = warp >> 2;
    const int cq = lane & 7, rg = lane >> 3;
    const int c0 = 32 * ch + 4 * cq;
    const int coff = (c0 & 7) * 2, cchunk = c0 >> 3;
    const int vb_own = I == 0 ? vb[0] : I == 1 ? vb[1] : I == 2 ? vb[2] : vb[3];

    float u4[4];
    ld4bf(reinterpret_cast<const uint8_t*>(D.u) + (h * N + c0) * 2, u4);

    // state S^T[j][i]: j = frow + 8*hh, i = fcol + 8*rep + e  ->  st[4*rep + 2*hh + e]
    float st[16];
    {
        const float* s0 = D.s0 ? D.s0 + (size_t)bh * N * N : nullptr;
#pragma unroll
        for (int rep = 0; rep < 4; rep++)
#pragma unroll
            for (int hh = 0; hh < 2; hh++)
#pragma unroll
                for (int e = 0; e < 2; e++)
                    st[4 * rep + 2 * hh + e] = s0 ? s0[(fcol + 8 * rep + e) * N + frow + 8 * hh] : 0.f;
    }
    auto store_sp = [&]() {
#pragma unroll
        for (int rep = 0; rep < 4; rep++)
#pragma unroll
            for (int hh = 0; hh < 2; hh++) {
                const int col = fcol + 8 * rep;
                *reinterpret_cast<uint32_t*>(sm.sp + sw_chunk(frow + 8 * hh, col >> 3) + (col & 7) * 2) =
                    pack_bf16(st[4 * rep + 2 * hh], st[4 * rep + 2 * hh + 1]);
            }
    };
    store_sp();
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = sm.tmem_base;
    const uint32_t tP = tbase, tY = tbase + 64, tD = tbase + 128;

    auto issue_rkw = [&](int c) {   // thread 0 only
        const int q = rev ? nchunks - 1 - c : c;
        mbar_arrive_expect_tx(&sm.bar_rkw, 3u * CL * 128);
        tma_load_3d(sm.raw_r, &D.mr, &sm.bar_rkw, h * N, q * CL, b);
        tma_load_3d(sm.raw_k, &D.mk, &sm.bar_rkw, h * N, q * CL, b);
        tma_load_3d(sm.raw_w, &D.mw, &sm.bar_rkw, h * N, q * CL, b);
    };
    auto issue_v = [&](int c) {
        const int q = rev ? nchunks - 1 - c : c;
        mbar_arrive_expect_tx(&sm.bar_v[c & 1], (uint32_t)CL * 128);
        tma_load_3d(sm.v[c & 1], &D.mv, &sm.bar_v[c & 1], h * N, q * CL, b);
    };
    if (tid == 0 && nchunks > 0) { issue_rkw(0); issue_v(0); }

    // UMMA smem descriptors built once; inside the loop: base + (byte offset >> 4)
    const uint64_t dk_rst = desc_k(smem_u32(sm.rst)), dk_kt = desc_k(smem_u32(sm.kt)), dk_roff = desc_k(smem_u32(sm.roff));
    const uint64_t dm_sp = desc_mn(smem_u32(sm.sp)), dm_kt = desc_mn(smem_u32(sm.kt)), dm_pt = desc_mn(smem_u32(sm.pt)),
                   dm_v0 = desc_mn(smem_u32(sm.v[0])), dm_v1 = desc_mn(smem_u32(sm.v[1]));
#define DOFF(bytes) ((uint64_t)((bytes) >> 4))
    constexpr uint32_t ID_G1 = umma_idesc(0, 0, 64, 16);
    constexpr uint32_t ID_G2 = umma_idesc(1, 1, 64, 64);
    constexpr uint32_t ID_G3 = umma_idesc(0, 1, 64, 64);

    for (int c = 0; c < nchunks; c++) {
        const int q = rev ? nchunks - 1 - c : c;
        const int t0 = q * CL;
        const uint32_t par = c & 1;
        const int c_ts = c;
        WKV6_TS(0, 0);
        if (tid == 0) tma_store_wait_read<0>();   // y staging tile (= pt) of the previous chunk has been read

        // ---------------- phase 1: decays, in-block cumulative sums, Rtil, Ktil(own block), bonus
        mbar_wait(&sm.bar_rkw, par);
        WKV6_TS(0, 1);
        float ro[4][4], ko[4][4];
        int rows[4];
        {
            float l2[4][4], bsum[4];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int pos = 16 * I + 4 * rg + r;
                const int row = rev ? 63 - pos : pos;
                rows[r] = row;
                const int off = sw_chunk(row, cchunk) + coff;
                float w4[4];
                ld4bf(sm.raw_r + off, ro[r]);
                ld4bf(sm.raw_k + off, ko[r]);
                ld4bf(sm.raw_w + off, w4);
                const bool valid = t0 + row < T;
                float bs = 0.f;
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    bs = fmaf(ro[r][e] * u4[e], ko[r][e], bs);
                    const float l = fmaxf(-ex2(w4[e] * kLog2e) * kLog2e, -kClampL2);
                    l2[r][e] = valid ? l : 0.f;
                }
                bsum[r] = bs;
            }
#pragma unroll
            for (int e = 0; e < 4; e++) {
                l2[1][e] += l2[0][e];
                l2[2][e] += l2[1][e];
                l2[3][e] += l2[2][e];
            }
            float excl[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                float x = l2[3][e];
                float y = __shfl_up_sync(0xffffffffu, x, 8);
                if (rg >= 1) x += y;
                y = __shfl_up_sync(0xffffffffu, x, 16);
                if (rg >= 2) x += y;
                excl[e] = x - l2[3][e];
                if (rg == 3) sm.blocktot[I][c0 + e] = x;
            }
#pragma unroll
            for (int r = 0; r < 4; r++) {
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const float aex = excl[e] + (r ? l2[r - 1][e] : 0.f);   // A_p  - a_I   (<= 0)
                    const float ain = excl[e] + l2[r][e];                   // A+_p - a_I   (<= 0)
                    ro[r][e] *= ex2(aex);
                    ko[r][e] *= ex2(-ain);
                }
                const int off = sw_chunk(rows[r], cchunk) + coff;
                st4bf(sm.roff + off, ro[r]);
                st4bf(sm.kt + vb_own + off, ko[r]);
                float bs = bsum[r];
                bs += __shfl_xor_sync(0xffffffffu, bs, 1);
                bs += __shfl_xor_sync(0xffffffffu, bs, 2);
                bs += __shfl_xor_sync(0xffffffffu, bs, 4);
                if (cq == 0) sm.bonus[ch][rows[r]] = bs;
            }
        }
        WKV6_TS(0, 2);
        __syncthreads();   // B1: blocktot / bonus visible, raw r/k/w consumed
        WKV6_TS(0, 3);
        if (tid == 0 && c + 1 < nchunks) { issue_rkw(c + 1); issue_v(c + 1); }

        // ---------------- phase 2: cross-block factors -> Rst, Ktil(I' > I) incl. Ktil(4)
        {
            float fst[4], fk[4][4], tot[4][4];
#pragma unroll
            for (int m = 0; m < 4; m++) {
                const float4 t = *reinterpret_cast<const float4*>(&sm.blocktot[m][c0]);
                tot[m][0] = t.x; tot[m][1] = t.y; tot[m][2] = t.z; tot[m][3] = t.w;
            }
#pragma unroll
            for (int e = 0; e < 4; e++) {
                float s = 0.f, acc = 0.f;
#pragma unroll
                for (int m = 0; m < 4; m++) {
                    const float t = tot[m][e];
                    if (m < I) s += t;
                    if (m >= I) acc += t;
                    fk[m][e] = ex2(acc);        // 2^(a_{m+1} - a_I), used for m >= I
                }
                fst[e] = ex2(s);                // 2^(a_I)
            }
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int off = sw_chunk(rows[r], cchunk) + coff;
                float x4[4];
#pragma unroll
                for (int e = 0; e < 4; e++) x4[e] = ro[r][e] * fst[e];
                st4bf(sm.rst + off, x4);
#pragma unroll
                for (int m = 0; m < 4; m++) {
                    if (m >= I) {
#pragma unroll
                        for (int e = 0; e < 4; e++) x4[e] = ko[r][e] * fk[m][e];
                        st4bf(sm.kt + vb[m + 1] + off, x4);
                    }
                }
            }
        }
        fence_proxy_async();
        tc_fence_before();
        WKV6_TS(0, 4);
        __syncthreads();   // B2: operand tiles complete
        WKV6_TS(0, 5);
        if (tid == 0 && D.ckpt) {   // S at the start of this chunk (bf16, [j][i]) -> checkpoint c of this sequence
            tma_store_2d(&D.mc, sm.sp, 0, (bh * nchunks + c) * N);
            tma_store_commit();
        }
        if (lane == 0 && warp < 5) {
            tc_fence_after();
            if (warp < 4) {          // G1 group `warp`
                const int i = warp;
                const int rowlo = rev ? 48 - 16 * i : 16 * i;
                const int vbi = i == 0 ? vb[0] : i == 1 ? vb[1] : i == 2 ? vb[2] : vb[3];
#pragma unroll
                for (int ks = 0; ks < 4; ks++)
                    umma(tP + rowlo, dk_kt + DOFF(vbi + ks * 32), dk_roff + DOFF(rowlo * 128 + ks * 32), ID_G1, ks > 0);
                umma_commit(&sm.bar_p);
            } else {                 // G3, G4 (and later G2) from one thread: G2 accumulates on top of G3
                const uint64_t dm_v = (c & 1) ? dm_v1 : dm_v0;
                mbar_wait(&sm.bar_v[c & 1], (c >> 1) & 1);
#pragma unroll
                for (int ks = 0; ks < 4; ks++) umma(tY, dk_rst + DOFF(ks * 32), dm_sp + DOFF(ks * 2048), ID_G3, ks > 0);
#pragma unroll
                for (int ks = 0; ks < 4; ks++) umma(tD, dm_kt + DOFF(vb[4] + ks * 2048), dm_v + DOFF(ks * 2048), ID_G2, ks > 0);
                umma_commit(&sm.bar_s);
            }
        }
        __syncwarp();
        WKV6_TS(0, 6);

        // ---------------- P^T readout: causal mask + bonus diagonal -> bf16 tile (MN-major A of G2)
        mbar_wait(&sm.bar_p, par);
        WKV6_TS(0, 7);
        tc_fence_after();
        {
            float x[16];
            tmem_ld_16x32(tP + lane_field + 32 * half, x);
#pragma unroll
            for (int rep = 0; rep < 4; rep++)
#pragma unroll
                for (int hh = 0; hh < 2; hh++) {
                    const int qq = frow + 8 * hh;
                    float y[2];
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        const int pp = fcol + 8 * rep + e;
                        const bool before = rev ? (qq > pp) : (qq < pp);
                        y[e] = before ? x[4 * rep + 2 * hh + e] : (qq == pp ? sm.bonus[0][pp] + sm.bonus[1][pp] : 0.f);
                    }
                    const int pc = fcol + 8 * rep;
                    *reinterpret_cast<uint32_t*>(sm.pt + sw_chunk(qq, pc >> 3) + (pc & 7) * 2) = pack_bf16(y[0], y[1]);
                }
        }
        fence_proxy_async();
        tc_fence_before();
        if (tid == 0 && D.ckpt) tma_store_wait_read<0>();   // sp is rewritten after B3
        WKV6_TS(0, 8);
        __syncthreads();   // B3: P^T tile complete
        WKV6_TS(0, 9);
        if (lane == 0 && warp == 4) {
            tc_fence_after();
            const uint64_t dm_v = (c & 1) ? dm_v1 : dm_v0;
#pragma unroll
            for (int ks = 0; ks < 4; ks++) umma(tY, dm_pt + DOFF(ks * 2048), dm_v + DOFF(ks * 2048), ID_G2, 1);
            umma_commit(&sm.bar_y);
        }
        __syncwarp();

        // ---------------- state update (overlaps G2):  S^T[j][i] = S^T[j][i] 2^(A_L[j]) + dS^T[j][i]
        mbar_wait(&sm.bar_s, par);
        WKV6_TS(0, 10);
        tc_fence_after();
        {
            float dv[16];
            tmem_ld_16x32(tD + lane_field + 32 * half, dv);
            const float f0 = ex2(sm.blocktot[0][frow] + sm.blocktot[1][frow] + sm.blocktot[2][frow] + sm.blocktot[3][frow]);
            const float f1 = ex2(sm.blocktot[0][frow + 8] + sm.blocktot[1][frow + 8] + sm.blocktot[2][frow + 8] + sm.blocktot[3][frow + 8]);
#pragma unroll
            for (int rep = 0; rep < 4; rep++) {
                st[4 * rep + 0] = fmaf(st[4 * rep + 0], f0, dv[4 * rep + 0]);
                st[4 * rep + 1] = fmaf(st[4 * rep + 1], f0, dv[4 * rep + 1]);
                st[4 * rep + 2] = fmaf(st[4 * rep + 2], f1, dv[4 * rep + 2]);
                st[4 * rep + 3] = fmaf(st[4 * rep + 3], f1, dv[4 * rep + 3]);
            }
            store_sp();
        }

        // ---------------- Y readout -> bf16 staging tile -> TMA store
        WKV6_TS(0, 11);
        mbar_wait(&sm.bar_y, par);
        WKV6_TS(0, 12);
        tc_fence_after();
        if (!p.skip_y) {
            float yv[16];
            tmem_ld_16x32(tY + lane_field + 32 * half, yv);
#pragma unroll
            for (int rep = 0; rep < 4; rep++)
#pragma unroll
                for (int hh = 0; hh < 2; hh++) {
                    const int col = fcol + 8 * rep;
                    *reinterpret_cast<uint32_t*>(sm.pt + sw_chunk(frow + 8 * hh, col >> 3) + (col & 7) * 2) =
                        pack_bf16(yv[4 * rep + 2 * hh], yv[4 * rep + 2 * hh + 1]);
                }
        }
        fence_proxy_async();
        tc_fence_before();
        WKV6_TS(0, 13);
        __syncthreads();   // B4: y tile and sp complete; all TMEM reads of this chunk done
        WKV6_TS(0, 14);
        if (tid == 0 && !p.skip_y) {
            tma_store_3d(&D.my, sm.pt, h * N, t0, b);
            tma_store_commit();
        }
    }

    if (tid == 0) tma_store_wait<0>();
    if (D.sT) {
        float* sT = D.sT + (size_t)bh * N * N;
#pragma unroll
        for (int rep = 0; rep < 4; rep++)
#pragma unroll
            for (int hh = 0; hh < 2; hh++)
#pragma unroll
                for (int e = 0; e < 2; e++) sT[(fcol + 8 * rep + e) * N + frow + 8 * hh] = st[4 * rep + 2 * hh + e];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tbase) : "memory");
}

int launch_fwd_chunk_lean(const Params& p, cudaStream_t st);   // wkv6_chunk_lean.cu
int launch_fwd_chunk_ws(const Params& p, cudaStream_t st);     // wkv6_chunk_ws.cu

int launch_fwd_chunk(const Params& p, cudaStream_t st) {
    static const bool use_v1 = getenv("WKV6_FWD_CHUNK_V1") != nullptr;   // A/B switch: the first 256-thread kernel
    // warp-specialised variant (dedicated issuer warp): faster for a CTA that has its SM to itself (31 vs 36 us on the
    // bench workload), slower when two CTAs share an SM (long tail, 55 us) -- opt-in until that is understood
    static const bool use_ws = getenv("WKV6_FWD_CHUNK_WS") != nullptr;
    if (use_ws) return launch_fwd_chunk_ws(p, st);
    if (!use_v1) return launch_fwd_chunk_lean(p, st);
    const size_t smem = sizeof(Chunk256Smem) + 1024;
    cudaError_t e = cudaFuncSetAttribute(wkv6_fwd_chunk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    Params q = p;
    q.dbg = debug_trace_buffer();
    wkv6_fwd_chunk_kernel<<<dim3(q.ndir * q.B * q.H), dim3(256), smem, st>>>(q);
    count_launch(1);
    return (int)cudaGetLastError();
}

}  // namespace wkv6
