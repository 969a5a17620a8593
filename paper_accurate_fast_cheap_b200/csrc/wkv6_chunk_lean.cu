// wkv6_chunk_lean.cu -- chunked (dense-contraction) WKV6 forward for bf16 I/O on tcgen05 / TMEM / TMA (sm_100a).
//
// Same function as the reference's kernel_forward (wenet/rwkv_v6/cuda/wkv6_cuda.cu:9-63; with an initial state:
// wkv6state_cuda.cu:7-65), restated per chunk of L = 64 steps so that the O(L*N*N) work runs on
// the 5th-gen tensor cores:
//     l_t = -exp(w_t) (log2 units, clamped at -5*log2e), A_t = exclusive cumsum, A+_t = inclusive cumsum
//     P[p][q]  = sum_j r_p[j] k_q[j] 2^(A_p[j] - A+_q[j])   (q before p),  P[p][p] = sum_j r_p u k_p
//     Y        = P V  +  (r_p 2^A_p) S^T                     S' = S diag(2^A_L) + V^T (k_q 2^(A_L - A+_q))
// Overflow-free factorisation: the chunk is cut into four 16-step sub-blocks; every factor is taken relative to
// the START of the row's sub-block I:  Rtil_p = r_p 2^(A_p - a_I) (<= |r|),  Ktil(I)_q = k_q 2^(a_I - A+_q)
// (<= |k| for earlier sub-blocks, <= 2^(16*7.2) inside the diagonal sub-block: finite in bf16/fp32).
//
// One CTA (256 threads) per (direction, batch, head) [x time segment]; two CTAs per SM overlap each other's latency.
//   G1 (x4): P^T[:, block I] = Ktil(I) [64 q x 64 j, K-major A]  x  Rtil[block I] [16 p x 64 j, K-major B]   M64 N16 K64
//   G3     : Y   = Rst [64 p x 64 j, K-major A] x S' [64 j x 64 i, MN-major B]                              M64 N64 K64
//   G2     : Y  += P^T tile [64 q x 64 p, MN-major A] x V [64 q x 64 i, MN-major B, straight from TMA]
//   G4     : dS^T[j][i] = Ktil(4) [64 q x 64 j, MN-major A] x V [MN-major B]
// All tiles are [64][64] bf16 with 128-byte rows and the 128B swizzle, so the same bytes serve as K-major or
// MN-major operands (conventions verified on hardware by tests/probes/umma_probe.cu).  Accumulators live in TMEM
// (P^T, Y, dS^T: 3 x 64 columns) and are read with tcgen05.ld.16x256b; the fp32 state S^T[j][i] stays in
// registers in that fragment layout for the whole sequence.  The Ktil variants are stored compactly (variant I holds
// the rows of sub-blocks <= I only: 28 KB instead of 40 KB, bases in ktbase()).
//
// How the kernel got here from a straightforward first version (~1150 instructions per thread and chunk, 52 % issue
// utilisation, 4 CTA-wide barriers per chunk; history in DESIGN.md):
//   * operand tiles are produced with packed bf16x2 arithmetic (HMUL2.BF16 on the raw TMA bits; the decay factors are
//     computed in fp32, rounded once to bf16x2): no bf16->fp32 unpacking of r/k, half the multiplies, no F2FP per tile;
//   * the causal mask is resolved per 16x16 block (warp-uniform: keep / zero / diagonal), only diagonal blocks pay
//     per-element selects;
//   * TMEM fragments go to the swizzled operand tiles with stmatrix.x4 (2 per 16x32 fragment instead of 8 STS.32);
//   * the chunk loop is software-pipelined: the tile building of chunk c+1 (which does not depend on the state) runs
//     between the G2 issue and the Y readout of chunk c, so G2 / G3 / G4 latency is covered and only G1 -> P^T readout
//     is exposed; 3 CTA-wide barriers per chunk.
// Log-decays are kept with the opposite sign here (m = -l >= 0) so that the clamp is one FMNMX.
#include "wkv6_fwd_tiles.cuh"

namespace wkv6 {

struct LeanSmem {
    alignas(1024) uint8_t raw_r[CL * 128];   // TMA landing tiles (bf16, swizzled)
    alignas(1024) uint8_t raw_k[CL * 128];
    alignas(1024) uint8_t raw_w[CL * 128];
    alignas(1024) uint8_t v[2][CL * 128];    // V: MMA operand straight from TMA
    alignas(1024) uint8_t roff[CL * 128];    // r 2^(A_p - a_I)
    alignas(1024) uint8_t rst[CL * 128];     // r 2^(A_p)
    alignas(1024) uint8_t kt[28 * 1024];     // Ktil(0..4), compact (see wkv6_chunk.cu)
    alignas(1024) uint8_t pt[CL * 128];      // masked P^T [q][p]
    alignas(1024) uint8_t ys[CL * 128];      // y staging tile [p][i]
    alignas(1024) uint8_t sp[N * 128];       // bf16 state operand S^T [j][i]; source of the checkpoint store
    float blocktot[2][4][N];                 // -log2 decay sums of the four sub-blocks per channel, double-buffered
    float bonus[2][CL];                      // sum_j r u k, one partial per channel half
    alignas(8) uint64_t bar_rkw, bar_v[2], bar_p, bar_s, bar_y;
    uint32_t tmem_base;
};

// REV is a RUNTIME flag on purpose: with the direction as a template parameter the two instantiations of co-resident CTAs
// thrashed the instruction cache (ncu: "no instruction" top stall; 50.8 us mixed vs 45.6 us single-direction launches)
// The CTA processes the chunks [cbeg, cbeg + nchunks) of its sequence (processing order), starting from the state *s0p
// (null: zero) and leaving the final state in *sTp (unless null); decay_out (unless null) receives the product of all
// per-step decays of that range per key channel.  Unsegmented: cbeg = 0, nchunks = all.
// kMode is compile-time, so that the normal instantiation carries none of the other modes' code (and vice versa):
//   kNormal     the whole range from *s0p, outputs (and checkpoints)
//   kStateOnly  pass 1 of the two-pass time-segmented forward: no outputs, no checkpoints, S_0 = 0, decay product out
//   kLocalOut   the one-pass time-segmented forward: outputs computed from S_0 = 0 (wkv6_fwd_corr_kernel adds the carried
//               state's share afterwards), local end state and decay product out, and per chunk the decay product from the
//               start of the segment to the start of the chunk (prefix_out[chunk][64])
constexpr int kNormal = 0, kStateOnly = 1, kLocalOut = 2;
template <int kMode>
__device__ __forceinline__ void lean_body(const Params& p, LeanSmem& sm, const DirArgs& D, const int b, const int h, const int bh,
                                          const bool REV, const int cbeg, const int nchunks, const float* s0p_, float* sTp,
                                          float* decay_out_, const bool skip_y_, const bool ckpt_on_, float* prefix_out_ = nullptr) {
    const float* const s0p = kMode == kNormal ? s0p_ : nullptr;
    float* const decay_out = kMode == kNormal ? nullptr : decay_out_;
    const bool skip_y = kMode == kStateOnly ? true : kMode == kLocalOut ? false : skip_y_;
    const bool ckpt_on = kMode == kNormal ? ckpt_on_ : false;
    float* const prefix_out = kMode == kLocalOut ? prefix_out_ : nullptr;
    const int tid = threadIdx.x, lane = tid & 31;
    const bool wkv6_trace_on = WKV6_TRACE_ON(p, tid);
    // warp index through a lane-0 broadcast: the compiler then keeps it (and everything derived from it: MMA operands,
    // descriptors, role branches) in uniform registers; a plain tid >> 5 makes every UTCHMMA a waterfall loop
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int T = p.T;
    const int ntot = (T + CL - 1) / CL;
    auto tileq = [&](int c) { return REV ? ntot - 1 - (cbeg + c) : cbeg + c; };   // tile index in memory of local chunk c
    // one elected lane of warp 5 issues every TMA load / store; one elected lane of warps 0..4 issues the MMAs
#define TMA_THR (warp == 5 && elect_one())

    if (tid == 0) {
        mbar_init(&sm.bar_p, 4);      // the four G1 groups are issued by four warps' lane 0 in parallel
        mbar_init(&sm.bar_s, 1);
        mbar_init(&sm.bar_y, 1);
        fence_barrier_init();
    }
    if (TMA_THR) {
        // the TMA lane owns the load barriers: it initialises them and puts the first chunk's loads in flight before the
        // rest of the prologue (TMEM allocation, shared-memory clears, state load) instead of after it
        mbar_init(&sm.bar_rkw, 1);
        mbar_init(&sm.bar_v[0], 1);
        mbar_init(&sm.bar_v[1], 1);
        fence_barrier_init();
        fence_proxy_async();
        tma_prefetch_desc(&D.mr); tma_prefetch_desc(&D.mk); tma_prefetch_desc(&D.mv); tma_prefetch_desc(&D.mw);
        if (!skip_y) tma_prefetch_desc(&D.my);
        pdl_wait();
        if (nchunks > 0) {
            const int q = tileq(0);
            mbar_arrive_expect_tx(&sm.bar_rkw, 3u * CL * 128);
            tma_load_3d(sm.raw_r, &D.mr, &sm.bar_rkw, h * N, q * CL, b);
            tma_load_3d(sm.raw_k, &D.mk, &sm.bar_rkw, h * N, q * CL, b);
            tma_load_3d(sm.raw_w, &D.mw, &sm.bar_rkw, h * N, q * CL, b);
            mbar_arrive_expect_tx(&sm.bar_v[0], (uint32_t)CL * 128);
            tma_load_3d(sm.v[0], &D.mv, &sm.bar_v[0], h * N, q * CL, b);
        }
    }
    __syncwarp();
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&sm.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // rows of a Ktil variant that are never written must still hold finite bf16 values
    for (int i = tid; i < 28 * 1024 / 16; i += 256) reinterpret_cast<uint4*>(sm.kt)[i] = make_uint4(0, 0, 0, 0);

    // TMEM fragment coordinates (16x256b.x4): rows frow, frow+8; columns fcol + 8*rep + {0,1}
    const int qtr = warp & 3, half = warp >> 2;
    const int frow = 16 * qtr + (lane >> 2);
    const int fcol = 32 * half + 2 * (lane & 3);
    const uint32_t lane_field = (uint32_t)(qtr * 32) << 16;
    // stmatrix row addresses of this lane for the two x4 stores of a 16x32 fragment: matrix mi = lane>>3 of store s is
    // (rep = 2s + (mi>>1), hh = mi&1), i.e. registers pk[4s .. 4s+3] in order
    uint32_t stoff[2];
    {
        const int mi = lane >> 3, rr = lane & 7;
        const int row = 16 * qtr + 8 * (mi & 1) + rr;
        stoff[0] = (uint32_t)sw_chunk(row, 4 * half + (mi >> 1));
        stoff[1] = (uint32_t)sw_chunk(row, 4 * half + 2 + (mi >> 1));
    }
    // row-major coordinates: sub-block I (in processing order), channels c0..c0+3, rows 4*rg..4*rg+3 of the sub-block
    const int I = warp & 3, ch = warp >> 2;
    const int cq = lane & 7, rg = lane >> 3;
    const int c0 = 32 * ch + 4 * cq;
    int offs[4];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const int pos = 16 * I + 4 * rg + r;
        const int row = REV ? 63 - pos : pos;
        offs[r] = sw_chunk(row, c0 >> 3) + (c0 & 7) * 2;
    }
    const int vb_own = I == 0 ? ktbase(REV, 0) : I == 1 ? ktbase(REV, 1) : I == 2 ? ktbase(REV, 2) : ktbase(REV, 3);
    pdl_wait();   // first global access below: the previous kernel of the stream has completed and its writes are visible
    const uint2 u2 = *reinterpret_cast<const uint2*>(reinterpret_cast<const uint8_t*>(D.u) + (h * N + c0) * 2);

    const uint32_t rawr_u32 = smem_u32(sm.raw_r), rawk_u32 = smem_u32(sm.raw_k), raww_u32 = smem_u32(sm.raw_w);
    const uint32_t roff_u32 = smem_u32(sm.roff), rst_u32 = smem_u32(sm.rst), kt_u32 = smem_u32(sm.kt);
    const uint32_t pt_u32 = smem_u32(sm.pt), ys_u32 = smem_u32(sm.ys), sp_u32 = smem_u32(sm.sp);

    // state S^T[j][i]: j = frow + 8*hh, i = fcol + 8*rep + e  ->  st[4*rep + 2*hh + e]
    float dprod0 = 1.f, dprod1 = 1.f;   // product of the chunk decays 2^(A_L[j]), j = frow, frow + 8 (segmented pass 1)
    float st[16];
    {
        const float* s0 = s0p;
#pragma unroll
        for (int rep = 0; rep < 4; rep++)
#pragma unroll
            for (int hh = 0; hh < 2; hh++)
#pragma unroll
                for (int e = 0; e < 2; e++)
                    st[4 * rep + 2 * hh + e] = s0 ? s0[(fcol + 8 * rep + e) * N + frow + 8 * hh] : 0.f;
    }
    auto store_frag = [&](uint32_t tile_u32, const float (&x)[16]) {   // fp32 fragment -> bf16 swizzled tile
        uint32_t pk[8];
#pragma unroll
        for (int i = 0; i < 8; i++) pk[i] = pack_bf16(x[2 * i], x[2 * i + 1]);
        stsm_x4(tile_u32 + stoff[0], pk[0], pk[1], pk[2], pk[3]);
        stsm_x4(tile_u32 + stoff[1], pk[4], pk[5], pk[6], pk[7]);
    };
    store_frag(sp_u32, st);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = __shfl_sync(0xffffffffu, sm.tmem_base, 0);
    const uint32_t tP = tbase, tY = tbase + 64, tD = tbase + 128;

    auto issue_rkw = [&](int c) {   // TMA_THR only
        const int q = tileq(c);
        mbar_arrive_expect_tx(&sm.bar_rkw, 3u * CL * 128);
        tma_load_3d(sm.raw_r, &D.mr, &sm.bar_rkw, h * N, q * CL, b);
        tma_load_3d(sm.raw_k, &D.mk, &sm.bar_rkw, h * N, q * CL, b);
        tma_load_3d(sm.raw_w, &D.mw, &sm.bar_rkw, h * N, q * CL, b);
    };
    auto issue_v = [&](int c) {
        const int q = tileq(c);
        mbar_arrive_expect_tx(&sm.bar_v[c & 1], (uint32_t)CL * 128);
        tma_load_3d(sm.v[c & 1], &D.mv, &sm.bar_v[c & 1], h * N, q * CL, b);
    };

    // UMMA smem descriptors built once; inside the loop: base + (byte offset >> 4)
    const uint64_t dk_rst = desc_k(rst_u32), dk_kt = desc_k(kt_u32), dk_roff = desc_k(roff_u32);
    const uint64_t dm_sp = desc_mn(sp_u32), dm_kt = desc_mn(kt_u32), dm_pt = desc_mn(pt_u32),
                   dm_v0 = desc_mn(smem_u32(sm.v[0])), dm_v1 = desc_mn(smem_u32(sm.v[1]));
#define DOFF(bytes) ((uint64_t)((bytes) >> 4))
    constexpr uint32_t ID_G1 = umma_idesc(0, 0, 64, 16);
    constexpr uint32_t ID_G2 = umma_idesc(1, 1, 64, 64);
    constexpr uint32_t ID_G3 = umma_idesc(0, 1, 64, 64);

    uint32_t ro2[4][2], ko2[4][2];   // Rtil / Ktil(own) of this thread's 4 rows x 4 channels, packed bf16x2
    // state-only mode (pass 1 of a time-segmented launch): no y, so no P^T (G1), no G2 / G3, no bonus and no R-side tiles;
    // what is left per chunk is the decays, Ktil relative to the chunk end, dS (G4) and the state update
    const bool so = decay_out != nullptr && skip_y && !ckpt_on;

    // ---- phase 1 of chunk c: decays, in-block cumulative sums, Rtil, Ktil(own block), bonus
    auto phase1 = [&](const int c) {
        const int q = tileq(c);
        const int t0 = q * CL;
        mbar_wait(&sm.bar_rkw, c & 1);
        float m[4][4];
        uint32_t rr[4][2], kk[4][2];
#pragma unroll
        for (int r = 0; r < 4; r++) {
            uint2 ww;
            asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(rr[r][0]), "=r"(rr[r][1]) : "r"(rawr_u32 + offs[r]));
            asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(kk[r][0]), "=r"(kk[r][1]) : "r"(rawk_u32 + offs[r]));
            asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=r"(ww.x), "=r"(ww.y) : "r"(raww_u32 + offs[r]));
            float w4[4];
            up4bf(ww, w4);
#pragma unroll
            for (int e = 0; e < 4; e++) m[r][e] = fminf(ex2(fmaf(w4[e], kLog2e, kLog2Log2e)), kClampL2);
        }
        if (t0 + CL > T) {   // ragged last chunk (CTA-uniform): steps at or beyond T neither decay nor contribute
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int pos = 16 * I + 4 * rg + r;
                const int row = REV ? 63 - pos : pos;
                if (t0 + row >= T) {
#pragma unroll
                    for (int e = 0; e < 4; e++) m[r][e] = 0.f;
                }
            }
        }
        // bonus sum_j r u k over this thread's 4 channels (packed bf16), then a transposing reduction over the 8 lanes
        if (!so) {
        float bs[4];
#pragma unroll
        for (int r = 0; r < 4; r++) {
            uint32_t a = bmul2(bmul2(rr[r][0], u2.x), kk[r][0]);
            a = bfma2(bmul2(rr[r][1], u2.y), kk[r][1], a);
            bs[r] = __uint_as_float(a << 16) + __uint_as_float(a & 0xffff0000u);
        }
        {
            const bool b4 = (cq & 4) != 0, b2 = (cq & 2) != 0;
            float k0 = b4 ? bs[2] : bs[0], k1 = b4 ? bs[3] : bs[1];
            const float s0_ = b4 ? bs[0] : bs[2], s1_ = b4 ? bs[1] : bs[3];
            k0 += __shfl_xor_sync(0xffffffffu, s0_, 4);
            k1 += __shfl_xor_sync(0xffffffffu, s1_, 4);
            float kk_ = b2 ? k1 : k0;
            const float ss_ = b2 ? k0 : k1;
            kk_ += __shfl_xor_sync(0xffffffffu, ss_, 2);
            kk_ += __shfl_xor_sync(0xffffffffu, kk_, 1);
            if ((cq & 1) == 0) {
                const int pos = 16 * I + 4 * rg + (b4 ? 2 : 0) + (b2 ? 1 : 0);
                sm.bonus[ch][REV ? 63 - pos : pos] = kk_;
            }
        }
        }   // !so
        // per-step decays 2^(-m) of the thread's four steps: the factors of steps 1..3 follow from those of the run's ends by
        // products (10 ex2 per channel instead of 12: the MUFU pipe, 16 lanes per clock and SM, paces this phase)
        float d[4][4];
#pragma unroll
        for (int r = 0; r < 4; r++)
#pragma unroll
            for (int e = 0; e < 4; e++) d[r][e] = ex2(-m[r][e]);
#pragma unroll
        for (int e = 0; e < 4; e++) {   // inclusive sums over this thread's 4 steps
            m[1][e] += m[0][e];
            m[2][e] += m[1][e];
            m[3][e] += m[2][e];
        }
        float excl[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {   // scan over the 4 step-groups of the sub-block (lane bits 3, 4)
            float x = m[3][e];
            float y = __shfl_up_sync(0xffffffffu, x, 8);
            if (rg >= 1) x += y;
            y = __shfl_up_sync(0xffffffffu, x, 16);
            if (rg >= 2) x += y;
            excl[e] = x - m[3][e];
            m[3][e] = x;                // inclusive total up to and including this thread's last step
        }
        if (rg == 3) *reinterpret_cast<float4*>(&sm.blocktot[c & 1][I][c0]) = make_float4(m[3][0], m[3][1], m[3][2], m[3][3]);
        // fk[r] = 2^-(A+_p - a_I) >= 1 (inclusive), fr[r] = 2^(A_p - a_I) <= 1 (exclusive) of the thread's steps r = 0..3
        float fk[4][4], fr[4][4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            fk[3][e] = ex2(m[3][e]);
            fk[2][e] = fk[3][e] * d[3][e];
            fk[1][e] = fk[2][e] * d[2][e];
            fk[0][e] = fk[1][e] * d[1][e];
            fr[0][e] = ex2(-excl[e]);
            fr[1][e] = fr[0][e] * d[0][e];
            fr[2][e] = fr[1][e] * d[1][e];
            fr[3][e] = fr[2][e] * d[2][e];
        }
#pragma unroll
        for (int r = 0; r < 4; r++) {
            ko2[r][0] = bmul2(kk[r][0], pack_bf16(fk[r][0], fk[r][1]));
            ko2[r][1] = bmul2(kk[r][1], pack_bf16(fk[r][2], fk[r][3]));
            if (!so) {
                ro2[r][0] = bmul2(rr[r][0], pack_bf16(fr[r][0], fr[r][1]));
                ro2[r][1] = bmul2(rr[r][1], pack_bf16(fr[r][2], fr[r][3]));
                sts64(roff_u32 + offs[r], ro2[r][0], ro2[r][1]);
                sts64(kt_u32 + vb_own + offs[r], ko2[r][0], ko2[r][1]);   // Ktil(own): operand of G1 only
            }
        }
    };
    auto phase2 = [&](const int c) {
        const float* bt = &sm.blocktot[c & 1][0][0];
        switch (I) {   // warp-uniform
            case 0: lean_phase2<0>(REV, rst_u32, kt_u32, bt, c0, offs, ro2, ko2, so); break;
            case 1: lean_phase2<1>(REV, rst_u32, kt_u32, bt, c0, offs, ro2, ko2, so); break;
            case 2: lean_phase2<2>(REV, rst_u32, kt_u32, bt, c0, offs, ro2, ko2, so); break;
            default: lean_phase2<3>(REV, rst_u32, kt_u32, bt, c0, offs, ro2, ko2, so); break;
        }
    };
    // ---- MMAs that do not need P^T: G1 (four groups, four issuing lanes), G3 + G4 (one lane); checkpoint store
    auto issue_front = [&](const int c) {
        if (warp < 5 && elect_one()) {
            tc_fence_after();
            if (warp < 4) {
                if (so) goto front_done;
                const int rowlo = REV ? 48 - 16 * warp : 16 * warp;
                const int vbi = warp == 0 ? ktbase(REV, 0) : warp == 1 ? ktbase(REV, 1) : warp == 2 ? ktbase(REV, 2) : ktbase(REV, 3);
#pragma unroll
                for (int ks = 0; ks < 4; ks++)
                    umma(tP + rowlo, dk_kt + DOFF(vbi + ks * 32), dk_roff + DOFF(rowlo * 128 + ks * 32), ID_G1, ks > 0);
                umma_commit(&sm.bar_p);
            } else {
                const uint64_t dm_v = (c & 1) ? dm_v1 : dm_v0;
                mbar_wait(&sm.bar_v[c & 1], (c >> 1) & 1);
                if (!so) {
#pragma unroll
                    for (int ks = 0; ks < 4; ks++) umma(tY, dk_rst + DOFF(ks * 32), dm_sp + DOFF(ks * 2048), ID_G3, ks > 0);
                }
#pragma unroll
                for (int ks = 0; ks < 4; ks++)
                    umma(tD, dm_kt + DOFF(ktbase(REV, 4) + ks * 2048), dm_v + DOFF(ks * 2048), ID_G2, ks > 0);
                umma_commit(&sm.bar_s);
            }
        front_done:;
        }
        if (TMA_THR && ckpt_on) {   // S at the start of chunk c (bf16, [j][i]) -> checkpoint c of this sequence
            tma_store_2d(&D.mc, sm.sp, 0, (bh * ntot + cbeg + c) * N);
            tma_store_commit();
        }
        __syncwarp();
    };

    if (nchunks > 0) {
        phase1(0);                    // the loads of chunk 0 were issued at the top of the prologue
        __syncthreads();
        if (TMA_THR && nchunks > 1) { issue_rkw(1); issue_v(1); }
        phase2(0);
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        issue_front(0);
    }

    for (int c = 0; c < nchunks; c++) {
        const int q = tileq(c);
        const int t0 = q * CL;
        const uint32_t par = c & 1;
        const bool more = c + 1 < nchunks;
        const int c_ts = c;
        WKV6_TS(0, 0);
#if WKV6_TRACE_PHASES
        if (p.dbg && blockIdx.x == 0 && tid == 0 && c < 32) p.dbg[32 + c] = clock64();   // loop-top stamp of every chunk
#endif

        // ---------------- P^T readout: causal mask + bonus diagonal -> bf16 tile (MN-major A of G2)
        if (!so) {
        mbar_wait(&sm.bar_p, par);
        WKV6_TS(0, 1);
        tc_fence_after();
        {
            float x[16];
            tmem_ld_16x32(tP + lane_field + 32 * half, x);
            uint32_t pk[8];
#pragma unroll
            for (int cb = 0; cb < 2; cb++) {   // the two 16-column blocks of this warp's 16x32 fragment
                const int cls = REV ? qtr - (2 * half + cb) : (2 * half + cb) - qtr;   // warp-uniform
                if (cls > 0) {          // every q before every p: keep
#pragma unroll
                    for (int i = 0; i < 4; i++) pk[4 * cb + i] = pack_bf16(x[8 * cb + 2 * i], x[8 * cb + 2 * i + 1]);
                } else if (cls < 0) {   // no q before any p
#pragma unroll
                    for (int i = 0; i < 4; i++) pk[4 * cb + i] = 0u;
                } else {                // diagonal block: strict triangle + bonus on the diagonal
                    const int dbase = REV ? (lane >> 2) - 2 * (lane & 3) : 2 * (lane & 3) - (lane >> 2);
                    float bon[2];
#pragma unroll
                    for (int hh = 0; hh < 2; hh++) bon[hh] = sm.bonus[0][frow + 8 * hh] + sm.bonus[1][frow + 8 * hh];
#pragma unroll
                    for (int rl = 0; rl < 2; rl++)
#pragma unroll
                        for (int hh = 0; hh < 2; hh++) {
                            float y[2];
#pragma unroll
                            for (int e = 0; e < 2; e++) {
                                const int d = dbase + (REV ? 8 * hh - 8 * rl - e : 8 * rl + e - 8 * hh);   // > 0: q before p
                                const float xv = x[4 * (2 * cb + rl) + 2 * hh + e];
                                y[e] = d > 0 ? xv : (d == 0 ? bon[hh] : 0.f);
                            }
                            pk[2 * (2 * cb + rl) + hh] = pack_bf16(y[0], y[1]);
                        }
                }
            }
            stsm_x4(pt_u32 + stoff[0], pk[0], pk[1], pk[2], pk[3]);
            stsm_x4(pt_u32 + stoff[1], pk[4], pk[5], pk[6], pk[7]);
        }
        fence_proxy_async();
        tc_fence_before();
        if (TMA_THR) tma_store_wait_read<0>();   // checkpoint c has left sp, y of chunk c-1 has left ys
        WKV6_TS(0, 2);
        __syncthreads();   // B3: P^T tile complete
        WKV6_TS(0, 3);
        if (warp == 4 && elect_one()) {
            tc_fence_after();
            const uint64_t dm_v = (c & 1) ? dm_v1 : dm_v0;
#pragma unroll
            for (int ks = 0; ks < 4; ks++) umma(tY, dm_pt + DOFF(ks * 2048), dm_v + DOFF(ks * 2048), ID_G2, 1);
            umma_commit(&sm.bar_y);
        }
        __syncwarp();
        }   // !so

        // ---------------- state update (under G2):  S^T[j][i] = S^T[j][i] 2^(A_L[j]) + dS^T[j][i]
        mbar_wait(&sm.bar_s, par);
        WKV6_TS(0, 4);
        tc_fence_after();
        {
            float dv[16];
            tmem_ld_16x32(tD + lane_field + 32 * half, dv);
            const float* bt = &sm.blocktot[par][0][0];
            const float f0 = ex2(-((bt[frow] + bt[N + frow]) + (bt[2 * N + frow] + bt[3 * N + frow])));
            const float f1 = ex2(-((bt[frow + 8] + bt[N + frow + 8]) + (bt[2 * N + frow + 8] + bt[3 * N + frow + 8])));
            if (kMode == kLocalOut && half == 0 && (lane & 3) == 0) {   // decay from the start of the segment to the start of chunk c
                prefix_out[c * N + frow] = dprod0;
                prefix_out[c * N + frow + 8] = dprod1;
            }
            dprod0 *= f0;
            dprod1 *= f1;
#pragma unroll
            for (int rep = 0; rep < 4; rep++) {
                st[4 * rep + 0] = fmaf(st[4 * rep + 0], f0, dv[4 * rep + 0]);
                st[4 * rep + 1] = fmaf(st[4 * rep + 1], f0, dv[4 * rep + 1]);
                st[4 * rep + 2] = fmaf(st[4 * rep + 2], f1, dv[4 * rep + 2]);
                st[4 * rep + 3] = fmaf(st[4 * rep + 3], f1, dv[4 * rep + 3]);
            }
            if (!so) store_frag(sp_u32, st);
        }
        WKV6_TS(0, 5);

        // ---------------- tile building of the NEXT chunk (independent of the state; covers G2)
        if (more) {
            phase1(c + 1);
            WKV6_TS(0, 6);
            __syncthreads();   // B1: blocktot / bonus visible, raw r/k/w consumed
            WKV6_TS(0, 7);
            if (TMA_THR && c + 2 < nchunks) issue_rkw(c + 2);
            phase2(c + 1);
            WKV6_TS(0, 8);
        }

        // ---------------- Y readout -> bf16 staging tile -> TMA store
        if (!so) mbar_wait(&sm.bar_y, par);
        WKV6_TS(0, 9);
        tc_fence_after();
        if (!skip_y) {
            float yv[16];
            tmem_ld_16x32(tY + lane_field + 32 * half, yv);
            store_frag(ys_u32, yv);
        }
        fence_proxy_async();
        tc_fence_before();
        WKV6_TS(0, 10);
        __syncthreads();   // B2: y tile, sp and the operand tiles of chunk c+1 complete; TMEM reads of chunk c done
        WKV6_TS(0, 11);
        if (TMA_THR) {
            if (!skip_y) {
                tma_store_3d(&D.my, sm.ys, h * N, t0, b);
                tma_store_commit();
            }
            if (c + 2 < nchunks) issue_v(c + 2);   // v[c&1] is free: G2 and G4 of chunk c have completed
        }
        if (more) issue_front(c + 1);
        WKV6_TS(0, 12);
    }

    __syncwarp();
    if (TMA_THR) tma_store_wait<0>();
    if (decay_out && half == 0 && (lane & 3) == 0) {
        decay_out[frow] = dprod0;
        decay_out[frow + 8] = dprod1;
    }
    if (sTp) {
        float* sT = sTp;
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
#undef DOFF
#undef TMA_THR
}

// Two kernels share this top level: wkv6_fwd_chunk_lean_kernel carries ONLY the normal body (whole sequences, and pass 2 of
// the two-pass segmented scheme), wkv6_fwd_chunk_seg_kernel the two segment-local bodies (state-only pass 1; local outputs of
// the one-pass scheme).  With all three bodies in one kernel (9 500 SASS instructions) the normal launch was 0.9 us slower
// (41.2 vs 40.3 us) -- instruction-cache footprint of code it never runs.
template <bool kSegKernel>
__device__ __forceinline__ void lean_top(const Params& p) {
    extern __shared__ __align__(1024) uint8_t smem_bytes[];
    // offset arithmetic on the __shared__ array (not a uintptr_t round trip) keeps the address space visible to the
    // compiler: LDS/STS instead of generic LD/ST
    LeanSmem& sm = *reinterpret_cast<LeanSmem*>(smem_bytes + ((1024u - (smem_u32(smem_bytes) & 1023u)) & 1023u));
    const int BH = p.B * p.H;
    const int nseg = p.nseg > 0 ? p.nseg : 1;
    const int seq = blockIdx.x / nseg, seg = blockIdx.x - seq * nseg;   // seq = direction * B*H + bh
    const int ds = seq / BH;
    const int bh = seq - ds * BH;
    const int b = bh / p.H, h = bh - b * p.H;
    pdl_launch_dependents();   // the next launch of the stream may take the SM slots this grid leaves free (it blocks in pdl_wait())
    const DirArgs& D = p.d[ds];
    if (p.dbg && threadIdx.x == 0 && blockIdx.x < 1024) {
        unsigned long long t;
        uint32_t smid;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        p.dbg[128 + 4 * blockIdx.x] = t;
        p.dbg[129 + 4 * blockIdx.x] = smid;
    }
    const int ntot = (p.T + CL - 1) / CL;
    // Two instantiations of the body: the normal one and the state-only pass 1 (a single call site with run-time mode flags
    // was measured 3 us per launch SLOWER on the bench: 43.6 vs 40.5 us -- the pass-1 code and flags sit in the chunk loop).
    // Unsegmented: the whole sequence from S_0 to S_T.  Time-segmented: pass 1 (seg_mode 1) leaves every
    // segment's local state (from S = 0) and decay product in the scratch, a small scan kernel turns them into the
    // segment-start states, pass 2 (seg_mode 2) recomputes every segment from its start state.
    const bool segd = p.nseg > 1, pass1 = segd && p.seg_mode != 2;      // seg_mode 1: state-only pass 1; 3: local outputs
    const int cbeg = segd ? seg * p.seg_chunks : 0;
    const int nloc = segd ? min(p.seg_chunks, ntot - cbeg) : ntot;
    float* const slot = segd ? p.seg_state + ((size_t)seq * nseg + seg) * N * N : nullptr;
    const float* const s0p = pass1 ? nullptr : segd ? slot : (D.s0 ? D.s0 + (size_t)bh * N * N : nullptr);
    float* const sTp = pass1 ? slot : ((!segd || seg == nseg - 1) && D.sT) ? D.sT + (size_t)bh * N * N : nullptr;
    float* const dec = pass1 ? p.seg_decay + ((size_t)seq * nseg + seg) * N : nullptr;
    if (kSegKernel) {
        if (p.seg_mode == 3)
            lean_body<kLocalOut>(p, sm, D, b, h, bh, D.reverse != 0, cbeg, nloc, nullptr, sTp, dec, false, false,
                                 p.seg_prefix + ((size_t)seq * ntot + cbeg) * N);
        else
            lean_body<kStateOnly>(p, sm, D, b, h, bh, D.reverse != 0, cbeg, nloc, nullptr, sTp, dec, true, false);
    } else {
        lean_body<kNormal>(p, sm, D, b, h, bh, D.reverse != 0, cbeg, nloc, s0p, sTp, nullptr, p.skip_y != 0, D.ckpt != nullptr);
    }
    if (p.dbg && threadIdx.x == 0 && blockIdx.x < 1024) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.dbg[130 + 4 * blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(256, 2) wkv6_fwd_chunk_lean_kernel(const __grid_constant__ Params p) { lean_top<false>(p); }
__global__ void __launch_bounds__(256, 2) wkv6_fwd_chunk_seg_kernel(const __grid_constant__ Params p) { lean_top<true>(p); }

// Segment scan: S_start[0] = S_0 (or 0), S_start[s+1][i][j] = S_start[s][i][j] * D_s[j] + dS_s[i][j]; the slots that pass 1
// filled with dS_s are overwritten with S_start[s].  One thread per state element, grid = sequences x 16: every thread
// first loads its element of all segments (independent loads, all in flight), then runs the recurrence from registers.
constexpr int kMaxSeg = 40;   // pick_segments clamps to this (nseq >= 8 gives <= 37 anyway)
__global__ void __launch_bounds__(256) wkv6_seg_scan_kernel(const __grid_constant__ Params p) {
    const int BH = p.B * p.H;
    const int seq = blockIdx.x >> 4, ds = seq / BH, bh = seq - ds * BH;
    const int e = ((blockIdx.x & 15) << 8) + threadIdx.x;   // element i * N + j of the state
    const int j = e & (N - 1);
    const DirArgs& D = p.d[ds];
    float* st = p.seg_state + (size_t)seq * p.nseg * N * N + e;
    const float* dec = p.seg_decay + (size_t)seq * p.nseg * N + j;
    pdl_launch_dependents();   // programmatic dependent launch (wkv6_common.cuh): the correction kernel's prologue may start
    pdl_wait();                // pass 1 has completed and its writes are visible
    float dS[kMaxSeg], dk[kMaxSeg];
#pragma unroll
    for (int s = 0; s < kMaxSeg; s++)
        if (s < p.nseg) {
            dS[s] = st[(size_t)s * N * N];
            dk[s] = dec[s * N];
        }
    float cur = D.s0 ? D.s0[(size_t)bh * N * N + e] : 0.f;
#pragma unroll
    for (int s = 0; s < kMaxSeg; s++)
        if (s < p.nseg) {
            st[(size_t)s * N * N] = cur;
            cur = fmaf(cur, dk[s], dS[s]);
        }
    if (p.seg_mode == 3 && D.sT) D.sT[(size_t)bh * N * N + e] = cur;   // one-pass scheme: nobody else computes S_T
}

// One-pass time-segmented forward, second kernel: y of every chunk outside the first segment was computed from S = 0 at the
// start of its segment; the state carried into the segment contributes  y_p += sum_j r_p[j] 2^(A_p[j]) Dp[j] S_start[.][j],
// Dp = decay from the start of the segment to the start of the chunk (written per chunk by the first kernel), S_start from
// the segment scan.  One CTA per (sequence, segment, piece of the segment): the bf16 copy of S_start is built ONCE and the
// CTA walks its run of chunks with the r, w and y tiles of the next two chunks in flight (TMA, two stages); per chunk: the
// r 2^(A_p) Dp tile rebuilt from r and w (decays, in-block cumulative sums -- the R side of the main kernel's tile building),
// one M64 N64 K64 MMA group against S_start, y tile updated in shared memory and stored back.  66 KB, 64 TMEM columns: three
// CTAs per SM.  (The first version ran one CTA per chunk -- barrier init, TMEM allocation, a 16 KB state load and a serial
// load -> compute -> store chain for 24 KB of traffic each: 75 us per configs[2] block at one utterance per GPU, twice the
// time of its bytes.)
struct CorrSmem {
    alignas(1024) uint8_t raw_r[2][CL * 128];
    alignas(1024) uint8_t raw_w[2][CL * 128];
    alignas(1024) uint8_t y[2][CL * 128];
    alignas(1024) uint8_t rc[CL * 128];      // r 2^(A_p) Dp, K-major [p][j]
    alignas(1024) uint8_t sp[N * 128];       // S_start^T [j][i] bf16 (MN-major B operand, as in the main kernel)
    float blocktot[4][N];
    alignas(8) uint64_t bar_in[2], bar_m;
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(256) wkv6_fwd_corr_kernel(const __grid_constant__ Params p) {
    extern __shared__ __align__(1024) uint8_t smem_bytes[];
    CorrSmem& sm = *reinterpret_cast<CorrSmem*>(smem_bytes + ((1024u - (smem_u32(smem_bytes) & 1023u)) & 1023u));
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    // corr_first: the first chunk that needs the correction -- seg_chunks (the first segment starts from S_0 = 0 and is
    // already right) or 0 (a caller-supplied S_0 is carried into the first segment like any other start state).
    // blockIdx = ((sequence * segments) + segment) * corr_pp + piece; a piece is corr_pc consecutive chunks of the segment.
    const int BH = p.B * p.H, ntot = (p.T + CL - 1) / CL;
    const int seg_first = p.corr_first / p.seg_chunks, nsegc = p.nseg - seg_first;
    const int piece = blockIdx.x % p.corr_pp, ss = blockIdx.x / p.corr_pp;
    const int seq = ss / nsegc, seg = seg_first + (ss - seq * nsegc);
    const int cfirst = seg * p.seg_chunks + piece * p.corr_pc;                              // chunks in processing order
    const int cend = min(min(cfirst + p.corr_pc, (seg + 1) * p.seg_chunks), ntot);
    const int n = cend - cfirst;
    pdl_launch_dependents();                                                               // (the next block's first kernel)
    if (n <= 0) return;                                                                    // (CTA-uniform)
    const int ds = seq / BH, bh = seq - ds * BH, b = bh / p.H, h = bh - b * p.H;
    const DirArgs& D = p.d[ds];
    const bool rev = D.reverse != 0;
    auto issue = [&](int i) {                  // thread 0: r, w, y tiles of the CTA's chunk i into stage i & 1
        const int c = cfirst + i, t0 = (rev ? ntot - 1 - c : c) * CL, st = i & 1;
        mbar_arrive_expect_tx(&sm.bar_in[st], 3u * CL * 128);
        tma_load_3d(sm.raw_r[st], &D.mr, &sm.bar_in[st], h * N, t0, b);
        tma_load_3d(sm.raw_w[st], &D.mw, &sm.bar_in[st], h * N, t0, b);
        tma_load_3d(sm.y[st], &D.my, &sm.bar_in[st], h * N, t0, b);
    };
    if (tid == 0) {
        mbar_init(&sm.bar_in[0], 1);
        mbar_init(&sm.bar_in[1], 1);
        mbar_init(&sm.bar_m, 1);
        fence_barrier_init();
        fence_proxy_async();
        pdl_wait();            // the segment scan (and with it the first kernel) has completed
        issue(0);
        if (n > 1) issue(1);
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&sm.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    pdl_wait();
    // S_start [i][j] fp32 -> bf16 tile [j][i] (16 elements per thread: i = tid >> 2, j = 16 * (tid & 3) ..), once per CTA
    {
        const float* S = p.seg_state + ((size_t)seq * p.nseg + seg) * N * N;
        const int i = tid >> 2, j0 = 16 * (tid & 3);
#pragma unroll
        for (int v4 = 0; v4 < 4; v4++) {
            const float4 f = *reinterpret_cast<const float4*>(S + i * N + j0 + 4 * v4);
            const float vals[4] = {f.x, f.y, f.z, f.w};
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const int j = j0 + 4 * v4 + e;
                *reinterpret_cast<__nv_bfloat16*>(sm.sp + sw_chunk(j, i >> 3) + (i & 7) * 2) = __float2bfloat16_rn(vals[e]);
            }
        }
    }
    const int qtr = warp & 3, half = warp >> 2;
    const int frow = 16 * qtr + (lane >> 2);
    const int fcol = 32 * half + 2 * (lane & 3);
    const uint32_t lane_field = (uint32_t)(qtr * 32) << 16;
    const int I = warp & 3, ch = warp >> 2;
    const int cq = lane & 7, rg = lane >> 3;
    const int c0 = 32 * ch + 4 * cq;
    const int coff = (c0 & 7) * 2, cchunk = c0 >> 3;
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();   // barriers, tmem_base and the S_start tile are visible
    tc_fence_after();
    const uint32_t tY = __shfl_sync(0xffffffffu, sm.tmem_base, 0);
    const uint64_t dk_rc = desc_k(smem_u32(sm.rc)), dm_sp = desc_mn(smem_u32(sm.sp));
    constexpr uint32_t ID = umma_idesc(0, 1, 64, 64);

    for (int i = 0; i < n; i++) {
        const int c = cfirst + i, st = i & 1;
        const int t0 = (rev ? ntot - 1 - c : c) * CL;
        float pre[4];
        {
            const float4 f = *reinterpret_cast<const float4*>(p.seg_prefix + ((size_t)seq * ntot + c) * N + c0);
            pre[0] = f.x; pre[1] = f.y; pre[2] = f.z; pre[3] = f.w;
        }
        mbar_wait(&sm.bar_in[st], (i >> 1) & 1);
        float ro[4][4];
        {
            float l2[4][4];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int pos = 16 * I + 4 * rg + r;
                const int row = rev ? 63 - pos : pos;
                const int off = sw_chunk(row, cchunk) + coff;
                float w4[4];
                ld4bf(sm.raw_r[st] + off, ro[r]);
                ld4bf(sm.raw_w[st] + off, w4);
                const bool valid = t0 + row < p.T;
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    const float x = ex2(w4[e] * kLog2e);
                    l2[r][e] = valid ? fmaxf(-x * kLog2e, -kClampL2) : 0.f;
                }
            }
#pragma unroll
            for (int e = 0; e < 4; e++) {
                l2[1][e] += l2[0][e];
                l2[2][e] += l2[1][e];
                l2[3][e] += l2[2][e];
            }
#pragma unroll
            for (int e = 0; e < 4; e++) {
                float x = l2[3][e];
                float y = __shfl_up_sync(0xffffffffu, x, 8);
                if (rg >= 1) x += y;
                y = __shfl_up_sync(0xffffffffu, x, 16);
                if (rg >= 2) x += y;
                const float excl = x - l2[3][e];
                if (rg == 3) sm.blocktot[I][c0 + e] = x;
#pragma unroll
                for (int r = 0; r < 4; r++) ro[r][e] *= ex2(excl + (r ? l2[r - 1][e] : 0.f));
            }
        }
        __syncthreads();   // S1: blocktot visible
        {
            float fst[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                float s_ = 0.f;
#pragma unroll
                for (int m = 0; m < 4; m++)
                    if (m < I) s_ += sm.blocktot[m][c0 + e];
                fst[e] = ex2(s_) * pre[e];
            }
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int pos = 16 * I + 4 * rg + r;
                const int row = rev ? 63 - pos : pos;
                float x4[4];
#pragma unroll
                for (int e = 0; e < 4; e++) x4[e] = ro[r][e] * fst[e];
                st4bf(sm.rc + sw_chunk(row, cchunk) + coff, x4);
            }
        }
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();   // S2: the operand tile is complete
        tc_fence_after();
        if (warp == 0 && elect_one()) {
#pragma unroll
            for (int ks = 0; ks < 4; ks++) umma(tY, dk_rc + (uint64_t)((ks * 32) >> 4), dm_sp + (uint64_t)((ks * 2048) >> 4), ID, ks > 0);
            umma_commit(&sm.bar_m);
        }
        __syncwarp();
        mbar_wait(&sm.bar_m, i & 1);
        tc_fence_after();
        {
            float x[16];
            tmem_ld_16x32(tY + lane_field + 32 * half, x);
#pragma unroll
            for (int rep = 0; rep < 4; rep++)
#pragma unroll
                for (int hh = 0; hh < 2; hh++) {
                    const int col = fcol + 8 * rep;
                    uint32_t* wp = reinterpret_cast<uint32_t*>(sm.y[st] + sw_chunk(frow + 8 * hh, col >> 3) + (col & 7) * 2);
                    const uint32_t old = *wp;
                    *wp = pack_bf16(__uint_as_float(old << 16) + x[4 * rep + 2 * hh], __uint_as_float(old & 0xffff0000u) + x[4 * rep + 2 * hh + 1]);
                }
        }
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();   // S3: y tile complete; TMEM reads, blocktot and operand-tile reads of this chunk done
        if (tid == 0) {
            tma_store_3d(&D.my, sm.y[st], h * N, t0, b);
            tma_store_commit();
            if (i + 2 < n) {               // refill this stage: the store must have read the y tile first
                tma_store_wait_read<0>();
                issue(i + 2);
            }
        }
    }
    if (tid == 0) tma_store_wait<0>();
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tY) : "memory");
}

// Number of time segments for a launch (0 = unsegmented): only when one CTA per sequence leaves SM slots empty and the
// sequences are long.  The two passes cost 2x the work of one for the exact scan and ~1.75x for the chunk kernel (its
// pass 1 is state-only), so at least min_seg segments (4 / 2), each at least 8 chunks long.
static int pick_segments(const Params& p, size_t scratch_bytes, int min_seg, int* seg_chunks) {
    const int nseq = p.ndir * p.B * p.H;
    const int ntot = (p.T + CL - 1) / CL;
    if (nseq <= 0 || nseq * min_seg > 296 || ntot < 32) return 0;
    int nseg = 296 / nseq;                       // 148 SMs x 2 resident CTAs
    if (nseg > ntot / 8) nseg = ntot / 8;
    if (nseg > kMaxSeg) nseg = kMaxSeg;          // wkv6_seg_scan_kernel keeps one register pair per segment
    if (nseg < min_seg) return 0;
    int sc = (ntot + nseg - 1) / nseg;
    nseg = (ntot + sc - 1) / sc;                 // no empty segments
    const size_t need = (size_t)nseq * (nseg * (N * N + N) + (size_t)ntot * N) * sizeof(float);   // + per-chunk decay prefixes
    if (need > scratch_bytes) return 0;
    *seg_chunks = sc;
    return nseg;
}

// Fills the segmentation fields of q (number of segments, chunks per segment, scratch layout); false: run unsegmented.
bool plan_segments(Params& q, void* scratch, size_t scratch_bytes, int min_seg) {
    int sc = 0;
    const int nseg = scratch ? pick_segments(q, scratch_bytes, min_seg, &sc) : 0;
    if (nseg == 0) return false;
    const int nseq = q.ndir * q.B * q.H;
    q.nseg = nseg;
    q.seg_chunks = sc;
    q.seg_state = static_cast<float*>(scratch);
    q.seg_decay = q.seg_state + (size_t)nseq * nseg * N * N;
    q.seg_prefix = q.seg_decay + (size_t)nseq * nseg * N;
    return true;
}

// Scratch bytes with which a launch of this shape runs time-segmented (0: it runs one CTA per sequence whatever it is given).
size_t segment_scratch_bytes(const Params& p, int min_seg) {
    int sc = 0;
    const int nseg = pick_segments(p, ~(size_t)0, min_seg, &sc);
    return nseg ? (size_t)p.ndir * p.B * p.H * (nseg * (N * N + N) + (size_t)((p.T + CL - 1) / CL) * N) * sizeof(float) : 0;
}

int launch_seg_scan(const Params& q, cudaStream_t st) {
    if (q.nseg > kMaxSeg) return (int)cudaErrorInvalidValue;
    const cudaError_t e = launch_pdl(wkv6_seg_scan_kernel, dim3(q.ndir * q.B * q.H * 16), dim3(256), 0, st, q);
    count_launch(1);
    return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

// Segmented forward: pass 1, scan, pass 2 on the caller's scratch buffer.  Returns -1 when the launch does not qualify.
int launch_fwd_chunk_segmented(const Params& p, void* scratch, size_t scratch_bytes, cudaStream_t st) {
    Params q = p;
    if (!plan_segments(q, scratch, scratch_bytes, 2)) return -1;
    bool any_ckpt = q.skip_y != 0;
    for (int d = 0; d < q.ndir; d++) any_ckpt = any_ckpt || q.d[d].ckpt != nullptr;
    if (!any_ckpt) {
        // inference: ONE pass over the inputs -- outputs from the zero-start state of every segment, the segment scan, then
        // the carried state's share added to y (wkv6_fwd_corr_kernel) -- instead of a state-only pass plus a full pass
        q.seg_mode = 3;
        int e = launch_fwd_chunk(q, st);
        if (e) return e;
        if ((e = launch_seg_scan(q, st))) return e;
        const size_t smem = sizeof(CorrSmem) + 1024;
        static std::atomic<unsigned long long> opted{0};
        int dev = 0;
        cudaError_t ce = cudaGetDevice(&dev);
        if (ce != cudaSuccess) return (int)ce;
        if (!(opted.load(std::memory_order_acquire) & (1ull << (dev & 63)))) {
            if ((ce = cudaFuncSetAttribute(wkv6_fwd_corr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return (int)ce;
            cudaFuncSetAttribute(wkv6_fwd_corr_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
            opted.fetch_or(1ull << (dev & 63), std::memory_order_release);
        }
        const int ntot = (q.T + CL - 1) / CL;
        Params qc = q;
        qc.dbg = nullptr;
        qc.corr_first = q.seg_chunks;
        for (int d = 0; d < q.ndir; d++)
            if (q.d[d].s0) qc.corr_first = 0;
        // pieces per segment: enough CTAs for three per SM, the fewest (waves x chunks per CTA), at least 4 chunks per CTA
        // where the segments allow it (the S_start tile is built once per CTA)
        const int nsegc = q.nseg - qc.corr_first / q.seg_chunks, nss = q.ndir * q.B * q.H * nsegc;
        int best_pp = 1;
        long best_cost = -1;
        for (int pp = 1; pp <= 16 && pp <= q.seg_chunks; pp++) {
            const int pc = (q.seg_chunks + pp - 1) / pp;
            if (pp > 1 && pc < 4) break;
            const long waves = ((long)nss * pp + 3 * 148 - 1) / (3 * 148);
            const long cost = waves * (pc + 1);                  // + 1: the CTA's prologue
            if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_pp = pp; }
        }
        qc.corr_pp = best_pp;
        qc.corr_pc = (q.seg_chunks + best_pp - 1) / best_pp;
        (void)ntot;
        ce = launch_pdl(wkv6_fwd_corr_kernel, dim3(nss * best_pp), dim3(256), smem, st, qc);
        count_launch(1);
        return ce != cudaSuccess ? (int)ce : (int)cudaGetLastError();
    }
    q.seg_mode = 1;
    int e = launch_fwd_chunk(q, st);
    if (e) return e;
    if ((e = launch_seg_scan(q, st))) return e;
    q.seg_mode = 2;
    return launch_fwd_chunk(q, st);
}

int launch_fwd_chunk(const Params& p, cudaStream_t st) {
    const size_t smem = sizeof(LeanSmem) + 1024;
    const bool seg_local = p.nseg > 1 && p.seg_mode != 2;          // pass 1 of either segmented scheme
    cudaError_t e = seg_local ? opt_in_smem_once(wkv6_fwd_chunk_seg_kernel, smem, 2) : opt_in_smem_once(wkv6_fwd_chunk_lean_kernel, smem, 0);
    if (e != cudaSuccess) return (int)e;
    Params q = p;
    q.dbg = debug_trace_buffer();
    const dim3 grid(q.ndir * q.B * q.H * (q.nseg > 0 ? q.nseg : 1));
    e = seg_local ? launch_pdl(wkv6_fwd_chunk_seg_kernel, grid, dim3(256), smem, st, q)
                  : launch_pdl(wkv6_fwd_chunk_lean_kernel, grid, dim3(256), smem, st, q);
    count_launch(1);
    return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

}  // namespace wkv6
