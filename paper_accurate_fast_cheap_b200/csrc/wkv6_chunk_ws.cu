// wkv6_chunk_ws.cu -- chunked WKV6 forward (bf16 I/O) on tcgen05 / TMEM / TMA, warp-specialised version (opt-in: WKV6_FWD_CHUNK_WS=1).
//
// Same function, tiles and MMA groups as wkv6_chunk.cu / wkv6_chunk_lean.cu (reference: wenet/rwkv_v6/cuda/
// wkv6_cuda.cu:9-63; chunked form SURVEY.md App. A.5).  Measured on B200 (scripts/trace_phases.py, tests/probes/
// umma_timing.cu): one tcgen05.mma costs the ISSUING thread ~37 cycles (+ a ~270-cycle restart bubble when the tensor
// pipe was idle), so the 28 MMAs of a chunk kept five of the eight compute warps of the previous kernel parked in
// __syncwarp() for 700-1000 of 4750 cycles per chunk, on the critical path.  Here:
//   * warps 0-7 (256 threads) only compute: P^T readout, tile building of the next chunk, Y readout, state update;
//   * warp 8 (one elected lane) issues every MMA, TMA load and TMA store;
//   * the two sides talk through mbarriers only: compute threads ARRIVE (never block) when a tile set is complete
//     (bar_pt, bar_raw, bar_tiles, bar_ys, bar_sp: 256 arrivals each); the issuer commits MMA groups to bar_p / bar_s /
//     bar_y, which compute threads wait on just before they need the accumulator.  One named barrier (256 threads,
//     for the sub-block decay totals) is the only CTA-wide rendezvous left in the chunk loop.
// Per chunk c (compute side):  wait G1(c) -> P^T readout -> [tiles of chunk c+1: phase 1, B1, phase 2] -> wait G2(c)
//   -> Y readout -> wait G4(c) (long done) -> state update.   Issuer side:  G2(c) | rkw(c+2) | G1(c+1) | y(c) store |
//   checkpoint(c+1), G3(c+1), G4(c+1) | v(c+2).
#include "wkv6_fwd_tiles.cuh"
#include <cstdio>
#include <cstdlib>

namespace wkv6 {

struct WsSmem {
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
    float bonus[2][2][CL];                   // [chunk parity][channel half][row]: sum_j r u k
    alignas(8) uint64_t bar_rkw, bar_v[2];   // TMA -> compute / issuer
    uint64_t bar_p, bar_s, bar_y;            // MMA groups done (tcgen05.commit): G1 | G3+G4 | G2
    uint64_t bar_pt, bar_raw, bar_tiles, bar_ys, bar_sp;   // compute -> issuer, 256 arrivals each
    uint32_t tmem_base;
};

constexpr int kWsCompute = 256;   // threads of warps 0-7

__device__ __forceinline__ void bar_sync_compute() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------------------------
// issuer warp (warp 8): one elected lane runs the whole loop
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ws_issuer(const Params& p, WsSmem& sm, const DirArgs& D, const int b, const int h, const int bh,
                                          const uint32_t tbase, const bool REV) {
    const int nchunks = (p.T + CL - 1) / CL;
    const uint32_t tP = tbase, tY = tbase + 64, tD = tbase + 128;
    const uint32_t rst_u32 = smem_u32(sm.rst), kt_u32 = smem_u32(sm.kt), roff_u32 = smem_u32(sm.roff);
    const uint64_t dk_rst = desc_k(rst_u32), dk_kt = desc_k(kt_u32), dk_roff = desc_k(roff_u32);
    const uint64_t dm_sp = desc_mn(smem_u32(sm.sp)), dm_kt = desc_mn(kt_u32), dm_pt = desc_mn(smem_u32(sm.pt)),
                   dm_v0 = desc_mn(smem_u32(sm.v[0])), dm_v1 = desc_mn(smem_u32(sm.v[1]));
#define DOFF(bytes) ((uint64_t)((bytes) >> 4))
    constexpr uint32_t ID_G1 = umma_idesc(0, 0, 64, 16);
    constexpr uint32_t ID_G2 = umma_idesc(1, 1, 64, 64);
    constexpr uint32_t ID_G3 = umma_idesc(0, 1, 64, 64);

    auto issue_rkw = [&](int c) {
        const int q = REV ? nchunks - 1 - c : c;
        mbar_arrive_expect_tx(&sm.bar_rkw, 3u * CL * 128);
        tma_load_3d(sm.raw_r, &D.mr, &sm.bar_rkw, h * N, q * CL, b);
        tma_load_3d(sm.raw_k, &D.mk, &sm.bar_rkw, h * N, q * CL, b);
        tma_load_3d(sm.raw_w, &D.mw, &sm.bar_rkw, h * N, q * CL, b);
    };
    auto issue_v = [&](int c) {
        const int q = REV ? nchunks - 1 - c : c;
        mbar_arrive_expect_tx(&sm.bar_v[c & 1], (uint32_t)CL * 128);
        tma_load_3d(sm.v[c & 1], &D.mv, &sm.bar_v[c & 1], h * N, q * CL, b);
    };
    // raw r/k/w of chunk k consumed -> load chunk k+1 (and, once, the second V tile)
    auto front_raw = [&](int k) {
        mbar_wait(&sm.bar_raw, k & 1);
        if (k + 1 < nchunks) {
            issue_rkw(k + 1);
            if (k == 0) issue_v(1);
        }
    };
    // operand tiles of chunk k complete -> G1 (P^T, four sub-block groups)
    auto front_g1 = [&](int k) {
        mbar_wait(&sm.bar_tiles, k & 1);
        tc_fence_after();
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int rowlo = REV ? 48 - 16 * i : 16 * i;
#pragma unroll
            for (int ks = 0; ks < 4; ks++)
                umma(tP + rowlo, dk_kt + DOFF(ktbase(REV, i) + ks * 32), dk_roff + DOFF(rowlo * 128 + ks * 32), ID_G1, ks > 0);
        }
        umma_commit(&sm.bar_p);
    };
    // state at the start of chunk k is in sp -> checkpoint k, G3 (Y = Rst S), G4 (dS = Ktil(4)^T V)
    auto front_s = [&](int k) {
        mbar_wait(&sm.bar_sp, k & 1);
        if (D.ckpt) {
            tma_store_2d(&D.mc, sm.sp, 0, (bh * nchunks + k) * N);
            tma_store_commit();
        }
        const uint64_t dm_v = (k & 1) ? dm_v1 : dm_v0;
        mbar_wait(&sm.bar_v[k & 1], (k >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < 4; ks++) umma(tY, dk_rst + DOFF(ks * 32), dm_sp + DOFF(ks * 2048), ID_G3, ks > 0);
#pragma unroll
        for (int ks = 0; ks < 4; ks++) umma(tD, dm_kt + DOFF(ktbase(REV, 4) + ks * 2048), dm_v + DOFF(ks * 2048), ID_G2, ks > 0);
        umma_commit(&sm.bar_s);
    };

    tma_prefetch_desc(&D.mr); tma_prefetch_desc(&D.mk); tma_prefetch_desc(&D.mv); tma_prefetch_desc(&D.mw);
    if (!p.skip_y) tma_prefetch_desc(&D.my);
    if (nchunks == 0) return;
    issue_rkw(0);
    issue_v(0);
    front_raw(0);
    front_g1(0);
    front_s(0);
    for (int c = 0; c < nchunks; c++) {
        const int q = REV ? nchunks - 1 - c : c;
        const bool more = c + 1 < nchunks;
        // G2: Y += P V, once the masked P^T tile of chunk c is in smem.  Every earlier TMA store (checkpoint c out of
        // sp, y of chunk c-1 out of ys) has been read by then: compute threads rewrite ys / sp only after bar_y(c).
        mbar_wait(&sm.bar_pt, c & 1);
        tma_store_wait_read<0>();
        tc_fence_after();
        {
            const uint64_t dm_v = (c & 1) ? dm_v1 : dm_v0;
#pragma unroll
            for (int ks = 0; ks < 4; ks++) umma(tY, dm_pt + DOFF(ks * 2048), dm_v + DOFF(ks * 2048), ID_G2, 1);
            umma_commit(&sm.bar_y);
        }
        if (more) {
            front_raw(c + 1);
            front_g1(c + 1);
        }
        mbar_wait(&sm.bar_ys, c & 1);
        if (!p.skip_y) {
            tma_store_3d(&D.my, sm.ys, h * N, q * CL, b);
            tma_store_commit();
        }
        if (more) front_s(c + 1);
        if (c + 2 < nchunks) {   // v[c&1] is free once G2(c) and G4(c) have completed (both long done here)
            mbar_wait(&sm.bar_y, c & 1);
            issue_v(c + 2);
        }
    }
    tma_store_wait<0>();
#undef DOFF
}

// ---------------------------------------------------------------------------------------------------------------
// compute warps (0-7)
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ws_compute(const Params& p, WsSmem& sm, const DirArgs& D, const int h, const int bh,
                                           const int warp, const uint32_t tbase, float (&st)[16], const uint32_t (&stoff)[2],
                                           const bool REV) {
    const int tid = threadIdx.x, lane = tid & 31;
    const bool wkv6_trace_on = WKV6_TRACE_ON(p, tid);
    const int T = p.T;
    const int nchunks = (T + CL - 1) / CL;
    const uint32_t tP = tbase, tY = tbase + 64, tD = tbase + 128;

    // TMEM fragment coordinates (16x256b.x4): rows frow, frow+8; columns fcol + 8*rep + {0,1}
    const int qtr = warp & 3, half = warp >> 2;
    const int frow = 16 * qtr + (lane >> 2);
    const uint32_t lane_field = (uint32_t)(qtr * 32) << 16;
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
    const uint2 u2 = *reinterpret_cast<const uint2*>(reinterpret_cast<const uint8_t*>(D.u) + (h * N + c0) * 2);

    const uint32_t rawr_u32 = smem_u32(sm.raw_r), rawk_u32 = smem_u32(sm.raw_k), raww_u32 = smem_u32(sm.raw_w);
    const uint32_t roff_u32 = smem_u32(sm.roff), rst_u32 = smem_u32(sm.rst), kt_u32 = smem_u32(sm.kt);
    const uint32_t pt_u32 = smem_u32(sm.pt), ys_u32 = smem_u32(sm.ys), sp_u32 = smem_u32(sm.sp);

    auto store_frag = [&](uint32_t tile_u32, const float (&x)[16]) {   // fp32 fragment -> bf16 swizzled tile
        uint32_t pk[8];
#pragma unroll
        for (int i = 0; i < 8; i++) pk[i] = pack_bf16(x[2 * i], x[2 * i + 1]);
        stsm_x4(tile_u32 + stoff[0], pk[0], pk[1], pk[2], pk[3]);
        stsm_x4(tile_u32 + stoff[1], pk[4], pk[5], pk[6], pk[7]);
    };

    uint32_t ro2[4][2], ko2[4][2];   // Rtil / Ktil(own) of this thread's 4 rows x 4 channels, packed bf16x2
    float fdec[2] = {1.f, 1.f}, fnext[2] = {1.f, 1.f};   // 2^(A_L[j]), j = frow, frow + 8: current chunk / chunk built last

    // ---- phase 1 of chunk c: decays, in-block cumulative sums, Rtil, Ktil(own block), bonus
    auto phase1 = [&](const int c) {
        const int q = REV ? nchunks - 1 - c : c;
        const int t0 = q * CL;
        mbar_wait(&sm.bar_rkw, c & 1);
        {
            const int c_ts = c - 1;
            WKV6_TS(0, 10);
        }
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
        // this thread's part of the raw tiles is in registers: the issuer may overwrite them with chunk c+1 (arrive has
        // release semantics: the loads above are performed before it); a whole iteration of prefetch distance
        mbar_arrive(&sm.bar_raw);
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
                sm.bonus[c & 1][ch][REV ? 63 - pos : pos] = kk_;
            }
        }
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
#pragma unroll
        for (int r = 0; r < 4; r++) {
            float fr[4], fk[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const float aex = r ? excl[e] + m[r - 1][e] : excl[e];          // -(A_p  - a_I) >= 0
                const float ain = r == 3 ? m[3][e] : excl[e] + m[r][e];         // -(A+_p - a_I) >= 0
                fr[e] = ex2(-aex);
                fk[e] = ex2(ain);
            }
            ro2[r][0] = bmul2(rr[r][0], pack_bf16(fr[0], fr[1]));
            ro2[r][1] = bmul2(rr[r][1], pack_bf16(fr[2], fr[3]));
            ko2[r][0] = bmul2(kk[r][0], pack_bf16(fk[0], fk[1]));
            ko2[r][1] = bmul2(kk[r][1], pack_bf16(fk[2], fk[3]));
            sts64(roff_u32 + offs[r], ro2[r][0], ro2[r][1]);
            sts64(kt_u32 + vb_own + offs[r], ko2[r][0], ko2[r][1]);
        }
    };
    // ---- phase 2 of chunk c (after the sub-block totals are visible): Rst, Ktil variants, whole-chunk decay
    auto phase2 = [&](const int c) {
        const float* bt = &sm.blocktot[c & 1][0][0];
        switch (I) {   // warp-uniform
            case 0: lean_phase2<0>(REV, rst_u32, kt_u32, bt, c0, offs, ro2, ko2, false); break;
            case 1: lean_phase2<1>(REV, rst_u32, kt_u32, bt, c0, offs, ro2, ko2, false); break;
            case 2: lean_phase2<2>(REV, rst_u32, kt_u32, bt, c0, offs, ro2, ko2, false); break;
            default: lean_phase2<3>(REV, rst_u32, kt_u32, bt, c0, offs, ro2, ko2, false); break;
        }
        fnext[0] = ex2(-((bt[frow] + bt[N + frow]) + (bt[2 * N + frow] + bt[3 * N + frow])));
        fnext[1] = ex2(-((bt[frow + 8] + bt[N + frow + 8]) + (bt[2 * N + frow + 8] + bt[3 * N + frow + 8])));
    };

    // c = -1 is the prologue iteration: only the tiles of chunk 0 are built (one instantiation of the tile code)
    for (int c = nchunks > 0 ? -1 : 0; c < nchunks; c++) {
        const uint32_t par = c & 1;
        const bool more = c + 1 < nchunks;
        const bool live = c >= 0;
        const int c_ts = c;
        WKV6_TS(0, 0);
        if (p.dbg && blockIdx.x == 0 && tid == 0 && live && c < 32) p.dbg[32 + c] = clock64();   // loop-top stamp of every chunk

        // ---------------- P^T readout: causal mask + bonus diagonal -> bf16 tile (MN-major A of G2)
        if (live) {
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
                    for (int hh = 0; hh < 2; hh++) bon[hh] = sm.bonus[par][0][frow + 8 * hh] + sm.bonus[par][1][frow + 8 * hh];
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
        mbar_arrive(&sm.bar_pt);   // -> issuer: G2(c)
        }
        WKV6_TS(0, 2);

        // ---------------- tile building of the NEXT chunk (independent of the state; covers G2)
        if (more) {
            phase1(c + 1);         // roff / Ktil(own) are free: G1(c) has completed (bar_p)
            WKV6_TS(0, 3);
            bar_sync_compute();    // B1: sub-block totals visible
            WKV6_TS(0, 4);
        }
        if (live) mbar_wait(&sm.bar_s, par); // G3(c), G4(c) done: rst, Ktil(4), sp may be rewritten; dS is in TMEM
        WKV6_TS(0, 5);
        if (more) {
            phase2(c + 1);
            fence_proxy_async();
            mbar_arrive(&sm.bar_tiles);   // -> issuer: G1(c+1)
        }
        WKV6_TS(0, 6);
        if (!live) {
            fdec[0] = fnext[0]; fdec[1] = fnext[1];
            continue;
        }

        // ---------------- Y readout -> bf16 staging tile -> TMA store (issuer)
        mbar_wait(&sm.bar_y, par);   // also: checkpoint c has left sp, y of chunk c-1 has left ys (issuer waited)
        WKV6_TS(0, 7);
        tc_fence_after();
        if (!p.skip_y) {
            float yv[16];
            tmem_ld_16x32(tY + lane_field + 32 * half, yv);
            store_frag(ys_u32, yv);
        }
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(&sm.bar_ys);     // -> issuer: y(c) store
        WKV6_TS(0, 8);

        // ---------------- state update:  S^T[j][i] = S^T[j][i] 2^(A_L[j]) + dS^T[j][i]
        {
            float dv[16];
            tmem_ld_16x32(tD + lane_field + 32 * half, dv);
#pragma unroll
            for (int rep = 0; rep < 4; rep++) {
                st[4 * rep + 0] = fmaf(st[4 * rep + 0], fdec[0], dv[4 * rep + 0]);
                st[4 * rep + 1] = fmaf(st[4 * rep + 1], fdec[0], dv[4 * rep + 1]);
                st[4 * rep + 2] = fmaf(st[4 * rep + 2], fdec[1], dv[4 * rep + 2]);
                st[4 * rep + 3] = fmaf(st[4 * rep + 3], fdec[1], dv[4 * rep + 3]);
            }
            store_frag(sp_u32, st);
            fdec[0] = fnext[0]; fdec[1] = fnext[1];
        }
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(&sm.bar_sp);     // -> issuer: checkpoint(c+1), G3(c+1), G4(c+1)
        WKV6_TS(0, 9);
    }
}

// REV is a runtime flag (one instantiation), see wkv6_chunk_lean.cu
__device__ __forceinline__ void ws_body(const Params& p, WsSmem& sm, const DirArgs& D, const int b, const int h, const int bh,
                                        const bool REV) {
    const int tid = threadIdx.x, lane = tid & 31;
    const bool wkv6_trace_on = WKV6_TRACE_ON(p, tid);
    // warp index through a lane-0 broadcast: everything derived from it stays in uniform registers
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);

    if (tid == 0) {
        mbar_init(&sm.bar_rkw, 1);
        mbar_init(&sm.bar_v[0], 1);
        mbar_init(&sm.bar_v[1], 1);
        mbar_init(&sm.bar_p, 1);
        mbar_init(&sm.bar_s, 1);
        mbar_init(&sm.bar_y, 1);
        mbar_init(&sm.bar_pt, kWsCompute);
        mbar_init(&sm.bar_raw, kWsCompute);
        mbar_init(&sm.bar_tiles, kWsCompute);
        mbar_init(&sm.bar_ys, kWsCompute);
        mbar_init(&sm.bar_sp, kWsCompute);
        fence_barrier_init();
    }
    if (warp == 8) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&sm.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // rows of a Ktil variant that are never written must still hold finite bf16 values
    for (int i = tid; i < 28 * 1024 / 16; i += 288) reinterpret_cast<uint4*>(sm.kt)[i] = make_uint4(0, 0, 0, 0);

    // state S^T[j][i] in the TMEM fragment layout of the compute warps: j = frow + 8*hh, i = fcol + 8*rep + e
    const int qtr = warp & 3, half = (warp >> 2) & 1;
    const int frow = 16 * qtr + (lane >> 2);
    const int fcol = 32 * half + 2 * (lane & 3);
    // stmatrix row addresses of this lane for the two x4 stores of a 16x32 fragment: matrix mi = lane>>3 of store s is
    // (rep = 2s + (mi>>1), hh = mi&1), i.e. registers pk[4s .. 4s+3] in order
    uint32_t stoff[2];
    {
        const int mi = lane >> 3, rr = lane & 7;
        const int row = 16 * qtr + 8 * (mi & 1) + rr;
        stoff[0] = (uint32_t)sw_chunk(row, 4 * half + (mi >> 1));
        stoff[1] = (uint32_t)sw_chunk(row, 4 * half + 2 + (mi >> 1));
    }
    float st[16];
    if (warp < 8) {
        const float* s0 = D.s0 ? D.s0 + (size_t)bh * N * N : nullptr;
        uint32_t pk[8];
#pragma unroll
        for (int rep = 0; rep < 4; rep++)
#pragma unroll
            for (int hh = 0; hh < 2; hh++) {
#pragma unroll
                for (int e = 0; e < 2; e++)
                    st[4 * rep + 2 * hh + e] = s0 ? s0[(fcol + 8 * rep + e) * N + frow + 8 * hh] : 0.f;
                pk[2 * rep + hh] = pack_bf16(st[4 * rep + 2 * hh], st[4 * rep + 2 * hh + 1]);
            }
        const uint32_t sp_u32 = smem_u32(sm.sp);
        stsm_x4(sp_u32 + stoff[0], pk[0], pk[1], pk[2], pk[3]);
        stsm_x4(sp_u32 + stoff[1], pk[4], pk[5], pk[6], pk[7]);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = __shfl_sync(0xffffffffu, sm.tmem_base, 0);

    if (warp == 8) {
        if (elect_one()) ws_issuer(p, sm, D, b, h, bh, tbase, REV);
        __syncwarp();
    } else {
        mbar_arrive(&sm.bar_sp);   // phase 0: the initial state is in sp
        ws_compute(p, sm, D, h, bh, warp, tbase, st, stoff, REV);
        if (D.sT) {
            float* sT = D.sT + (size_t)bh * N * N;
#pragma unroll
            for (int rep = 0; rep < 4; rep++)
#pragma unroll
                for (int hh = 0; hh < 2; hh++)
#pragma unroll
                    for (int e = 0; e < 2; e++) sT[(fcol + 8 * rep + e) * N + frow + 8 * hh] = st[4 * rep + 2 * hh + e];
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tbase) : "memory");
}

__global__ void __launch_bounds__(288, 2) wkv6_fwd_chunk_ws_kernel(const __grid_constant__ Params p) {
    extern __shared__ __align__(1024) uint8_t smem_bytes[];
    // offset arithmetic on the __shared__ array (not a uintptr_t round trip) keeps the address space visible to the
    // compiler: LDS/STS instead of generic LD/ST
    WsSmem& sm = *reinterpret_cast<WsSmem*>(smem_bytes + ((1024u - (smem_u32(smem_bytes) & 1023u)) & 1023u));
    const int BH = p.B * p.H;
    // direction from the block index PARITY: CTAs i and i + 148 share an SM, so both run the same template
    // instantiation (ncu: "no instruction" was the top stall reason with mixed pairs thrashing the instruction cache)
    const bool il = p.ndir == 2 && p.dir_interleave;
    const int ds = il ? (int)(blockIdx.x & 1) : (int)(blockIdx.x / BH);
    const int bh = il ? (int)(blockIdx.x >> 1) : (int)(blockIdx.x - ds * BH);
    const int b = bh / p.H, h = bh - b * p.H;
    const DirArgs& D = p.d[ds];
    if (p.dbg && threadIdx.x == 0 && blockIdx.x < 1024) {
        unsigned long long t;
        uint32_t smid;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        p.dbg[128 + 4 * blockIdx.x] = t;
        p.dbg[129 + 4 * blockIdx.x] = smid;
    }
    ws_body(p, sm, D, b, h, bh, D.reverse != 0);
    if (p.dbg && threadIdx.x == 0 && blockIdx.x < 1024) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.dbg[130 + 4 * blockIdx.x] = t;
    }
}

int launch_fwd_chunk_ws(const Params& p, cudaStream_t st) {
    const size_t smem = sizeof(WsSmem) + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(wkv6_fwd_chunk_ws_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        cudaFuncSetAttribute(wkv6_fwd_chunk_ws_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        if (getenv("WKV6_DEBUG_OCC")) {
            int nb = -1;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, wkv6_fwd_chunk_ws_kernel, 288, smem);
            fprintf(stderr, "[wkv6] fwd_chunk_ws: smem %zu B, resident CTAs/SM %d\n", smem, nb);
        }
        attr_set = true;
    }
    Params q = p;
    q.dbg = debug_trace_buffer();
    static const bool il = getenv("WKV6_DIR_INTERLEAVE") != nullptr;
    q.dir_interleave = il ? 1 : 0;
    wkv6_fwd_chunk_ws_kernel<<<dim3(q.ndir * q.B * q.H), dim3(288), smem, st>>>(q);
    count_launch(1);
    return (int)cudaGetLastError();
}

}  // namespace wkv6
