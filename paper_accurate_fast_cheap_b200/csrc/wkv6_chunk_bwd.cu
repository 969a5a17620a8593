// wkv6_chunk_bwd.cu -- chunked (dense-contraction) WKV6 backward for bf16 I/O on tcgen05 / TMEM / TMA (sm_100a).
//
// Same function as the reference's four backward kernels (wenet/rwkv_v6/cuda/wkv6_cuda.cu:66-263; with initial state
// and gs: wkv6state_cuda.cu:68-293), restated per chunk of L = 64 steps so that all O(L*N*N) work is dense MMA.
// ONE sweep against the direction's time order; the forward state at every chunk start comes from bf16 checkpoints
// written by the chunked forward (wkv6_chunk.cu), the reverse state C[i][j] = sum_{p>t} (prod d) gy_p[i] r_p[j] is
// carried in registers.  Notation (log2 domain): l_p = -exp(w_p) log2e, A_p = exclusive cumsum
// in the chunk, A+_p = A_p + l_p, a_I = A at the start of 16-step sub-block I, A_L = chunk total.
//
//   Q[p][q]   = gy_p . v_q                                  (contraction over the value channel i)
//   P[p][q]   = sum_j r_p k_q 2^(A_p - A+_q)   (q before p),   P[p][p] = sum_j r_p u k_p          (as in the forward)
//   gv_q      = sum_p P[p][q] gy_p  +  sum_j k_q 2^(A_L - A+_q) C_L[.][j]
//   gr_p[j]   = 2^(A_p - a_I) [ sum_{q<p} Q[p][q] k_q 2^(a_I - A+_q)  +  2^(a_I) sum_i gy_p[i] S_0[i][j] ] + u k_p (v_p.gy_p)
//   gk_q[j]   = 2^(a_{J+1} - A+_q) [ sum_{p>q} Q[p][q] r_p 2^(A_p - a_{J+1}) + 2^(A_L - a_{J+1}) sum_i v_q[i] C_L[i][j] ]
//               + u r_q (v_q.gy_q)
//   C_0       = C_L 2^(A_L) + sum_p gy_p (r_p 2^(A_p))^T
//   gw_p[j]   = -exp(w_p) H_p,   H_p = sum_i S_0[i][j] C_0[i][j] + sum_{q<p} (k_q gkC_q - r_q grS_q) - r_p grS_p
// (grS / gkC = gr / gk without their u terms).  The last line is the reference's gw (kernel_backward_201) written as
// a chunk-local prefix sum re-anchored at every chunk start by S_0 . C_0: no O(T) array, no T limit, and rounding
// errors cannot accumulate beyond 64 steps.  Every exponent that reaches an MMA operand is relative to the start
// (forward factors) or the end (gk factors) of the row's own 16-step sub-block: bounded by 2^115, finite in bf16.
//
// MMA groups per chunk (tcgen05.mma kind::f16, bf16 operands, fp32 accumulators in TMEM, one issuing thread):
//   phase 1:  Q = G V^T;  P^T (4 x N16, as forward G1);  RS^T[j][p] = S0^T G^T;  KC^T[j][q] = C^T V^T
//   phase 2:  R^T[j][p in I] = Ktil(I)^T Qm^T (4 x N16);  K^T[j][q in J] = Rtil'(J)^T Qm (4 x N16);
//             GV[q][i] = P^T G + Kend C^T;  dC^T[j][i] = Rst^T G
// 8 accumulators x 64 columns = all 512 TMEM columns -> one CTA per SM: 512 threads (16 warps, 128 registers each), so
// that every scheduler of the SM has four warps to hide the latencies of the elementwise phases with.
#include <chrono>
#include "wkv6_fwd_tiles.cuh"

namespace wkv6 {

constexpr int XSTR = 68;   // row stride (floats) of the fp32 transposition buffers: conflict-free scalar stores

struct BwdChunkSmem {
    alignas(1024) uint8_t raw_r[CL * 128];   // TMA landing tiles (bf16, 128B swizzle), rows = time, cols = channel
    alignas(1024) uint8_t raw_k[CL * 128];
    alignas(1024) uint8_t raw_w[CL * 128];
    alignas(1024) uint8_t v[2][CL * 128];    // V and G: MMA operands straight from TMA, double-buffered (loaded a
    alignas(1024) uint8_t g[2][CL * 128];    // whole chunk ahead)
    alignas(1024) uint8_t sp[N * 128];       // S_0^T [j][i] bf16 (checkpoint, TMA)
    alignas(1024) uint8_t cp[N * 128];       // C^T   [j][i] bf16 copy of the reverse state
    alignas(1024) uint8_t roff[CL * 128];    // r 2^(A_p - a_I)
    alignas(1024) uint8_t rst[CL * 128];     // r 2^(A_p)
    alignas(1024) uint8_t rst_lo[CL * 128];  // low-order halves (see st4bf_hilo)
    alignas(1024) uint8_t kt_lo[20 * 1024];
    alignas(1024) uint8_t rp_lo[20 * 1024];
    alignas(1024) uint8_t kt[28 * 1024];     // Ktil(0..3) compact (20 KB) + Kend (8 KB); later: fp32 buffer xr
    alignas(1024) uint8_t rp[20 * 1024];     // Rtil'(0..3) compact; later: fp32 buffer xk
    alignas(1024) uint8_t qm[CL * 128];      // masked Q   [p][q];   after the phase-2 MMAs: staging tile of gr
    alignas(1024) uint8_t qmt[CL * 128];     // masked Q^T [q][p];   after the phase-2 MMAs: staging tile of gk
    alignas(1024) uint8_t pt[CL * 128];      // masked P^T [q][p] with the u-bonus diagonal; later: staging tile of gv
    alignas(1024) uint8_t o_gw[CL * 128];    // staging tile of gw (TMA store)
    float blocktot[4][N];                    // log2 decay sums of the four sub-blocks (direction order), per channel
    float dtot[4][N];                        // sub-block sums of (k gkC - r grS), per channel
    float hst[4][N];                         // S_0 . C_0 partial sums (four column quarters)
    float bonus[4][CL];                      // sum_j r u k per row (four channel quarters)
    float vg[CL];                            // v_p . gy_p per row
    float gured[16][N / 4];
    alignas(8) uint64_t bar_rkw, bar_vg[2], bar_sp, bar_m1, bar_mp, bar_m2, bar_m2b;
    uint32_t tmem_base;
};
static_assert(sizeof(BwdChunkSmem) <= 226 * 1024, "shared memory budget");
static_assert(CL * XSTR * 4 <= 20 * 1024, "xk must fit into the rp region");

constexpr int kBwdThreads = 512;
// Which operand tiles carry low-order halves (st2bf_hilo: two MMAs into the same accumulator).  Measured on one B200 box
// (scripts/gpu_variants.sh, cfg2 shape): everything on 132.4 us, gw <= 4.7e-3 on every test shape; without the halves of
// the ADJACENT sub-block's variants 128.6 us, but gw of the strong-decay case ("stress", w ~ U(-8, 1.5)) 2.5e-2 -- over the
// 2e-2 tolerance; without the halves of r 2^(A_p) (reverse-state update) another 1.4 us, gs0 3e-6 -> 2e-3.  Both stay on.
#ifndef WKV6_ADJLO
#define WKV6_ADJLO 1
#endif
#ifndef WKV6_RSTLO
#define WKV6_RSTLO 1
#endif
constexpr bool kAdjLo = WKV6_ADJLO != 0;
constexpr bool kRstLo = WKV6_RSTLO != 0;   // low-order halves also for the adjacent sub-block's variant (experiment)

__global__ void __launch_bounds__(kBwdThreads, 1) wkv6_bwd_chunk_kernel(const __grid_constant__ Params p) {
    extern __shared__ __align__(1024) uint8_t smem_bytes[];
    // offset arithmetic on the __shared__ array (not a uintptr_t round trip) keeps the address space visible to the
    // compiler: LDS/STS instead of generic LD/ST
    BwdChunkSmem& sm = *reinterpret_cast<BwdChunkSmem*>(smem_bytes + ((1024u - (smem_u32(smem_bytes) & 1023u)) & 1023u));

    const int tid = threadIdx.x, lane = tid & 31;
    const bool wkv6_trace_on = WKV6_TRACE_ON(p, tid);
    // warp index through a lane-0 broadcast: keeps MMA operands / role branches in uniform registers (a plain tid >> 5
    // turns every UTCHMMA into a waterfall loop of R2UR broadcasts)
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int BH = p.B * p.H;
    const int T = p.T;
    const int nchunks = (T + CL - 1) / CL;
    // ---- the CTA's schedule: a list of "pieces" (runs of consecutive chunks of one sequence, in PROCESSING order: chunk 0 is
    // the one the reverse sweep visits first), walked by ONE flat loop over chunks so that the loads of the next chunk are
    // in flight across piece boundaries.  Three launch shapes:
    //  * one CTA per sequence (grid = sequences): a single piece of kind W (whole sequence);
    //  * time-segmented (launch_bwd_chunk_segmented below; nseg > 1): CTA = (sequence, segment), a single piece of kind SEG
    //    that starts from the segment scan's state and leaves a partial gu;
    //  * balanced persistent grid (launch_bwd_chunk_balanced; bal_ncta > 0): every direction's (sequence, chunk) units are cut
    //    into bal_ncta equal contiguous ranges, one CTA each.  A range = [B: rest of a sequence another CTA began]
    //    [whole sequences] [A: first part of a sequence the next CTA finishes]; the CTA runs A FIRST and hands the reverse
    //    state and the partial gu over through a record in the scratch (data, __threadfence, flag = the launch's epoch with
    //    release semantics), then the whole sequences, and B LAST (acquire on the previous CTA's record, which was written
    //    right at that CTA's start): nobody waits for long, and with blocks dispatched in index order never forever.
    // The unit (sequence, chunk) of iteration `it` is a closed-form function of the loop counter and CTA constants -- no
    // loop-carried schedule state, divisions by multiplication with host-made reciprocals -- so that ptxas keeps every TMA
    // coordinate in uniform registers (loop-carried state ended up in vector registers: every TMA issue a waterfall loop).
    enum { kA = 0, kW = 1, kB = 2, kSEG = 3 };
    const int nseg = p.nseg > 1 ? p.nseg : 1;
    int ds, seq = 0, seg = 0, nunits;
    int bhA = 0, lenA = 0, bhW = 0, nW = 0, bhB = 0, cB0 = 0;
    if (p.bal_ncta > 0) {
        ds = blockIdx.x / p.bal_ncta;
        const unsigned j = blockIdx.x - ds * p.bal_ncta, U = (unsigned)(BH * nchunks);    // (j + 1) U < 2^31: pick_balanced
        const int gs = (int)(j * U / (unsigned)p.bal_ncta), ge = (int)((j + 1) * U / (unsigned)p.bal_ncta);
        const int gsu = min((gs + nchunks - 1) / nchunks * nchunks, ge), ged = max(ge / nchunks * nchunks, gsu);
        bhA = ged / nchunks; lenA = ge - ged;
        bhW = gsu / nchunks; nW = (ged - gsu) / nchunks;
        bhB = gs / nchunks;  cB0 = gs - bhB * nchunks;          // B exists when gsu > gs
        nunits = ge - gs;
    } else {
        seq = blockIdx.x / nseg; seg = blockIdx.x - seq * nseg;
        ds = seq / BH;
        bhB = seq - ds * BH;                                    // one piece, in the B slot: chunks cB0 .. cB0 + nunits - 1
        cB0 = nseg > 1 ? seg * p.seg_chunks : 0;
        nunits = nseg > 1 ? min(p.seg_chunks, nchunks - cB0) : nchunks;
    }
    const int lenW = nW * nchunks;
    const int stageB = p.bal_ncta > 0 ? kB : nseg > 1 ? kSEG : kW;
    struct Unit { int stage, bh, b, h, c; };                    // c: chunk of the sequence in processing order
    auto udiv = [](int x, unsigned magic) { return (int)(((unsigned long long)(unsigned)x * magic) >> 31); };   // x / d, magic = 2^31 / d + 1
    auto unit_at = [&](int it) {
        Unit u;
        const int x = it - lenA;
        const int iw = udiv(max(x, 0), p.bal_magic_n);
        u.stage = x < 0 ? kA : x < lenW ? kW : stageB;
        u.bh = x < 0 ? bhA : x < lenW ? bhW + iw : bhB;
        u.c = x < 0 ? it : x < lenW ? x - iw * nchunks : cB0 + (x - lenW);
        u.b = udiv(u.bh, p.bal_magic_h);
        u.h = u.bh - u.b * p.H;
        return u;
    };
    const Unit cu0 = unit_at(0);
    const DirArgs& D = p.d[ds];
    const bool rev = D.reverse != 0;
    // hand-over records of the balanced grid: [64 i][64 j] reverse state, [64] partial gu, flag
    constexpr size_t kRecBytes = (size_t)N * N * 4 + 512;
    uint8_t* const rec_in = p.bal_rec + (size_t)blockIdx.x * kRecBytes;    // written by CTA blockIdx.x - 1 (its piece A)
    uint8_t* const rec_out = rec_in + kRecBytes;

    // compact variant tiles: virtual base (row 0) of every variant inside sm.kt / sm.rp
    const int vbK[4] = {rev ? 12288 : 0, rev ? 10240 : 2048, 6144, rev ? 0 : 12288};   // Ktil(I): sub-blocks <= I
    const int vbR[4] = {0, rev ? 8192 : 6144, rev ? 14336 : 10240, rev ? 18432 : 12288};   // Rtil'(J): sub-blocks >= J
    uint8_t* const kend = sm.kt + 20 * 1024;
    float* const xr = reinterpret_cast<float*>(sm.kt);   // [p][XSTR] fp32, valid after the phase-2 MMAs
    float* const xk = reinterpret_cast<float*>(sm.rp);
    uint8_t* const o_gr = sm.qm;    // staging tiles alias operand tiles that are dead after the phase-2 MMAs; the next
    uint8_t* const o_gk = sm.qmt;   // chunk rewrites them only after thread 0 has waited for the TMA stores
    uint8_t* const o_gv = sm.pt;

    if (p.dbg && tid == 0 && blockIdx.x < 1024) {   // debug trace only: per-CTA {start, SM id, end, 0}
        unsigned long long t;
        uint32_t smid;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        p.dbg[128 + 4 * blockIdx.x] = t;
        p.dbg[129 + 4 * blockIdx.x] = smid;
    }
    // Programmatic dependent launch: the next kernel of the stream may become resident from here on (its prologue overlaps
    // this kernel's tail).  Triggering later, e.g. at the last chunk, was tried: the instruction inside the chunk loop costs
    // 50 R2UR / 120 instructions of code generation and 2.5 us per launch.
    pdl_launch_dependents();
    if (tid == 0) {
        mbar_init(&sm.bar_rkw, 1);
        mbar_init(&sm.bar_vg[0], 1);
        mbar_init(&sm.bar_vg[1], 1);
        mbar_init(&sm.bar_sp, 1);
        mbar_init(&sm.bar_mp, 4);     // MMA groups are issued by several warps' lane 0 in parallel (an issue costs
        mbar_init(&sm.bar_m1, 1);     // ~50 cycles of dependent uniform-datapath work); every issuer commits (bar_m1: Q, warp 0)
        mbar_init(&sm.bar_m2, 4);       // R^T / K^T groups: four issuing lanes
        mbar_init(&sm.bar_m2b, 2);      // GV, dC
        fence_barrier_init();
        tma_prefetch_desc(&D.mr); tma_prefetch_desc(&D.mk); tma_prefetch_desc(&D.mv); tma_prefetch_desc(&D.mw);
        tma_prefetch_desc(&D.mg); tma_prefetch_desc(&D.mc);
        tma_prefetch_desc(&D.mgr); tma_prefetch_desc(&D.mgk); tma_prefetch_desc(&D.mgv); tma_prefetch_desc(&D.mgw);
        // the first chunk's loads go out here, before the rest of the prologue (TMEM allocation, shared-memory clears,
        // reverse-state load), not after it
        fence_proxy_async();
        pdl_wait();
        if (nunits > 0) {
            const int q0 = rev ? cu0.c : nchunks - 1 - cu0.c;   // tile of the chunk this CTA's reverse sweep starts with
            mbar_arrive_expect_tx(&sm.bar_rkw, 3u * CL * 128);
            tma_load_3d(sm.raw_r, &D.mr, &sm.bar_rkw, cu0.h * N, q0 * CL, cu0.b);
            tma_load_3d(sm.raw_k, &D.mk, &sm.bar_rkw, cu0.h * N, q0 * CL, cu0.b);
            tma_load_3d(sm.raw_w, &D.mw, &sm.bar_rkw, cu0.h * N, q0 * CL, cu0.b);
            mbar_arrive_expect_tx(&sm.bar_vg[0], 2u * CL * 128);
            tma_load_3d(sm.v[0], &D.mv, &sm.bar_vg[0], cu0.h * N, q0 * CL, cu0.b);
            tma_load_3d(sm.g[0], &D.mg, &sm.bar_vg[0], cu0.h * N, q0 * CL, cu0.b);
            mbar_arrive_expect_tx(&sm.bar_sp, (uint32_t)N * 128);
            tma_load_2d(sm.sp, &D.mc, &sm.bar_sp, 0, (cu0.bh * nchunks + (nchunks - 1 - cu0.c)) * N);
        }
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&sm.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = tid; i < 28 * 1024 / 16; i += kBwdThreads) reinterpret_cast<uint4*>(sm.kt)[i] = make_uint4(0, 0, 0, 0);
    for (int i = tid; i < 20 * 1024 / 16; i += kBwdThreads) reinterpret_cast<uint4*>(sm.rp)[i] = make_uint4(0, 0, 0, 0);

    // ---- fragment coordinates (TMEM 16x256b.x2): warp = (lane quarter qtr, column quarter cqt): one 16x16 block of an
    // accumulator; rows frow, frow+8; columns fcol + 8*rep + {0,1}, rep = 0..1
    const int qtr = warp & 3, cqt = warp >> 2;
    const int frow = 16 * qtr + (lane >> 2);
    const int fcol = 16 * cqt + 2 * (lane & 3);
    const uint32_t lane_field = (uint32_t)(qtr * 32) << 16;
    // stmatrix row address of this lane for the 16x16 fragment (one x4 store; matrix k = lane>>3 is rep = k>>1,
    // hh = k&1): stoff = the fragment as it is, ttoff = its transpose (rows = fragment columns)
    uint32_t stoff, ttoff;
    {
        const int k = lane >> 3, r8 = lane & 7;
        stoff = (uint32_t)sw_chunk(16 * qtr + 8 * (k & 1) + r8, 2 * cqt + (k >> 1));
        ttoff = (uint32_t)sw_chunk(16 * cqt + 8 * (k >> 1) + r8, 2 * qtr + (k & 1));
    }
    const int sgn = rev ? -1 : 1;
    const int dfrag = sgn * ((lane >> 2) - 2 * (lane & 3));   // sgn * (row - column) of fragment element (rep 0, hh 0, e 0)

    // ---- row-major coordinates (preprocessing / final epilogue): sub-block I (direction order), 4 rows x 2 channels
    const int I = warp & 3, ch = warp >> 2;
    const int cq = lane & 7, rg = lane >> 3;
    const int c0 = 16 * ch + 2 * cq;
    const int coff = (c0 & 7) * 2, cchunk = c0 >> 3;
    const int row0 = rev ? 63 - (16 * I + 4 * rg) : 16 * I + 4 * rg;   // the thread's rows: row0 + sgn * r, r = 0..3
    // Loop-invariant shared-memory offsets.  offs: the thread's four (row, channel pair) words of a row-major tile;
    // fo[2 * rep + hh]: word (row frow + 8 hh, columns fcol + 8 rep ..+1) of a fragment-layout tile; xbase: its fp32
    // transposition-buffer slot.  (ptxas rematerialises some of this swizzle arithmetic inside the chunk loop instead of
    // keeping it in registers; pinning the values with empty asm statements was measured 1.2 % SLOWER.)
    int offs[4];
    uint32_t fo[4];
#pragma unroll
    for (int r = 0; r < 4; r++) {
        offs[r] = sw_chunk(row0 + sgn * r, cchunk) + coff;
    }
#pragma unroll
    for (int rep = 0; rep < 2; rep++)
#pragma unroll
        for (int hh = 0; hh < 2; hh++) {
            const int col = fcol + 8 * rep;
            fo[2 * rep + hh] = (uint32_t)(sw_chunk(frow + 8 * hh, col >> 3) + (col & 7) * 2);
        }
    const int xbase = fcol * XSTR + frow;
    const int vbK_own = I == 0 ? vbK[0] : I == 1 ? vbK[1] : I == 2 ? vbK[2] : vbK[3];

    pdl_wait();   // first global access below (see wkv6_common.cuh)
    float u4[2];
    float gu4[2];
    // reverse state C^T[j][i]: j = frow + 8*hh, i = fcol + 8*rep + e  ->  cs[4*rep + 2*hh + e]
    float cs[8];
    auto store_cp = [&]() {
#pragma unroll
        for (int rep = 0; rep < 2; rep++)
#pragma unroll
            for (int hh = 0; hh < 2; hh++)
                *reinterpret_cast<uint32_t*>(sm.cp + fo[2 * rep + hh]) = pack_bf16(cs[4 * rep + 2 * hh], cs[4 * rep + 2 * hh + 1]);
    };
    // start of a piece: the reverse state (gradient into S_T / what the segment scan left in the slot / the record the
    // previous CTA handed over), u of the head, gu = 0, bf16 copy of the state for the first chunk's MMAs
    auto piece_start = [&](const Unit& u) {
        const float* g0;
        if (u.stage == kSEG) g0 = p.seg_state + ((size_t)seq * nseg + seg) * N * N;
        else if (u.stage == kB) {
            if (tid == 0) {
                const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(rec_in + (size_t)N * N * 4 + 256);
                unsigned long long v = 0;
                long long spins = 0;
                for (;;) {
                    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
                    if (v == p.bal_epoch) break;
                    if (++spins > (1ll << 25)) __trap();          // seconds: the producer died
                    __nanosleep(100);
                }
            }
            __syncthreads();
            g0 = reinterpret_cast<const float*>(rec_in);
        } else g0 = D.gsT ? D.gsT + (size_t)u.bh * N * N : nullptr;
#pragma unroll
        for (int rep = 0; rep < 2; rep++)
#pragma unroll
            for (int hh = 0; hh < 2; hh++)
#pragma unroll
                for (int e = 0; e < 2; e++)
                    cs[4 * rep + 2 * hh + e] = g0 ? __ldcg(g0 + (fcol + 8 * rep + e) * N + frow + 8 * hh) : 0.f;
        up2bf(*reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(D.u) + (u.h * N + c0) * 2), u4);
        gu4[0] = gu4[1] = 0.f;
        store_cp();
        fence_proxy_async();
    };
    // end of a piece: gu (final, per-segment partial, or partial into the hand-over record), the reverse state (gs0 when the
    // sweep reached the sequence's first step, or the hand-over record followed by its flag)
    auto piece_end = [&](const Unit& u) {
        // gu[b][h*64 + c] = sum over rows of r k (v.gy): row groups by shuffle, sub-blocks through smem
#pragma unroll
        for (int e = 0; e < 2; e++) {
            gu4[e] += __shfl_xor_sync(0xffffffffu, gu4[e], 8);
            gu4[e] += __shfl_xor_sync(0xffffffffu, gu4[e], 16);
        }
        if (rg == 0) {
#pragma unroll
            for (int e = 0; e < 2; e++) sm.gured[warp][2 * cq + e] = gu4[e];
        }
        float* sout = nullptr;
        if (u.stage == kA) sout = reinterpret_cast<float*>(rec_out);
        else if (D.gs0 && (u.stage != kSEG || seg == nseg - 1)) sout = D.gs0 + (size_t)u.bh * N * N;   // the sequence's first step
        if (sout) {
#pragma unroll
            for (int rep = 0; rep < 2; rep++)
#pragma unroll
                for (int hh = 0; hh < 2; hh++)
#pragma unroll
                    for (int e = 0; e < 2; e++) sout[(fcol + 8 * rep + e) * N + frow + 8 * hh] = cs[4 * rep + 2 * hh + e];
        }
        __syncthreads();
        if (tid < N) {
            const int hc = tid >> 4, cc = tid & 15;   // warp = sub-block + 4 * (16-channel quarter)
            float g = sm.gured[4 * hc + 0][cc] + sm.gured[4 * hc + 1][cc] + sm.gured[4 * hc + 2][cc] + sm.gured[4 * hc + 3][cc];
            if (u.stage == kB) g += __ldcg(reinterpret_cast<const float*>(rec_in + (size_t)N * N * 4) + tid);
            if (u.stage == kA) reinterpret_cast<float*>(rec_out + (size_t)N * N * 4)[tid] = g;
            else if (u.stage == kSEG) p.seg_gu[((size_t)seq * nseg + seg) * N + tid] = g;      // summed by wkv6_bwd_seg_gu_kernel
            else reinterpret_cast<__nv_bfloat16*>(D.gu)[(size_t)u.b * p.C + u.h * N + tid] = __float2bfloat16_rn(g);
        }
        // (the barrier is unconditional: a CTA barrier under the stage test cost ptxas the uniformity of the whole chunk loop)
        if (u.stage == kA) __threadfence();
        __syncthreads();
        if (tid == 0) {
            if (u.stage == kA) {
                unsigned long long* flag = reinterpret_cast<unsigned long long*>(rec_out + (size_t)N * N * 4 + 256);
                asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(flag), "l"(p.bal_epoch) : "memory");
            } else if (u.stage == kB) {
                // consumed: the flag goes back to 0, so that a replay of the same launch (CUDA graph: same epoch) waits again
                *reinterpret_cast<volatile unsigned long long*>(rec_in + (size_t)N * N * 4 + 256) = 0ull;
            }
        }
    };
    if (nunits > 0) piece_start(cu0);

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tbase = __shfl_sync(0xffffffffu, sm.tmem_base, 0);
    const uint32_t tQ = tbase, tP = tbase + 64, tRS = tbase + 128, tKC = tbase + 192, tR = tbase + 256, tK = tbase + 320,
                   tGV = tbase + 384, tDC = tbase + 448;

    // UMMA smem descriptors of every operand tile, built once: inside the chunk loop a descriptor is base + (byte
    // offset >> 4), one 64-bit add on the uniform datapath instead of a dozen dependent uniform ops per MMA.
    const uint64_t dk_g0 = desc_k(smem_u32(sm.g[0])), dk_v0 = desc_k(smem_u32(sm.v[0])), dk_sp = desc_k(smem_u32(sm.sp)),
                   dk_cp = desc_k(smem_u32(sm.cp)), dk_kt = desc_k(smem_u32(sm.kt)), dk_roff = desc_k(smem_u32(sm.roff)),
                   dk_qm = desc_k(smem_u32(sm.qm)), dk_qmt = desc_k(smem_u32(sm.qmt)), dk_pt = desc_k(smem_u32(sm.pt)),
                   dk_kend = desc_k(smem_u32(kend));
    const uint64_t dm_kt = desc_mn(smem_u32(sm.kt)), dm_ktl = desc_mn(smem_u32(sm.kt_lo)), dm_rp = desc_mn(smem_u32(sm.rp)),
                   dm_rpl = desc_mn(smem_u32(sm.rp_lo)), dm_g0 = desc_mn(smem_u32(sm.g[0])), dm_cp = desc_mn(smem_u32(sm.cp)),
                   dm_rst = desc_mn(smem_u32(sm.rst)), dm_rstl = desc_mn(smem_u32(sm.rst_lo));
#define DOFF(bytes) ((uint64_t)((bytes) >> 4))

    // Loads of the next chunk are issued as soon as their landing tiles are dead: r/k/w after the operand tiles are
    // built (r, k live on in registers), v/gy after the phase-2 MMAs, the checkpoint after the anchor sum.
    // (cb = local chunk counter of this CTA, 0 .. nloc - 1; cbeg + cb = chunk in the sequence's processing order)
    // (it = chunk counter of this CTA, 0 .. nunits - 1; Unit::c = chunk in its sequence's processing order)
    auto tile_q = [&](const Unit& u) { return rev ? u.c : nchunks - 1 - u.c; };
    auto issue_rkw = [&](const Unit& u) {   // thread 0 only
        const int q = tile_q(u);
        mbar_arrive_expect_tx(&sm.bar_rkw, 3u * CL * 128);
        tma_load_3d(sm.raw_r, &D.mr, &sm.bar_rkw, u.h * N, q * CL, u.b);
        tma_load_3d(sm.raw_k, &D.mk, &sm.bar_rkw, u.h * N, q * CL, u.b);
        tma_load_3d(sm.raw_w, &D.mw, &sm.bar_rkw, u.h * N, q * CL, u.b);
    };
    auto issue_vg = [&](const Unit& u, int buf) {
        const int q = tile_q(u);
        mbar_arrive_expect_tx(&sm.bar_vg[buf], 2u * CL * 128);
        tma_load_3d(sm.v[buf], &D.mv, &sm.bar_vg[buf], u.h * N, q * CL, u.b);
        tma_load_3d(sm.g[buf], &D.mg, &sm.bar_vg[buf], u.h * N, q * CL, u.b);
    };
    auto issue_sp = [&](const Unit& u) {
        mbar_arrive_expect_tx(&sm.bar_sp, (uint32_t)N * 128);
        tma_load_2d(sm.sp, &D.mc, &sm.bar_sp, 0, (u.bh * nchunks + (nchunks - 1 - u.c)) * N);
    };
    // (the loads of the first chunk were issued at the top of the prologue)

    constexpr uint32_t ID_KK64 = umma_idesc(0, 0, 64, 64);
    constexpr uint32_t ID_KK16 = umma_idesc(0, 0, 64, 16);
    constexpr uint32_t ID_MK16 = umma_idesc(1, 0, 64, 16);
    constexpr uint32_t ID_KM64 = umma_idesc(0, 1, 64, 64);
    constexpr uint32_t ID_MM64 = umma_idesc(1, 1, 64, 64);

    for (int cb = 0; cb < nunits; cb++) {
        const bool more = cb + 1 < nunits;
        // what every warp needs of the schedule: the chunk's position in its sequence and whether its piece ends here; the
        // full units (batch row, head: two more reciprocal divisions each) are computed where they are used -- by the lanes
        // that issue TMA and at the piece boundaries
        const int xw = cb - lenA;                        // position in the run of whole sequences
        const int c_cur = xw < 0 ? cb : xw < lenW ? xw - udiv(xw, p.bal_magic_n) * nchunks : cB0 + (xw - lenW);
        const bool pend = !more || cb + 1 == lenA || (xw >= 0 && xw < lenW && c_cur == nchunks - 1);   // (uniform)
        const int t0 = (rev ? c_cur : nchunks - 1 - c_cur) * CL;   // first step of the chunk's tile in memory
        const uint32_t par = cb & 1;
        const uint64_t vgoff = (cb & 1) ? DOFF(CL * 128) : 0;   // V / G buffer of this chunk (v[1] - v[0] = g[1] - g[0] = 8 KB)
        const uint64_t dk_g = dk_g0 + vgoff, dk_v = dk_v0 + vgoff, dm_g = dm_g0 + vgoff;
        const int c_ts = cb;
#if WKV6_TRACE_PHASES
        if (wkv6_trace_on && tid == 0 && cb < 32) p.dbg[96 + cb] = clock64();   // loop-top stamp of every chunk (debug trace)
#endif
        WKV6_TS(64, 0);
        // staging tiles of the previous chunk have been read (they alias operand tiles that are rewritten after the next
        // CTA-wide barrier).  The store thread is warp 7 lane 0: it issues no MMAs, so this wait delays nobody.
        if (tid == 224) tma_store_wait_read<0>();
        if (tid == 0 && more) issue_vg(unit_at(cb + 1), (cb + 1) & 1);   // other V / G buffer: its readers finished last iteration
        if (warp == 0 && elect_one()) {
            // ---- MMA phase 1a: Q needs no preprocessing (V, G are raw tiles).  Issue only: a tcgen05.commit right behind the
            // MMAs parks the issuing thread until the tensor pipe has drained them (measured: 430 cycles for this group, by which
            // the issuing warps reached barrier B1 late); the commit follows after the R-side preprocessing, when it is free.
            mbar_wait(&sm.bar_vg[cb & 1], (cb >> 1) & 1);
            tc_fence_after();
#pragma unroll
            for (int ks = 0; ks < 4; ks++) umma(tQ, dk_g + DOFF(ks * 32), dk_v + DOFF(ks * 32), ID_KK64, ks > 0);
        }
        __syncwarp();

        // ================= phase 1: decays, in-block cumulative sums, Rtil, Ktil(own), bonus
        mbar_wait(&sm.bar_rkw, par);
        WKV6_TS(64, 1);
        float eo[4][2], ei[4][2], ew[4][2], ebt[2];           // ebt = 2^(log2 decay total of the own sub-block)
        uint32_t rpk[4], kpk[4];                         // this thread's r and k (bf16 x2 per row) for the later phases
        {
            float rr[4][2], kk[4][2], l2[4][2], dstep[4][2], bsum[4];
            float xmax = 0.f;
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int row = row0 + sgn * r;
                const int off = offs[r];
                float w4[2];
                rpk[r] = *reinterpret_cast<const uint32_t*>(sm.raw_r + off);
                kpk[r] = *reinterpret_cast<const uint32_t*>(sm.raw_k + off);
                up2bf(rpk[r], rr[r]);
                up2bf(kpk[r], kk[r]);
                up2bf(*reinterpret_cast<const uint32_t*>(sm.raw_w + off), w4);
                float bs = 0.f;
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    bs = fmaf(rr[r][e] * u4[e], kk[r][e], bs);
                    const float x = ex2(w4[e] * kLog2e);            // exp(w)
                    ew[r][e] = x;
                    xmax = fmaxf(xmax, x);
                    l2[r][e] = fmaxf(-x * kLog2e, -kClampL2);
                }
                bsum[r] = bs;
            }
            if (t0 + CL > T) {        // ragged last tile of the sequence (uniform): steps at or beyond T do not decay
#pragma unroll
                for (int r = 0; r < 4; r++)
                    if (t0 + row0 + sgn * r >= T) l2[r][0] = l2[r][1] = 0.f;
            }
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int e = 0; e < 2; e++) dstep[r][e] = ex2(l2[r][e]);   // the step's decay
            // d l / d w where the decay floor is active (exp(w) > 5; never with trained decays, so the whole warp skips it):
            // the recurrence ran with 2^l instead of 2^lt, so the H_p computed below is (floored decay) x dL/d(decay); the
            // true step has (true decay) x the same derivative to first order: rescale exp(w) by 2^(lt - l)
            if (__any_sync(0xffffffffu, xmax * kLog2e > kClampL2)) {
#pragma unroll
                for (int r = 0; r < 4; r++)
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        const float x = fminf(ew[r][e], 1e30f);
                        ew[r][e] = x * ex2(fminf(kClampL2 - x * kLog2e, 0.f));
                    }
            }
#pragma unroll
            for (int e = 0; e < 2; e++) {
                l2[1][e] += l2[0][e];
                l2[2][e] += l2[1][e];
                l2[3][e] += l2[2][e];
            }
            float excl[2];
#pragma unroll
            for (int e = 0; e < 2; e++) {
                float x = l2[3][e];
                float y = __shfl_up_sync(0xffffffffu, x, 8);
                if (rg >= 1) x += y;
                y = __shfl_up_sync(0xffffffffu, x, 16);
                if (rg >= 2) x += y;
                excl[e] = x - l2[3][e];
                if (rg == 3) sm.blocktot[I][c0 + e] = x;
            }
            // eo[r] = 2^(A_p - a_I), ei[r] = 2^-(A+_p - a_I) of the thread's four consecutive steps: one ex2 at each end of the run
            // and products of the step decays in between (the MUFU pipe, 16 lanes per clock per SM, bounds this phase: 6 ex2
            // per channel instead of 8)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                eo[0][e] = ex2(excl[e]);
                eo[1][e] = eo[0][e] * dstep[0][e];
                eo[2][e] = eo[1][e] * dstep[1][e];
                eo[3][e] = eo[2][e] * dstep[2][e];
                ei[3][e] = ex2(-(excl[e] + l2[3][e]));
                ei[2][e] = ei[3][e] * dstep[3][e];
                ei[1][e] = ei[2][e] * dstep[2][e];
                ei[0][e] = ei[1][e] * dstep[1][e];
            }
#pragma unroll
            for (int r = 0; r < 4; r++) {
                float ro[2], ko[2];
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    ro[e] = rr[r][e] * eo[r][e];
                    ko[e] = kk[r][e] * ei[r][e];
                }
                st2bf(sm.roff + offs[r], ro);
                st2bf_hilo(sm.kt + vbK_own + offs[r], sm.kt_lo + vbK_own + offs[r], ko);
                float bs = bsum[r];
                bs += __shfl_xor_sync(0xffffffffu, bs, 1);
                bs += __shfl_xor_sync(0xffffffffu, bs, 2);
                bs += __shfl_xor_sync(0xffffffffu, bs, 4);
                if (cq == 0) sm.bonus[ch][row0 + sgn * r] = bs;
            }
        }
        WKV6_TS(64, 2);
        __syncthreads();   // B1: blocktot / bonus visible

        // ================= phase 1b: cross-block factors -> Rst, Ktil(I' > I), Kend, Rtil'(J <= I)
        {
            // Six factors per channel: fst = 2^(a_I);  fk[m] = 2^(a_{m+1} - a_I), m >= I (m = 3: chunk end);  fr[J] = 2^(a_I - a_{J+1}),
            // J <= I.  The four row groups of a warp work on the same 16 channels: lane l computes three of the six of channel
            // 16 ch + (l & 15) and the warp exchanges them through 384 bytes of the gw staging tile (idle until the final
            // epilogue; its TMA store was waited for at the top of the chunk): 3 ex2 per lane instead of 12 -- the MUFU pipe
            // bounds this phase.  Slots [6][16 channels]: m < I: fr[m];  m >= I: fk[m];  4: fr[I];  5: fst.
            float* const fac = reinterpret_cast<float*>(sm.o_gw) + warp * 96;
            {
                const int cw = lane & 15, part = lane >> 4;
                float tot[4];
#pragma unroll
                for (int m = 0; m < 4; m++) tot[m] = sm.blocktot[m][16 * ch + cw];
                // slot m < 4: m >= I: sum_{x = I..m} tot[x];  m < I: sum_{x = m+1..I-1} tot[x]
                auto slot_exp = [&](int m) {
                    float a = 0.f;
#pragma unroll
                    for (int x = 0; x < 4; x++)
                        if (m >= I ? (x >= I && x <= m) : (x > m && x < I)) a += tot[x];
                    return a;
                };
                float pre = 0.f, totI = tot[0];
#pragma unroll
                for (int m = 0; m < 4; m++) {
                    if (m < I) pre += tot[m];
                    if (m == I) totI = tot[m];
                }
                float* const f = fac + cw;
                if (part == 0) {
                    f[0 * 16] = ex2(slot_exp(0));
                    f[1 * 16] = ex2(slot_exp(1));
                    f[4 * 16] = ex2(-totI);
                } else {
                    f[2 * 16] = ex2(slot_exp(2));
                    f[3 * 16] = ex2(slot_exp(3));
                    f[5 * 16] = ex2(pre);
                }
            }
            __syncwarp();
            auto ldfac = [&](int slot, float (&v)[2]) {
                const float2 t = *reinterpret_cast<const float2*>(fac + slot * 16 + 2 * cq);
                v[0] = t.x; v[1] = t.y;
            };
            ldfac(I, ebt);                      // 2^(total of the own sub-block)
            {                                   // K side: needed by the P^T MMAs
                float ko[4][2];
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    up2bf(kpk[r], ko[r]);
#pragma unroll
                    for (int e = 0; e < 2; e++) ko[r][e] *= ei[r][e];
                }
#pragma unroll
                for (int m = 0; m < 4; m++) {
                    if (m >= I) {               // target variant I' = m + 1 (I' = 4 -> Kend)
                        float f[2], x4[2];
                        ldfac(m, f);
#pragma unroll
                        for (int r = 0; r < 4; r++) {
#pragma unroll
                            for (int e = 0; e < 2; e++) x4[e] = ko[r][e] * f[e];
                            // low-order halves only where the gw cancellation needs them: variant I' = I + 1 (and the own
                            // variant, phase 1); farther variants are damped by at least one whole sub-block of decay
                            if (m == 3) st2bf(kend + offs[r], x4);
                            else if (kAdjLo && m == I) st2bf_hilo(sm.kt + vbK[m + 1] + offs[r], sm.kt_lo + vbK[m + 1] + offs[r], x4);
                            else st2bf(sm.kt + vbK[m + 1] + offs[r], x4);
                        }
                    }
                }
            }
            fence_proxy_async();
            tc_fence_before();
            WKV6_TS(64, 3);
            __syncthreads();   // B2: Rtil, Ktil variants complete; raw r / k / w tiles are dead
            WKV6_TS(64, 4);

            // ================= MMA phase 1b: P^T (runs under the R-side preprocessing below)
            if (warp < 4 && elect_one()) {
                tc_fence_after();
                const int i = warp;
                const int rowlo = rev ? 48 - 16 * i : 16 * i;
                const int vbi = i == 0 ? vbK[0] : i == 1 ? vbK[1] : i == 2 ? vbK[2] : vbK[3];
#pragma unroll
                for (int ks = 0; ks < 4; ks++)
                    umma(tP + rowlo, dk_kt + DOFF(vbi + ks * 32), dk_roff + DOFF(rowlo * 128 + ks * 32), ID_KK16, ks > 0);
                // RS^T = S_0^T G^T and KC^T = C^T V^T (raw / carried tiles) are read by E2 only: they queue up behind P^T and are
                // covered by their issuers' phase-2 commits (a commit tracks every earlier MMA of its thread)
                if (warp == 1) {
                    mbar_wait(&sm.bar_sp, par);
                    tc_fence_after();
#pragma unroll
                    for (int ks = 0; ks < 4; ks++) umma(tRS, dk_sp + DOFF(ks * 32), dk_g + DOFF(ks * 32), ID_KK64, ks > 0);
                } else if (warp == 2) {
#pragma unroll
                    for (int ks = 0; ks < 4; ks++) umma(tKC, dk_cp + DOFF(ks * 32), dk_v + DOFF(ks * 32), ID_KK64, ks > 0);
                }
            }
            __syncwarp();
            if (tid == 0 && more) issue_rkw(unit_at(cb + 1));
            {                                   // R side: needed by the phase-2 MMAs only
                float ro[4][2], fst[2];
                ldfac(5, fst);
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    float x4[2];
                    up2bf(rpk[r], ro[r]);
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        ro[r][e] *= eo[r][e];
                        x4[e] = ro[r][e] * fst[e];
                    }
                    if (kRstLo) st2bf_hilo(sm.rst + offs[r], sm.rst_lo + offs[r], x4);
                    else st2bf(sm.rst + offs[r], x4);
                }
#pragma unroll
                for (int J = 0; J < 4; J++) {
                    if (J <= I) {
                        float f[2], x4[2];
                        ldfac(J == I ? 4 : J, f);
#pragma unroll
                        for (int r = 0; r < 4; r++) {
#pragma unroll
                            for (int e = 0; e < 2; e++) x4[e] = ro[r][e] * f[e];
                            if (kAdjLo ? J + 1 >= I : J == I) st2bf_hilo(sm.rp + vbR[J] + offs[r], sm.rp_lo + vbR[J] + offs[r], x4);   // own / previous sub-block
                            else st2bf(sm.rp + vbR[J] + offs[r], x4);
                        }
                    }
                }
            }
        }

        if (warp < 4 && elect_one()) {            // (the same lanes that issued)
            umma_commit(&sm.bar_mp);
            if (warp == 0) umma_commit(&sm.bar_m1);
        }
        __syncwarp();
        WKV6_TS(64, 5);
        // ================= E1: Q -> masked Q, Q^T, vg;  P^T -> masked P^T with the bonus diagonal
        mbar_wait(&sm.bar_m1, par);
        WKV6_TS(64, 6);
        tc_fence_after();
        {
            // The causal mask is resolved per 16x16 block (warp-uniform: keep / zero / diagonal); the masked fragment goes
            // to the tiles with stmatrix (Q^T with the transposing form) instead of per-element stores.
            float x[8];
            uint32_t pk[4];
            tmem_ld_16x16(tQ + lane_field + 16 * cqt, x);   // Q[p][q]: rows p (block qtr), columns q (block cqt); keep q before p
            {
                const int cls = sgn * (qtr - cqt);
                if (cls > 0) {
#pragma unroll
                    for (int i = 0; i < 4; i++) pk[i] = pack_bf16(x[2 * i], x[2 * i + 1]);
                } else if (cls < 0) {
#pragma unroll
                    for (int i = 0; i < 4; i++) pk[i] = 0u;
                } else {
#pragma unroll
                    for (int rl = 0; rl < 2; rl++)
#pragma unroll
                        for (int hh = 0; hh < 2; hh++) {
                            float y[2];
#pragma unroll
                            for (int e = 0; e < 2; e++) {
                                const int d = dfrag + sgn * (8 * hh - 8 * rl - e);   // sgn * (p - q)
                                const float xv = x[4 * rl + 2 * hh + e];
                                y[e] = d > 0 ? xv : 0.f;
                                if (d == 0) sm.vg[frow + 8 * hh] = xv;
                            }
                            pk[2 * rl + hh] = pack_bf16(y[0], y[1]);
                        }
                }
            }
            stsm_x4(smem_u32(sm.qm) + stoff, pk[0], pk[1], pk[2], pk[3]);
            stsm_x4_trans(smem_u32(sm.qmt) + ttoff, pk[0], pk[1], pk[2], pk[3]);
            mbar_wait(&sm.bar_mp, par);
            WKV6_TS(64, 7);
            tc_fence_after();
            tmem_ld_16x16(tP + lane_field + 16 * cqt, x);   // P^T[q][p]: rows q (block qtr), columns p (block cqt); keep q before p
            {
                const int cls = sgn * (cqt - qtr);
                if (cls > 0) {
#pragma unroll
                    for (int i = 0; i < 4; i++) pk[i] = pack_bf16(x[2 * i], x[2 * i + 1]);
                } else if (cls < 0) {
#pragma unroll
                    for (int i = 0; i < 4; i++) pk[i] = 0u;
                } else {
                    float bon[2];
#pragma unroll
                    for (int hh = 0; hh < 2; hh++)
                        bon[hh] = (sm.bonus[0][frow + 8 * hh] + sm.bonus[1][frow + 8 * hh]) + (sm.bonus[2][frow + 8 * hh] + sm.bonus[3][frow + 8 * hh]);
#pragma unroll
                    for (int rl = 0; rl < 2; rl++)
#pragma unroll
                        for (int hh = 0; hh < 2; hh++) {
                            float y[2];
#pragma unroll
                            for (int e = 0; e < 2; e++) {
                                const int d = -dfrag + sgn * (8 * rl + e - 8 * hh);   // sgn * (p - q)
                                y[e] = d > 0 ? x[4 * rl + 2 * hh + e] : (d == 0 ? bon[hh] : 0.f);
                            }
                            pk[2 * rl + hh] = pack_bf16(y[0], y[1]);
                        }
                }
            }
            stsm_x4(smem_u32(sm.pt) + stoff, pk[0], pk[1], pk[2], pk[3]);
        }
        fence_proxy_async();
        tc_fence_before();
        WKV6_TS(64, 8);
        __syncthreads();   // B3: masked tiles complete
        WKV6_TS(64, 9);

        // ================= MMA phase 2: four issuing lanes.  Warp w issues the R^T and K^T groups w (10 MMAs) and commits
        // them to bar_m2; warps 0 / 1 then add GV / dC (8 MMAs each, bar_m2b): in the tensor-pipe queue those come after
        // the R^T / K^T groups, whose drain (the long part of E2) then overlaps them.
        if (warp < 4 && elect_one()) {
            tc_fence_after();
            {   // R^T[j][p in sub-block i] = sum over q in sub-blocks <= i
                const int i = warp;
                const int rb = rev ? 3 - i : i;
                const int vbi = i == 0 ? vbK[0] : i == 1 ? vbK[1] : i == 2 ? vbK[2] : vbK[3];
                for (int s_ = 0; s_ <= i; s_++) {
                    const int kb = rev ? 3 - s_ : s_;
                    umma(tR + 16 * rb, dm_kt + DOFF(vbi + kb * 2048), dk_qm + DOFF(rb * 2048 + kb * 32), ID_MK16, s_ > 0);
                    if (kAdjLo ? s_ + 1 >= i : s_ == i)   // low-order halves exist for the diagonal and the adjacent sub-block only
                        umma(tR + 16 * rb, dm_ktl + DOFF(vbi + kb * 2048), dk_qm + DOFF(rb * 2048 + kb * 32), ID_MK16, 1);
                }
            }
            {   // K^T[j][q in sub-block j] = sum over p in sub-blocks >= j
                const int j = warp;
                const int rb = rev ? 3 - j : j;
                const int vbj = j == 0 ? vbR[0] : j == 1 ? vbR[1] : j == 2 ? vbR[2] : vbR[3];
                for (int s_ = j; s_ < 4; s_++) {
                    const int kb = rev ? 3 - s_ : s_;
                    umma(tK + 16 * rb, dm_rp + DOFF(vbj + kb * 2048), dk_qmt + DOFF(rb * 2048 + kb * 32), ID_MK16, s_ > j);
                    if (kAdjLo ? s_ <= j + 1 : s_ == j)
                        umma(tK + 16 * rb, dm_rpl + DOFF(vbj + kb * 2048), dk_qmt + DOFF(rb * 2048 + kb * 32), ID_MK16, 1);
                }
            }
            umma_commit(&sm.bar_m2);
            if (warp == 0) {
#pragma unroll
                for (int ks = 0; ks < 4; ks++) umma(tGV, dk_pt + DOFF(ks * 32), dm_g + DOFF(ks * 2048), ID_KM64, ks > 0);
#pragma unroll
                for (int ks = 0; ks < 4; ks++) umma(tGV, dk_kend + DOFF(ks * 32), dm_cp + DOFF(ks * 2048), ID_KM64, 1);
                umma_commit(&sm.bar_m2b);
            } else if (warp == 1) {
#pragma unroll
                for (int ks = 0; ks < 4; ks++) umma(tDC, dm_rst + DOFF(ks * 2048), dm_g + DOFF(ks * 2048), ID_MM64, ks > 0);
                if (kRstLo) {
#pragma unroll
                    for (int ks = 0; ks < 4; ks++) umma(tDC, dm_rstl + DOFF(ks * 2048), dm_g + DOFF(ks * 2048), ID_MM64, 1);
                }
                umma_commit(&sm.bar_m2b);
            }
        }
        __syncwarp();

        // ================= E2: drain the phase-2 accumulators
        WKV6_TS(64, 10);
        mbar_wait(&sm.bar_m2, par);
        WKV6_TS(64, 11);
        tc_fence_after();
        {
            // per-thread decay factors of its two rows j (log2 sums of the sub-block totals)
            float t0j[4], t1j[4];
#pragma unroll
            for (int m = 0; m < 4; m++) {
                t0j[m] = sm.blocktot[m][frow];
                t1j[m] = sm.blocktot[m][frow + 8];
            }
            // columns of this thread lie in row-block cqt -> sub-block (direction order) iA
            const int iA = rev ? 3 - cqt : cqt;
            float sR[2], sK[2];   // [row hh]: 2^(a_I) and 2^(A_L - a_{I+1})
#pragma unroll
            for (int hh = 0; hh < 2; hh++) {
                const float* tj = hh ? t1j : t0j;
                float preA = 0.f, sufA = 0.f;
#pragma unroll
                for (int m = 0; m < 4; m++) {
                    if (m < iA) preA += tj[m];
                    if (m > iA) sufA += tj[m];
                }
                sR[hh] = ex2(preA);
                sK[hh] = ex2(sufA);
            }
            float x[8], y[8];
            tmem_ld_16x16(tR + lane_field + 16 * cqt, x);
            tmem_ld_16x16(tRS + lane_field + 16 * cqt, y);
#pragma unroll
            for (int rep = 0; rep < 2; rep++)
#pragma unroll
                for (int hh = 0; hh < 2; hh++)
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        xr[xbase + (8 * rep + e) * XSTR + 8 * hh] = fmaf(sR[hh], y[4 * rep + 2 * hh + e], x[4 * rep + 2 * hh + e]);
                    }
            tmem_ld_16x16(tK + lane_field + 16 * cqt, x);
            tmem_ld_16x16(tKC + lane_field + 16 * cqt, y);
#pragma unroll
            for (int rep = 0; rep < 2; rep++)
#pragma unroll
                for (int hh = 0; hh < 2; hh++)
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        xk[xbase + (8 * rep + e) * XSTR + 8 * hh] = fmaf(sK[hh], y[4 * rep + 2 * hh + e], x[4 * rep + 2 * hh + e]);
                    }
            // gv tile [q][i]
            mbar_wait(&sm.bar_m2b, par);
            tc_fence_after();
            tmem_ld_16x16(tGV + lane_field + 16 * cqt, x);
#pragma unroll
            for (int rep = 0; rep < 2; rep++)
#pragma unroll
                for (int hh = 0; hh < 2; hh++)
                    *reinterpret_cast<uint32_t*>(o_gv + fo[2 * rep + hh]) = pack_bf16(x[4 * rep + 2 * hh], x[4 * rep + 2 * hh + 1]);
            // reverse state: C <- C 2^(A_L[j]) + dC;  then the anchor  hst[j] = sum_i S_0[i][j] C_0[i][j]
            tmem_ld_16x16(tDC + lane_field + 16 * cqt, x);
            const float f0 = ex2(t0j[0] + t0j[1] + t0j[2] + t0j[3]);
            const float f1 = ex2(t1j[0] + t1j[1] + t1j[2] + t1j[3]);
            float h0 = 0.f, h1 = 0.f;
#pragma unroll
            for (int rep = 0; rep < 2; rep++) {
                cs[4 * rep + 0] = fmaf(cs[4 * rep + 0], f0, x[4 * rep + 0]);
                cs[4 * rep + 1] = fmaf(cs[4 * rep + 1], f0, x[4 * rep + 1]);
                cs[4 * rep + 2] = fmaf(cs[4 * rep + 2], f1, x[4 * rep + 2]);
                cs[4 * rep + 3] = fmaf(cs[4 * rep + 3], f1, x[4 * rep + 3]);
                const uint32_t s0w = *reinterpret_cast<const uint32_t*>(sm.sp + fo[2 * rep]);
                const uint32_t s1w = *reinterpret_cast<const uint32_t*>(sm.sp + fo[2 * rep + 1]);
                h0 = fmaf(__uint_as_float(s0w << 16), cs[4 * rep + 0], h0);
                h0 = fmaf(__uint_as_float(s0w & 0xffff0000u), cs[4 * rep + 1], h0);
                h1 = fmaf(__uint_as_float(s1w << 16), cs[4 * rep + 2], h1);
                h1 = fmaf(__uint_as_float(s1w & 0xffff0000u), cs[4 * rep + 3], h1);
            }
            h0 += __shfl_xor_sync(0xffffffffu, h0, 1);
            h0 += __shfl_xor_sync(0xffffffffu, h0, 2);
            h1 += __shfl_xor_sync(0xffffffffu, h1, 1);
            h1 += __shfl_xor_sync(0xffffffffu, h1, 2);
            if ((lane & 3) == 0) {
                sm.hst[cqt][frow] = h0;
                sm.hst[cqt][frow + 8] = h1;
            }
            store_cp();
        }
        WKV6_TS(64, 12);
        __syncthreads();   // B4: xr / xk / hst / vg visible; all TMEM reads of this chunk done
        WKV6_TS(64, 13);
        if (tid == 0 && more) issue_sp(unit_at(cb + 1));   // S_0 tile is dead after the anchor sum

        // ================= final epilogue (row-major): gr, gk, gu, gw
        {
            float dl[4][2], av[4][2];   // delta = k gkC - r grS,  a = r grS
            float hs[2];
#pragma unroll
            for (int e = 0; e < 2; e++) hs[e] = (sm.hst[0][c0 + e] + sm.hst[1][c0 + e]) + (sm.hst[2][c0 + e] + sm.hst[3][c0 + e]);
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int row = row0 + sgn * r;
                float rr[2], kk[2], gr4[2], gk4[2];
                up2bf(rpk[r], rr);
                up2bf(kpk[r], kk);
                const float vg = sm.vg[row];
                const float2 a4 = *reinterpret_cast<const float2*>(&xr[row * XSTR + c0]);
                const float2 b4 = *reinterpret_cast<const float2*>(&xk[row * XSTR + c0]);
                const float xa[2] = {a4.x, a4.y}, xb[2] = {b4.x, b4.y};
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const float grS = eo[r][e] * xa[e];
                    const float gkC = ei[r][e] * ebt[e] * xb[e];
                    gr4[e] = fmaf(u4[e] * kk[e], vg, grS);
                    gk4[e] = fmaf(u4[e] * rr[e], vg, gkC);
                    gu4[e] = fmaf(rr[e] * kk[e], vg, gu4[e]);
                    av[r][e] = rr[e] * grS;
                    dl[r][e] = fmaf(kk[e], gkC, -av[r][e]);
                }
                st2bf(o_gr + offs[r], gr4);
                st2bf(o_gk + offs[r], gk4);
            }
            // exclusive prefix of delta over the chunk in direction order: thread (4 rows) -> row groups -> sub-blocks
            float pre[4][2];
#pragma unroll
            for (int e = 0; e < 2; e++) {
                pre[0][e] = 0.f;
                pre[1][e] = dl[0][e];
                pre[2][e] = pre[1][e] + dl[1][e];
                pre[3][e] = pre[2][e] + dl[2][e];
                const float tot = pre[3][e] + dl[3][e];
                float x = tot;
                float y = __shfl_up_sync(0xffffffffu, x, 8);
                if (rg >= 1) x += y;
                y = __shfl_up_sync(0xffffffffu, x, 16);
                if (rg >= 2) x += y;
                const float ex = x - tot;
#pragma unroll
                for (int r = 0; r < 4; r++) pre[r][e] += ex;
                if (rg == 3) sm.dtot[I][c0 + e] = x;
            }
            WKV6_TS(64, 14);
            __syncthreads();   // B5: dtot visible
            float base[2];
#pragma unroll
            for (int e = 0; e < 2; e++) {
                float s = hs[e];
#pragma unroll
                for (int m = 0; m < 4; m++)
                    if (m < I) s += sm.dtot[m][c0 + e];
                base[e] = s;
            }
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int t = t0 + row0 + sgn * r;
                // exact zeros the reference hard-codes (wkv6_cuda.cu:237,262): nothing precedes the direction's first
                // step when S_0 = 0, nothing follows its last step when no gradient enters S_T
                const bool first = !D.s0 && (rev ? t == T - 1 : t == 0);
                const bool last = !D.gsT && (rev ? t == 0 : t == T - 1);
                float gw4[2];
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const float H = base[e] + pre[r][e] - av[r][e];
                    gw4[e] = (first || last) ? 0.f : -ew[r][e] * H;
                }
                st2bf(sm.o_gw + offs[r], gw4);
            }
        }
        fence_proxy_async();
        tc_fence_before();
        WKV6_TS(64, 15);
        __syncthreads();   // B6: staging tiles complete; raw tiles and xr/xk no longer needed
        if (tid == 224) {
            const Unit cu = unit_at(cb);
            tma_store_3d(&D.mgr, o_gr, cu.h * N, t0, cu.b);
            tma_store_3d(&D.mgk, o_gk, cu.h * N, t0, cu.b);
            tma_store_3d(&D.mgv, o_gv, cu.h * N, t0, cu.b);
            tma_store_3d(&D.mgw, sm.o_gw, cu.h * N, t0, cu.b);
            tma_store_commit();
        }
        if (pend) {                      // piece boundary
            piece_end(unit_at(cb));
            if (more) {
                piece_start(unit_at(cb + 1));
                tc_fence_before();
                __syncthreads();         // the new piece's bf16 state tile is visible to the chunk-start MMAs
                tc_fence_after();
            }
        }
    }

    if (tid == 224) tma_store_wait<0>();
    tc_fence_before();
    __syncthreads();
    if (p.dbg && tid == 0 && blockIdx.x < 1024) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.dbg[130 + 4 * blockIdx.x] = t;
        p.dbg[131 + 4 * blockIdx.x] = 0;
    }
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tbase) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------
// Time-segmented backward (few, long sequences: long-form fine-tuning, conf/rwkv/*FT-LFXL*.yaml -- one CTA per sequence
// would leave most SMs idle and the sweep is serial in the reverse state C).  C is linear in itself across a run of chunks,
// C_out = C_in * D + dC with D[j] = product of the run's per-step decays of key channel j, exactly like the forward state
// (wkv6_chunk_lean.cu), so: pass 1 (wkv6_bwd_state_kernel, below) sweeps every segment from C = 0 in a state-only mode --
// decays, the r 2^(A_p) tiles and the dC MMAs, about a fifth of a full chunk -- and leaves dC and D in the scratch;
// wkv6_seg_scan_kernel turns them into every segment's starting C (from gsT); pass 2 is the normal kernel started from
// those (segment parameters at the top of wkv6_bwd_chunk_kernel), with gu as per-segment partials that
// wkv6_bwd_seg_gu_kernel sums.  Four launches; the forward states come from the checkpoints as always.
struct BwdStateSmem {
    alignas(1024) uint8_t raw_r[CL * 128];
    alignas(1024) uint8_t raw_w[CL * 128];
    alignas(1024) uint8_t g[2][CL * 128];
    alignas(1024) uint8_t rst[CL * 128];
    alignas(1024) uint8_t rst_lo[CL * 128];
    float blocktot[4][N];
    alignas(8) uint64_t bar_rw, bar_g[2], bar_m;
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(256) wkv6_bwd_state_kernel(const __grid_constant__ Params p) {
    extern __shared__ __align__(1024) uint8_t smem_bytes[];
    BwdStateSmem& sm = *reinterpret_cast<BwdStateSmem*>(smem_bytes + ((1024u - (smem_u32(smem_bytes) & 1023u)) & 1023u));
    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int BH = p.B * p.H, nseg = p.nseg;
    const int seq = blockIdx.x / nseg, seg = blockIdx.x - seq * nseg;
    const int ds = seq / BH, bh = seq - ds * BH;
    const int b = bh / p.H, h = bh - b * p.H;
    const DirArgs& D = p.d[ds];
    const bool rev = D.reverse != 0;
    const int T = p.T, nchunks = (T + CL - 1) / CL;
    const int cbeg = seg * p.seg_chunks, nloc = min(p.seg_chunks, nchunks - cbeg);
    auto tile_q = [&](int cb) { const int c = nchunks - 1 - (cbeg + cb); return rev ? nchunks - 1 - c : c; };
    auto issue_rw = [&](int cb) {
        const int q = tile_q(cb);
        mbar_arrive_expect_tx(&sm.bar_rw, 2u * CL * 128);
        tma_load_3d(sm.raw_r, &D.mr, &sm.bar_rw, h * N, q * CL, b);
        tma_load_3d(sm.raw_w, &D.mw, &sm.bar_rw, h * N, q * CL, b);
    };
    auto issue_g = [&](int cb) {
        mbar_arrive_expect_tx(&sm.bar_g[cb & 1], 1u * CL * 128);
        tma_load_3d(sm.g[cb & 1], &D.mg, &sm.bar_g[cb & 1], h * N, tile_q(cb) * CL, b);
    };
    if (tid == 0) {
        mbar_init(&sm.bar_rw, 1);
        mbar_init(&sm.bar_g[0], 1);
        mbar_init(&sm.bar_g[1], 1);
        mbar_init(&sm.bar_m, 1);
        fence_barrier_init();
        tma_prefetch_desc(&D.mr); tma_prefetch_desc(&D.mw); tma_prefetch_desc(&D.mg);
        fence_proxy_async();
        if (nloc > 0) { issue_rw(0); issue_g(0); }
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&sm.tmem_base)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // fragment coordinates (TMEM 16x256b.x4) and row-major coordinates: as in wkv6_bwd_chunk_kernel
    const int qtr = warp & 3, half = warp >> 2;
    const int frow = 16 * qtr + (lane >> 2);
    const int fcol = 32 * half + 2 * (lane & 3);
    const uint32_t lane_field = (uint32_t)(qtr * 32) << 16;
    const int I = warp & 3, ch = warp >> 2;
    const int cq = lane & 7, rg = lane >> 3;
    const int c0 = 32 * ch + 4 * cq;
    const int coff = (c0 & 7) * 2, cchunk = c0 >> 3;
    float cs[16];
#pragma unroll
    for (int i = 0; i < 16; i++) cs[i] = 0.f;
    float lsum0 = 0.f, lsum1 = 0.f;              // log2 of the decay product of key channels frow, frow + 8
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tDC = __shfl_sync(0xffffffffu, sm.tmem_base, 0);
    const uint64_t dm_rst = desc_mn(smem_u32(sm.rst)), dm_rstl = desc_mn(smem_u32(sm.rst_lo)), dm_g0 = desc_mn(smem_u32(sm.g[0]));
    constexpr uint32_t ID_MM64 = umma_idesc(1, 1, 64, 64);

    for (int cb = 0; cb < nloc; cb++) {
        const int t0 = tile_q(cb) * CL;
        const uint32_t par = cb & 1;
        if (tid == 0 && cb + 1 < nloc) issue_g(cb + 1);
        mbar_wait(&sm.bar_rw, par);
        // decays, in-block cumulative sums, r 2^(A_p - a_I)  (phase 1 of the full kernel without its K side)
        float ro[4][4];
        {
            float l2[4][4];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int pos = 16 * I + 4 * rg + r;
                const int row = rev ? 63 - pos : pos;
                const int off = sw_chunk(row, cchunk) + coff;
                float w4[4];
                ld4bf(sm.raw_r + off, ro[r]);
                ld4bf(sm.raw_w + off, w4);
                const bool valid = t0 + row < T;
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
        __syncthreads();   // blocktot visible
        {
            float fst[4];
#pragma unroll
            for (int e = 0; e < 4; e++) {
                float s_ = 0.f;
#pragma unroll
                for (int m = 0; m < 4; m++)
                    if (m < I) s_ += sm.blocktot[m][c0 + e];
                fst[e] = ex2(s_);
            }
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int pos = 16 * I + 4 * rg + r;
                const int row = rev ? 63 - pos : pos;
                const int off = sw_chunk(row, cchunk) + coff;
                float x4[4];
#pragma unroll
                for (int e = 0; e < 4; e++) x4[e] = ro[r][e] * fst[e];
                st4bf_hilo(sm.rst + off, sm.rst_lo + off, x4);
            }
        }
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();   // tiles complete; raw r / w are dead
        if (tid == 0 && cb + 1 < nloc) issue_rw(cb + 1);
        if (warp == 1 && elect_one()) {
            tc_fence_after();
            mbar_wait(&sm.bar_g[cb & 1], (cb >> 1) & 1);
            const uint64_t dm_g = dm_g0 + ((cb & 1) ? (uint64_t)((CL * 128) >> 4) : 0);
#pragma unroll
            for (int ks = 0; ks < 4; ks++) umma(tDC, dm_rst + (uint64_t)((ks * 2048) >> 4), dm_g + (uint64_t)((ks * 2048) >> 4), ID_MM64, ks > 0);
#pragma unroll
            for (int ks = 0; ks < 4; ks++) umma(tDC, dm_rstl + (uint64_t)((ks * 2048) >> 4), dm_g + (uint64_t)((ks * 2048) >> 4), ID_MM64, 1);
            umma_commit(&sm.bar_m);
        }
        __syncwarp();
        mbar_wait(&sm.bar_m, par);
        tc_fence_after();
        {
            float x[16];
            tmem_ld_16x32(tDC + lane_field + 32 * half, x);
            const float s0 = (sm.blocktot[0][frow] + sm.blocktot[1][frow]) + (sm.blocktot[2][frow] + sm.blocktot[3][frow]);
            const float s1 = (sm.blocktot[0][frow + 8] + sm.blocktot[1][frow + 8]) + (sm.blocktot[2][frow + 8] + sm.blocktot[3][frow + 8]);
            const float f0 = ex2(s0), f1 = ex2(s1);
            lsum0 += s0;
            lsum1 += s1;
#pragma unroll
            for (int rep = 0; rep < 4; rep++) {
                cs[4 * rep + 0] = fmaf(cs[4 * rep + 0], f0, x[4 * rep + 0]);
                cs[4 * rep + 1] = fmaf(cs[4 * rep + 1], f0, x[4 * rep + 1]);
                cs[4 * rep + 2] = fmaf(cs[4 * rep + 2], f1, x[4 * rep + 2]);
                cs[4 * rep + 3] = fmaf(cs[4 * rep + 3], f1, x[4 * rep + 3]);
            }
        }
        tc_fence_before();
        __syncthreads();   // TMEM and blocktot reads of this chunk done
    }
    {
        float* slot = p.seg_state + ((size_t)seq * nseg + seg) * N * N;       // [i][j], like gs0
#pragma unroll
        for (int rep = 0; rep < 4; rep++)
#pragma unroll
            for (int hh = 0; hh < 2; hh++)
#pragma unroll
                for (int e = 0; e < 2; e++) slot[(fcol + 8 * rep + e) * N + frow + 8 * hh] = cs[4 * rep + 2 * hh + e];
        if (half == 0 && (lane & 3) == 0) {
            float* dec = p.seg_decay + ((size_t)seq * nseg + seg) * N;
            dec[frow] = ex2(lsum0);
            dec[frow + 8] = ex2(lsum1);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tDC) : "memory");
}

// gu[b][h*64 + c] = sum over the segments' partials (pass 2 of a time-segmented backward)
__global__ void __launch_bounds__(64) wkv6_bwd_seg_gu_kernel(const __grid_constant__ Params p) {
    const int BH = p.B * p.H, seq = blockIdx.x, ds = seq / BH, bh = seq - ds * BH;
    const int b = bh / p.H, h = bh - b * p.H;
    float g = 0.f;
    for (int s_ = 0; s_ < p.nseg; s_++) g += p.seg_gu[((size_t)seq * p.nseg + s_) * N + threadIdx.x];
    reinterpret_cast<__nv_bfloat16*>(p.d[ds].gu)[(size_t)b * p.C + h * N + threadIdx.x] = __float2bfloat16_rn(g);
}

constexpr int kMaxSegBwd = 40;   // = kMaxSeg of wkv6_seg_scan_kernel (wkv6_chunk_lean.cu)

// Segments per sequence of a backward launch (0: unsegmented): one CTA per SM is resident, so segmentation pays when the
// launch has at most 74 sequences; every segment at least 8 chunks long.
static int pick_segments_bwd(const Params& p, size_t scratch_bytes, int* seg_chunks) {
    const int nseq = p.ndir * p.B * p.H;
    const int ntot = (p.T + CL - 1) / CL;
    if (nseq <= 0 || nseq * 2 > 148 || ntot < 32) return 0;
    int nseg = 148 / nseq;
    if (nseg > ntot / 8) nseg = ntot / 8;
    if (nseg > kMaxSegBwd) nseg = kMaxSegBwd;
    if (nseg < 2) return 0;
    const int sc = (ntot + nseg - 1) / nseg;
    nseg = (ntot + sc - 1) / sc;
    if ((size_t)nseq * nseg * (N * N + 2 * N) * sizeof(float) > scratch_bytes) return 0;
    *seg_chunks = sc;
    return nseg;
}

// ---- balanced persistent grid (more sequences than SMs): see the schedule at the top of wkv6_bwd_chunk_kernel
constexpr size_t kBalRecBytes = (size_t)N * N * 4 + 512;      // = kRecBytes of the kernel

static int sm_count() {                                       // of the current device, cached per device
    static std::atomic<int> cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    int n = cached[dev & 63].load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        cached[dev & 63].store(n, std::memory_order_relaxed);
    }
    return n;
}

// CTAs per direction of a balanced launch (0: not applicable): the launch has more sequences than the GPU has SMs, i.e. the
// one-CTA-per-sequence grid would run in two or more waves (one CTA is resident per SM) with an idle tail in the last one.
static int pick_balanced(const Params& p) {
    const int nsm = sm_count();
    if (nsm <= 0 || p.ndir <= 0 || p.T <= 0) return 0;
    const long long nseq = (long long)p.ndir * p.B * p.H;
    const int ncta = nsm / p.ndir;
    if (nseq <= nsm || ncta <= 0 || (long long)p.B * p.H < ncta) return 0;
    const long long nch = (p.T + CL - 1) / CL;
    if ((long long)p.B * p.H * nch * (nch > p.H ? nch : p.H) >= (1ll << 31)) return 0;   // exactness of the reciprocal divisions
    if ((long long)p.B * p.H * nch * (ncta + 1) >= (1ll << 31)) return 0;               // 32-bit range arithmetic in the kernel
    return ncta;
}

static size_t balanced_scratch_bytes(const Params& p) {
    const int ncta = pick_balanced(p);
    return ncta ? ((size_t)ncta * p.ndir + 1) * kBalRecBytes : 0;
}

size_t bwd_segment_scratch_bytes(const Params& p) {
    int sc = 0;
    const int nseg = pick_segments_bwd(p, ~(size_t)0, &sc);
    return nseg ? (size_t)p.ndir * p.B * p.H * nseg * (N * N + 2 * N) * sizeof(float) : balanced_scratch_bytes(p);
}

// reciprocals for the kernel's in-loop divisions: x / d = (x * magic) >> 31, exact while x * d < 2^31
static void set_magics(Params& q) {
    q.bal_magic_n = (unsigned)((1ull << 31) / (unsigned)((q.T + CL - 1) / CL)) + 1u;
    q.bal_magic_h = (unsigned)((1ull << 31) / (unsigned)q.H) + 1u;
}

static int launch_bwd_main(const Params& q, int grid, cudaStream_t st) {
    const size_t smem = sizeof(BwdChunkSmem) + 1024;
    cudaError_t e = opt_in_smem_once(wkv6_bwd_chunk_kernel, smem, 1);
    if (e != cudaSuccess) return (int)e;
    e = launch_pdl(wkv6_bwd_chunk_kernel, dim3(grid), dim3(kBwdThreads), smem, st, q);
    count_launch(1);
    return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

int launch_bwd_chunk(const Params& p, cudaStream_t st) {
    Params q = p;
    q.dbg = debug_trace_buffer();
    q.nseg = 0;
    q.bal_ncta = 0;
    set_magics(q);
    return launch_bwd_main(q, q.ndir * q.B * q.H, st);
}

// -1: not applicable (or the scratch is missing / too small): the caller launches launch_bwd_chunk
int launch_bwd_chunk_balanced(const Params& p, void* scratch, size_t scratch_bytes, cudaStream_t st) {
    const int ncta = scratch ? pick_balanced(p) : 0;
    if (ncta == 0 || scratch_bytes < balanced_scratch_bytes(p)) return -1;
    // flags of the hand-over records carry the launch's epoch: 64 bits, never 0, different for every launch of the process
    // (and from whatever the caller's scratch held before, up to a 2^-64 coincidence); consumed flags are reset to 0 by the
    // kernel, which is what makes a CUDA-graph replay (same epoch every time) wait again
    static std::atomic<unsigned long long> counter{0};
    static const unsigned long long salt = [] {
        unsigned long long x = (unsigned long long)std::chrono::steady_clock::now().time_since_epoch().count();
        x ^= (unsigned long long)(uintptr_t)&counter * 0x9E3779B97F4A7C15ull;
        x ^= x >> 29; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 32;
        return x | 1ull;
    }();
    Params q = p;
    q.dbg = debug_trace_buffer();
    q.nseg = 0;
    q.bal_ncta = ncta;
    set_magics(q);
    q.bal_rec = static_cast<unsigned char*>(scratch);
    q.bal_epoch = (salt << 32) ^ (salt >> 32) ^ ((counter.fetch_add(1, std::memory_order_relaxed) + 1) * 0x9E3779B97F4A7C15ull);
    if (q.bal_epoch == 0) q.bal_epoch = 1;
    return launch_bwd_main(q, ncta * q.ndir, st);
}

// -1: the shape is not segmented (or the scratch is missing / too small): the caller launches launch_bwd_chunk
int launch_bwd_chunk_segmented(const Params& p, void* scratch, size_t scratch_bytes, cudaStream_t st) {
    int sc = 0;
    const int nseg = scratch ? pick_segments_bwd(p, scratch_bytes, &sc) : 0;
    if (nseg == 0) return -1;
    Params q = p;
    q.dbg = debug_trace_buffer();
    const int nseq = q.ndir * q.B * q.H;
    q.nseg = nseg;
    q.bal_ncta = 0;
    set_magics(q);
    q.seg_chunks = sc;
    q.seg_state = static_cast<float*>(scratch);
    q.seg_decay = q.seg_state + (size_t)nseq * nseg * N * N;
    q.seg_gu = q.seg_decay + (size_t)nseq * nseg * N;
    const size_t smem1 = sizeof(BwdStateSmem) + 1024;
    static std::atomic<unsigned long long> opted{0};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (!(opted.load(std::memory_order_acquire) & (1ull << (dev & 63)))) {
        if ((e = cudaFuncSetAttribute(wkv6_bwd_state_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1)) != cudaSuccess) return (int)e;
        opted.fetch_or(1ull << (dev & 63), std::memory_order_release);
    }
    wkv6_bwd_state_kernel<<<dim3(nseq * nseg), dim3(256), smem1, st>>>(q);
    count_launch(1);
    if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
    Params qs = q;                                   // the segment scan starts from "s0": here the gradient into S_T
    for (int d = 0; d < q.ndir; d++) qs.d[d].s0 = q.d[d].gsT;
    int rc = launch_seg_scan(qs, st);
    if (rc) return rc;
    if ((rc = launch_bwd_main(q, nseq * nseg, st))) return rc;
    wkv6_bwd_seg_gu_kernel<<<dim3(nseq), dim3(64), 0, st>>>(q);
    count_launch(1);
    return (int)cudaGetLastError();
}

}  // namespace wkv6
