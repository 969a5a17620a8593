/*
 * wkv6_oracle_impl.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the reference WKV6 CUDA kernels, included twice by
 * wkv6_oracle.c (REAL=float: same operation order and FMA contraction as the
 * GPU kernel; REAL=double: "truth").  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may call into this.
 *
 * Reference files restated (paths relative to /root/reference):
 *   wenet/rwkv_v6/cuda/wkv6_cuda.cu:9-63     kernel_forward
 *   wenet/rwkv_v6/cuda/wkv6_cuda.cu:66-109   kernel_backward_101 (gr, gu)
 *   wenet/rwkv_v6/cuda/wkv6_cuda.cu:112-151  kernel_backward_102 (gk)
 *   wenet/rwkv_v6/cuda/wkv6_cuda.cu:154-195  kernel_backward_103 (gv)
 *   wenet/rwkv_v6/cuda/wkv6_cuda.cu:198-263  kernel_backward_201 (gw)
 *   wenet/rwkv_v6/cuda/wkv6state_cuda.cu:7-65     kernel_forward (initial state)
 *   wenet/rwkv_v6/cuda/wkv6state_cuda.cu:68-188   kernel_backward_111 (gr,gk,gv,gu,gs)
 *   wenet/rwkv_v6/cuda/wkv6state_cuda.cu:191-293  kernel_backward_222 (gw with state)
 *
 * The reference maps one CUDA thread to one channel `i` of a head and keeps a
 * private float state[_N_] per thread.  Here the thread index becomes the
 * innermost (vectorisable) loop and the per-thread arrays become [j][i]
 * matrices, so that every "thread" still accumulates over j in exactly the
 * reference order (j = 0.._N_-1) with the same FMA contraction nvcc applies
 * under its default -fmad=true.
 *
 * PARITY PINNING: the reference ships no tests or golden vectors for this path
 * (SURVEY.md F7).  This restatement is pinned by (1) torch.autograd of a plain
 * fp64 recurrence (tests/test_oracle.py), (2) outputs of the reference's own
 * unmodified CUDA kernels run on a B200 (oracle/_ref, fixtures under
 * tests/golden/refcuda_*.npz made by tests/golden/make_refcuda_golden.py),
 * (3) the reference's own Python wrappers imported in the build container
 * (tests/golden/make_refpy_golden.py).
 */

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, SUF)

#ifndef WKV6_ORACLE_MAXN
#define WKV6_ORACLE_MAXN 256
#endif

/* d = exp(-exp(w))  -- wkv6_cuda.cu:28 (the GPU uses __expf). */
static inline REAL FN(decay_)(REAL w) { return EXP(-EXP(w)); }

/* ------------------------------------------------------------------------- */
/* Forward.  wkv6_cuda.cu:9-63 (S0 = 0) and wkv6state_cuda.cu:7-65 (S0 = s).  */
/* Extensions defined by the recurrence: final state    */
/* out (sT) and `rev` (run t = T-1..0, i.e. flip-free reversed direction).    */
/* State layout s[b][h][i_value][j_key]  (wkv6state_cuda.cu:15,24).           */
/* ------------------------------------------------------------------------- */
void FN(wkv6_oracle_fwd_)(int B, int T, int C, int H,
                          const REAL *r, const REAL *k, const REAL *v, const REAL *w,
                          const REAL *u, const REAL *s0, REAL *y, REAL *sT, int rev)
{
    const int N = C / H;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int b = 0; b < B; b++)
    for (int h = 0; h < H; h++) {
        REAL *S = (REAL *)malloc(sizeof(REAL) * N * N);   /* S[j][i] */
        REAL d[WKV6_ORACLE_MAXN], yacc[WKV6_ORACLE_MAXN];
        const REAL *uu = u + h * N;
        for (int j = 0; j < N; j++)
            for (int i = 0; i < N; i++)
                S[j * N + i] = s0 ? s0[(((size_t)b * H + h) * N + i) * N + j] : (REAL)0;
        for (int tt = 0; tt < T; tt++) {
            const int t = rev ? (T - 1 - tt) : tt;
            const size_t o = ((size_t)b * T + t) * C + (size_t)h * N;
            const REAL *rr = r + o, *kk = k + o, *vv = v + o, *ww = w + o;
            for (int j = 0; j < N; j++) d[j] = FN(decay_)(ww[j]);
            for (int i = 0; i < N; i++) yacc[i] = 0;
            for (int j = 0; j < N; j++) {
                const REAL rj = rr[j], kj = kk[j], uj = uu[j], dj = d[j];
                REAL *Sj = S + j * N;
                for (int i = 0; i < N; i++) {
                    const REAL x = kj * vv[i];                    /* x = k*v          */
                    yacc[i] = FMA(rj, FMA(uj, x, Sj[i]), yacc[i]); /* y += r*(u*x+s)   */
                    Sj[i] = FMA(Sj[i], dj, x);                     /* s = s*w + x      */
                }
            }
            for (int i = 0; i < N; i++) y[o + i] = yacc[i];
        }
        if (sT)
            for (int j = 0; j < N; j++)
                for (int i = 0; i < N; i++)
                    sT[(((size_t)b * H + h) * N + i) * N + j] = S[j * N + i];
        free(S);
    }
}

/* ------------------------------------------------------------------------- */
/* Literal backward, S0 = 0.  wkv6_cuda.cu:66-263.  gu is [B][C] (per-batch   */
/* partials; the caller sums over B, model.py:151).  Requires T >= 2 exactly  */
/* like the reference (kernel_backward_201 writes gw[t_1]); T == 1 gives gw=0.*/
/* ------------------------------------------------------------------------- */
void FN(wkv6_oracle_bwd_ref_)(int B, int T, int C, int H,
                              const REAL *r, const REAL *k, const REAL *v, const REAL *w,
                              const REAL *u, const REAL *gy,
                              REAL *gr, REAL *gk, REAL *gv, REAL *gw, REAL *gu)
{
    const int N = C / H;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int b = 0; b < B; b++)
    for (int h = 0; h < H; h++) {
        REAL *S  = (REAL *)malloc(sizeof(REAL) * N * N);
        REAL *S2 = (REAL *)malloc(sizeof(REAL) * N * N);
        REAL *sb = (REAL *)calloc((size_t)(T > 2 ? T : 2) * N, sizeof(REAL)); /* sbbbb[t][i] */
        REAL acc[WKV6_ORACLE_MAXN], acc2[WKV6_ORACLE_MAXN], d[WKV6_ORACLE_MAXN], guacc[WKV6_ORACLE_MAXN];
        const REAL *uu = u + h * N;
        const size_t o0 = (size_t)b * T * C + (size_t)h * N;
#define AT(p, t) ((p) + o0 + (size_t)(t) * C)

        /* ---- kernel_backward_101: thread i = key channel, state[j] over value j ---- */
        memset(S, 0, sizeof(REAL) * N * N);
        for (int i = 0; i < N; i++) guacc[i] = 0;
        for (int t = 0; t < T; t++) {
            const REAL *kk = AT(k, t), *vv = AT(v, t), *ww = AT(w, t), *gg = AT(gy, t), *rr = AT(r, t);
            for (int i = 0; i < N; i++) { d[i] = FN(decay_)(ww[i]); acc[i] = 0; acc2[i] = 0; }
            for (int j = 0; j < N; j++) {
                REAL *Sj = S + j * N; const REAL vj = vv[j], gj = gg[j];
                for (int i = 0; i < N; i++) {
                    const REAL x = kk[i] * vj;
                    acc[i]  = FMA(FMA(uu[i], x, Sj[i]), gj, acc[i]);   /* gr += (u*x+s)*gy[j] */
                    acc2[i] = FMA(x, gj, acc2[i]);                     /* gu_ += x*gy[j]      */
                    Sj[i] = FMA(Sj[i], d[i], x);                       /* s = s*w + x         */
                }
            }
            for (int i = 0; i < N; i++) {
                AT(gr, t)[i] = acc[i];
                guacc[i] = FMA(rr[i], acc2[i], guacc[i]);              /* gu += r*gu_         */
            }
        }
        for (int i = 0; i < N; i++) gu[(size_t)b * C + h * N + i] = guacc[i];

        /* ---- kernel_backward_102: thread i = key channel, scccc[j] over value j ---- */
        memset(S, 0, sizeof(REAL) * N * N);
        for (int t = T - 1; t >= 0; t--) {
            const REAL *rr = AT(r, t), *vv = AT(v, t), *ww = AT(w, t), *gg = AT(gy, t);
            for (int i = 0; i < N; i++) { d[i] = FN(decay_)(ww[i]); acc[i] = 0; }
            for (int j = 0; j < N; j++) {
                REAL *Sj = S + j * N; const REAL vj = vv[j], gj = gg[j];
                for (int i = 0; i < N; i++) {
                    const REAL x = rr[i] * gj;
                    acc[i] = FMA(FMA(uu[i], x, Sj[i]), vj, acc[i]);    /* gk += (u*x+s)*v[j]  */
                    Sj[i] = FMA(Sj[i], d[i], x);                       /* s = x + s*w         */
                }
            }
            for (int i = 0; i < N; i++) AT(gk, t)[i] = acc[i];
        }

        /* ---- kernel_backward_103: thread i = value channel, sdddd[j] over key j ---- */
        memset(S, 0, sizeof(REAL) * N * N);
        for (int t = T - 1; t >= 0; t--) {
            const REAL *rr = AT(r, t), *kk = AT(k, t), *ww = AT(w, t), *gg = AT(gy, t);
            for (int j = 0; j < N; j++) d[j] = FN(decay_)(ww[j]);
            for (int i = 0; i < N; i++) acc[i] = 0;
            for (int j = 0; j < N; j++) {
                REAL *Sj = S + j * N; const REAL rj = rr[j], kj = kk[j], uj = uu[j], dj = d[j];
                for (int i = 0; i < N; i++) {
                    const REAL x = gg[i] * rj;
                    acc[i] = FMA(FMA(uj, x, Sj[i]), kj, acc[i]);       /* gv += (u_[j]*x+s)*k[j] */
                    Sj[i] = FMA(Sj[i], dj, x);                         /* s = x + s*w_[j]        */
                }
            }
            for (int i = 0; i < N; i++) AT(gv, t)[i] = acc[i];
        }

        /* ---- kernel_backward_201: thread i = key channel; saaaa/scccc[j] over value j ---- */
        memset(S, 0, sizeof(REAL) * N * N);    /* saaaa */
        memset(S2, 0, sizeof(REAL) * N * N);   /* scccc */
        for (int t = T - 1; t > 1; t--) {      /* wkv6_cuda.cu:214-234 */
            const REAL *gg = AT(gy, t), *vv = AT(v, t - 2), *rr = AT(r, t), *ww = AT(w, t - 1), *kk = AT(k, t - 2);
            for (int i = 0; i < N; i++) { d[i] = FN(decay_)(ww[i]); acc[i] = 0; }
            for (int j = 0; j < N; j++) {
                REAL *Sj = S + j * N; const REAL vj = vv[j], gj = gg[j];
                for (int i = 0; i < N; i++) {
                    Sj[i] = FMA(rr[i], gj, Sj[i]) * d[i];              /* x = r*gy[j]; s = (s+x)*w */
                    acc[i] = FMA(Sj[i], vj, acc[i]);                   /* sum += s*v[j] */
                }
            }
            for (int i = 0; i < N; i++) sb[(size_t)(t - 2) * N + i] = acc[i] * kk[i];
        }
        {
            REAL sss[WKV6_ORACLE_MAXN];
            for (int i = 0; i < N; i++) sss[i] = sb[i];
            for (int i = 0; i < N; i++) AT(gw, 0)[i] = 0;              /* wkv6_cuda.cu:237 */
            if (T > 1)
                for (int i = 0; i < N; i++) AT(gw, 1)[i] = sss[i] * -EXP(AT(w, 1)[i]);
            for (int t = 2; t < T - 1; t++) {                           /* wkv6_cuda.cu:240-261 */
                const REAL *gg = AT(gy, t), *vv = AT(v, t - 2), *rr = AT(r, t), *ww = AT(w, t - 1), *kk = AT(k, t - 2);
                for (int i = 0; i < N; i++) { d[i] = FN(decay_)(ww[i]); acc[i] = 0; }
                for (int j = 0; j < N; j++) {
                    REAL *Sj = S2 + j * N; const REAL vj = vv[j], gj = gg[j];
                    for (int i = 0; i < N; i++) {
                        Sj[i] = FMA(kk[i], vj, Sj[i]) * d[i];          /* x = k*v[j]; s = (s+x)*w */
                        acc[i] = FMA(Sj[i], gj, acc[i]);
                    }
                }
                for (int i = 0; i < N; i++) {
                    sss[i] += FMA(-acc[i], rr[i], sb[(size_t)(t - 1) * N + i]);  /* sss += sbbbb - sum*r */
                    AT(gw, t)[i] = sss[i] * -EXP(AT(w, t)[i]);
                }
            }
            if (T > 1)
                for (int i = 0; i < N; i++) AT(gw, T - 1)[i] = 0;      /* wkv6_cuda.cu:262 */
        }
#undef AT
        free(S); free(S2); free(sb);
    }
}

/* ------------------------------------------------------------------------- */
/* Literal backward with initial state.  wkv6state_cuda.cu:68-293.            */
/* gs is [B][H][N_v][N_k] (per-batch; caller sums over B, model.py:97-98).    */
/* Requires T >= 2 like the reference.                                        */
/* ------------------------------------------------------------------------- */
void FN(wkv6_oracle_bwd_state_ref_)(int B, int T, int C, int H,
                                    const REAL *r, const REAL *k, const REAL *v, const REAL *w,
                                    const REAL *u, const REAL *s0, const REAL *gy,
                                    REAL *gr, REAL *gk, REAL *gv, REAL *gw, REAL *gu, REAL *gs)
{
    const int N = C / H;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int b = 0; b < B; b++)
    for (int h = 0; h < H; h++) {
        REAL *S  = (REAL *)malloc(sizeof(REAL) * N * N);
        REAL *S2 = (REAL *)malloc(sizeof(REAL) * N * N);
        REAL *St = (REAL *)malloc(sizeof(REAL) * N * N);   /* state[j][i] = s0[j_v][i_k] */
        REAL *Wp = (REAL *)malloc(sizeof(REAL) * N * N);
        REAL *sb = (REAL *)calloc((size_t)(T + 2) * N, sizeof(REAL));
        REAL acc[WKV6_ORACLE_MAXN], acc2[WKV6_ORACLE_MAXN], d[WKV6_ORACLE_MAXN], guacc[WKV6_ORACLE_MAXN];
        const REAL *uu = u + h * N;
        const size_t o0 = (size_t)b * T * C + (size_t)h * N;
        const REAL *s0h = s0 + ((size_t)b * H + h) * N * N;
#define AT(p, t) ((p) + o0 + (size_t)(t) * C)
        /* state[j] = _s[j*_N_ + i]  (thread i = key): St[j][i] = s0[j][i]  (wkv6state_cuda.cu:88) */
        for (int j = 0; j < N; j++) for (int i = 0; i < N; i++) St[j * N + i] = s0h[j * N + i];

        /* gr, gu (wkv6state_cuda.cu:96-122) */
        memcpy(S, St, sizeof(REAL) * N * N);
        for (int i = 0; i < N; i++) guacc[i] = 0;
        for (int t = 0; t < T; t++) {
            const REAL *kk = AT(k, t), *vv = AT(v, t), *ww = AT(w, t), *gg = AT(gy, t), *rr = AT(r, t);
            for (int i = 0; i < N; i++) { d[i] = FN(decay_)(ww[i]); acc[i] = 0; acc2[i] = 0; }
            for (int j = 0; j < N; j++) {
                REAL *Sj = S + j * N; const REAL vj = vv[j], gj = gg[j];
                for (int i = 0; i < N; i++) {
                    const REAL x = kk[i] * vj;
                    acc[i]  = FMA(FMA(uu[i], x, Sj[i]), gj, acc[i]);
                    acc2[i] = FMA(x, gj, acc2[i]);
                    Sj[i] = FMA(Sj[i], d[i], x);
                }
            }
            for (int i = 0; i < N; i++) { AT(gr, t)[i] = acc[i]; guacc[i] = FMA(rr[i], acc2[i], guacc[i]); }
        }
        for (int i = 0; i < N; i++) gu[(size_t)b * C + h * N + i] = guacc[i];

        /* gk (wkv6state_cuda.cu:124-146) */
        memset(S, 0, sizeof(REAL) * N * N);
        for (int t = T - 1; t >= 0; t--) {
            const REAL *rr = AT(r, t), *vv = AT(v, t), *ww = AT(w, t), *gg = AT(gy, t);
            for (int i = 0; i < N; i++) { d[i] = FN(decay_)(ww[i]); acc[i] = 0; }
            for (int j = 0; j < N; j++) {
                REAL *Sj = S + j * N; const REAL vj = vv[j], gj = gg[j];
                for (int i = 0; i < N; i++) {
                    const REAL x = rr[i] * gj;
                    acc[i] = FMA(FMA(uu[i], x, Sj[i]), vj, acc[i]);
                    Sj[i] = FMA(Sj[i], d[i], x);
                }
            }
            for (int i = 0; i < N; i++) AT(gk, t)[i] = acc[i];
        }

        /* gv (wkv6state_cuda.cu:148-170): thread i = value, j = key */
        memset(S, 0, sizeof(REAL) * N * N);
        for (int t = T - 1; t >= 0; t--) {
            const REAL *rr = AT(r, t), *kk = AT(k, t), *ww = AT(w, t), *gg = AT(gy, t);
            for (int j = 0; j < N; j++) d[j] = FN(decay_)(ww[j]);
            for (int i = 0; i < N; i++) acc[i] = 0;
            for (int j = 0; j < N; j++) {
                REAL *Sj = S + j * N; const REAL rj = rr[j], kj = kk[j], uj = uu[j], dj = d[j];
                for (int i = 0; i < N; i++) {
                    const REAL x = gg[i] * rj;
                    acc[i] = FMA(FMA(uj, x, Sj[i]), kj, acc[i]);
                    Sj[i] = FMA(Sj[i], dj, x);
                }
            }
            for (int i = 0; i < N; i++) AT(gv, t)[i] = acc[i];
        }

        /* gs (wkv6state_cuda.cu:172-191): thread i = value, sssss[j]/swwww[j] over key j */
        memset(S, 0, sizeof(REAL) * N * N);
        for (int q = 0; q < N * N; q++) Wp[q] = 1;
        for (int t = 0; t < T; t++) {
            const REAL *rr = AT(r, t), *ww = AT(w, t), *gg = AT(gy, t);
            for (int j = 0; j < N; j++) d[j] = FN(decay_)(ww[j]);
            for (int j = 0; j < N; j++) {
                REAL *Sj = S + j * N, *Wj = Wp + j * N; const REAL rj = rr[j], dj = d[j];
                for (int i = 0; i < N; i++) {
                    Sj[i] = FMA(gg[i] * Wj[i], rj, Sj[i]);             /* sssss += gyy*w*r[j] */
                    Wj[i] *= dj;
                }
            }
        }
        for (int i = 0; i < N; i++) for (int j = 0; j < N; j++)
            gs[(((size_t)b * H + h) * N + i) * N + j] = S[j * N + i];

        /* gw (wkv6state_cuda.cu:194-293): thread i = key; arrays over value j */
        memset(S, 0, sizeof(REAL) * N * N);    /* saaaa */
        memset(S2, 0, sizeof(REAL) * N * N);   /* scccc */
        for (int t = T - 1; t > 1; t--) {
            const REAL *gg = AT(gy, t), *vv = AT(v, t - 2), *rr = AT(r, t), *ww = AT(w, t - 1), *kk = AT(k, t - 2);
            for (int i = 0; i < N; i++) { d[i] = FN(decay_)(ww[i]); acc[i] = 0; }
            for (int j = 0; j < N; j++) {
                REAL *Sj = S + j * N; const REAL vj = vv[j], gj = gg[j];
                for (int i = 0; i < N; i++) {
                    Sj[i] = FMA(rr[i], gj, Sj[i]) * d[i];              /* s = (s + r*gy[j])*w */
                    acc[i] = FMA(Sj[i], vj, acc[i]);
                }
            }
            for (int i = 0; i < N; i++) sb[(size_t)(t - 1) * N + i] = acc[i] * kk[i];
        }
        {
            REAL sss[WKV6_ORACLE_MAXN];
            if (T > 1) {
                const REAL *gg = AT(gy, 1), *rr = AT(r, 1), *ww = AT(w, 0);
                for (int i = 0; i < N; i++) { d[i] = FN(decay_)(ww[i]); acc[i] = 0; }
                for (int j = 0; j < N; j++) {
                    REAL *Sj = S + j * N; const REAL gj = gg[j];
                    for (int i = 0; i < N; i++) {
                        Sj[i] = FMA(rr[i], gj, Sj[i]) * d[i];
                        acc[i] = FMA(Sj[i], St[j * N + i], acc[i]);
                    }
                }
                for (int i = 0; i < N; i++) sb[i] = acc[i];
            }
            for (int i = 0; i < N; i++) { sss[i] = sb[i]; AT(gw, 0)[i] = sss[i] * -EXP(AT(w, 0)[i]); }
            if (T > 1) {
                const REAL *gg = AT(gy, 1), *rr = AT(r, 1), *ww = AT(w, 0);
                for (int i = 0; i < N; i++) { d[i] = FN(decay_)(ww[i]); acc[i] = 0; }
                for (int j = 0; j < N; j++) {
                    REAL *Sj = S2 + j * N; const REAL gj = gg[j];
                    for (int i = 0; i < N; i++) {
                        Sj[i] = (Sj[i] + St[j * N + i]) * d[i];
                        acc[i] = FMA(Sj[i], gj, acc[i]);
                    }
                }
                for (int i = 0; i < N; i++) {
                    sss[i] += FMA(-acc[i], rr[i], sb[(size_t)1 * N + i]);
                    AT(gw, 1)[i] = sss[i] * -EXP(AT(w, 1)[i]);
                }
            }
            for (int t = 2; t < T - 1; t++) {
                const REAL *gg = AT(gy, t), *vv = AT(v, t - 2), *rr = AT(r, t), *ww = AT(w, t - 1), *kk = AT(k, t - 2);
                for (int i = 0; i < N; i++) { d[i] = FN(decay_)(ww[i]); acc[i] = 0; }
                for (int j = 0; j < N; j++) {
                    REAL *Sj = S2 + j * N; const REAL vj = vv[j], gj = gg[j];
                    for (int i = 0; i < N; i++) {
                        Sj[i] = FMA(kk[i], vj, Sj[i]) * d[i];
                        acc[i] = FMA(Sj[i], gj, acc[i]);
                    }
                }
                for (int i = 0; i < N; i++) {
                    sss[i] += FMA(-acc[i], rr[i], sb[(size_t)t * N + i]);
                    AT(gw, t)[i] = sss[i] * -EXP(AT(w, t)[i]);
                }
            }
            if (T > 1)
                for (int i = 0; i < N; i++) AT(gw, T - 1)[i] = 0;      /* wkv6state_cuda.cu:292 */
        }
#undef AT
        free(S); free(S2); free(St); free(Wp); free(sb);
    }
}

/* ------------------------------------------------------------------------- */
/* General backward (extension).  Supports an initial    */
/* state s0, a gradient gsT flowing into the final state, and `rev`.  Not in  */
/* the reference: defined by differentiating the recurrence, and checked      */
/* against (a) the literal restatements above when s0 = gsT = 0 / gsT = 0 and */
/* (b) torch.autograd on an fp64 recurrence (tests/test_oracle.py).           */
/*   gw_t[j] = -exp(w_t[j]) * d_t[j] * sum_i S_t[i][j] * C_t[i][j]            */
/* Stores the S_t history: O(T*N*N) memory per head, test sizes only.         */
/* ------------------------------------------------------------------------- */
void FN(wkv6_oracle_bwd_)(int B, int T, int C, int H,
                          const REAL *r, const REAL *k, const REAL *v, const REAL *w,
                          const REAL *u, const REAL *s0, const REAL *gy, const REAL *gsT,
                          REAL *gr, REAL *gk, REAL *gv, REAL *gw, REAL *gu, REAL *gs0, int rev)
{
    const int N = C / H;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int b = 0; b < B; b++)
    for (int h = 0; h < H; h++) {
        REAL *Sh = (REAL *)malloc(sizeof(REAL) * (size_t)T * N * N);  /* S_t[i][j] before update */
        REAL *S = (REAL *)malloc(sizeof(REAL) * N * N);
        REAL *Cm = (REAL *)malloc(sizeof(REAL) * N * N);
        REAL d[WKV6_ORACLE_MAXN], guacc[WKV6_ORACLE_MAXN];
        const REAL *uu = u + h * N;
        const size_t so = ((size_t)b * H + h) * N * N;
        for (int q = 0; q < N * N; q++) S[q] = s0 ? s0[so + q] : (REAL)0;
        for (int j = 0; j < N; j++) guacc[j] = 0;
        for (int tt = 0; tt < T; tt++) {
            const int t = rev ? (T - 1 - tt) : tt;
            const size_t o = ((size_t)b * T + t) * C + (size_t)h * N;
            memcpy(Sh + (size_t)tt * N * N, S, sizeof(REAL) * N * N);
            REAL vg = 0;
            for (int i = 0; i < N; i++) vg += v[o + i] * gy[o + i];
            for (int j = 0; j < N; j++) {
                REAL a = 0;
                for (int i = 0; i < N; i++) a += gy[o + i] * S[i * N + j];
                gr[o + j] = a + uu[j] * k[o + j] * vg;
                guacc[j] += r[o + j] * k[o + j] * vg;
                d[j] = FN(decay_)(w[o + j]);
            }
            for (int i = 0; i < N; i++)
                for (int j = 0; j < N; j++)
                    S[i * N + j] = S[i * N + j] * d[j] + k[o + j] * v[o + i];
        }
        for (int j = 0; j < N; j++) gu[(size_t)b * C + h * N + j] = guacc[j];
        for (int q = 0; q < N * N; q++) Cm[q] = gsT ? gsT[so + q] : (REAL)0;
        for (int tt = T - 1; tt >= 0; tt--) {
            const int t = rev ? (T - 1 - tt) : tt;
            const size_t o = ((size_t)b * T + t) * C + (size_t)h * N;
            const REAL *St = Sh + (size_t)tt * N * N;
            REAL vg = 0;
            for (int i = 0; i < N; i++) vg += v[o + i] * gy[o + i];
            for (int j = 0; j < N; j++) {
                REAL a = 0, sc = 0;
                for (int i = 0; i < N; i++) { a += v[o + i] * Cm[i * N + j]; sc += St[i * N + j] * Cm[i * N + j]; }
                d[j] = FN(decay_)(w[o + j]);
                gk[o + j] = a + uu[j] * r[o + j] * vg;
                gw[o + j] = -EXP(w[o + j]) * d[j] * sc;
            }
            for (int i = 0; i < N; i++) {
                REAL a = 0;
                for (int j = 0; j < N; j++) a += k[o + j] * (uu[j] * r[o + j] * gy[o + i] + Cm[i * N + j]);
                gv[o + i] = a;
            }
            for (int i = 0; i < N; i++)
                for (int j = 0; j < N; j++)
                    Cm[i * N + j] = Cm[i * N + j] * d[j] + gy[o + i] * r[o + j];
        }
        if (gs0) for (int q = 0; q < N * N; q++) gs0[so + q] = Cm[q];
        free(Sh); free(S); free(Cm);
    }
}

#undef FN
#undef CAT
#undef CAT_
