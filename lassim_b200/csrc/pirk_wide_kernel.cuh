// Stage-parallel time-series kernel for 17 .. 64 genes on sm_100a: one trajectory per
// CTA, the S stages of the Gauss collocation pair side by side in the CTA's warps.
//
// Same method, error control and controller as grn_series_pirk_kernel (pirk_kernel.cuh:
// continued-collocation predictor, M = S corrections Y_i <- y + h sum_j A_ij f(Y_j), error
// estimate = last correction of the solution, logistic-swing limit), laid out for systems
// that do not fit one warp: thread = stage * NPAD + gene, one gene per lane, NPAD = 32 or
// 64 lanes (WPS = 1 or 2 warps) per stage, the gene's row of W as ELL (in-degree <= 8) or
// CSR in place.  An attempt costs M + 1 sequential right-hand sides where the 8(5,3) pair
// of the standard kernel needs 12; what a batch of 50-gene candidates waits for is one time
// series of 25,000 attempts (BASELINE config 3).
//
// Exchanges per correction: every thread needs f(Y_j) of ITS gene from all S stages (one
// CTA-wide barrier; the lines are double-buffered so that a fast stage cannot overwrite
// what a slow one still reads), then the whole line Y_i of ITS stage for W Y_i (a barrier
// among the WPS warps of the stage only).
//
// Serves the time series a hand-over evaluation moves out of the standard kernel
// (B.resume = 1, work items from B.long_list) or all of them (B.resume = 0).
#pragma once
#include "grn_kernels.cuh"
#include "pirk_kernel.cuh"

namespace lassim {

__host__ __device__ constexpr int pirk_wide_threads(int S, int WPS) { return S * WPS * 32; }
// exp2 table + scratch, F lines (two buffers), Y lines, logistic values (floats), one slot
constexpr size_t pirk_wide_smem_doubles(int S, int WPS) {
    return (size_t)kSmemHeader + (size_t)(2 * S + S) * (WPS * 32) + (size_t)(S * WPS * 32 + 1) / 2 + 2;
}

template <int S, int WPS, class WT, bool FAST>
__global__ void __launch_bounds__(pirk_wide_threads(S, WPS), 1)
grn_series_pirk_wide_kernel(const DevProblem P, const DevBatch B) {
    static_assert(S == 4 || S == 5, "predictor polynomial written out for 4 and 5 stages");
    static_assert(WPS == 1 || WPS == 2, "one or two warps per stage");
    constexpr int NPAD = WPS * 32;
    constexpr int THREADS = S * NPAD;
    constexpr int NWARPS = THREADS / 32;
    constexpr int M = S;
    constexpr float kExpo = -1.0f / (2 * S);
    using T = GaussTable<S>;
    extern __shared__ double smem[];

    const int tid = threadIdx.x;
    const int stage = tid / NPAD, g = tid - stage * NPAD;
    const int lane = tid & 31, wid = tid >> 5;
    double *tab = smem;
    double *red = smem + kTabDoubles;                     // 32 doubles of reduction scratch
    double *fline = smem + kSmemHeader;                   // [2][S][NPAD]
    double *yline = fline + 2 * S * NPAD;                 // [S][NPAD]
    float *sline = reinterpret_cast<float *>(yline + S * NPAD);     // [S][NPAD]
    unsigned *slot = reinterpret_cast<unsigned *>(sline + S * NPAD + (S * NPAD & 1));

    auto stage_barrier = [&]() {
        if constexpr (WPS == 1) __syncwarp();
        else asm volatile("bar.sync %0, %1;" ::"r"(1 + stage), "n"(NPAD) : "memory");
    };
    // max of a non-negative float / sum of a double over the whole CTA (all threads call)
    auto cta_max = [&](float v) {
        unsigned u = __reduce_max_sync(0xffffffffu, __float_as_uint(v));
        float *rs = reinterpret_cast<float *>(red);
        if (lane == 0) rs[wid] = __uint_as_float(u);
        __syncthreads();
        float m = 0.0f;
#pragma unroll
        for (int i = 0; i < NWARPS; ++i) m = fmaxf(m, rs[i]);
        __syncthreads();
        return m;
    };
    auto cta_sum = [&](double v) {
        v = warp_sum(v);
        if (lane == 0) red[wid] = v;
        __syncthreads();
        double m = 0.0;
#pragma unroll
        for (int i = 0; i < NWARPS; ++i) m += red[i];
        __syncthreads();
        return m;
    };
    auto cta_fmax = [&](double v) {
        v = warp_max(v);
        if (lane == 0) red[wid] = v;
        __syncthreads();
        double m = red[0];
#pragma unroll
        for (int i = 1; i < NWARPS; ++i) m = fmax(m, red[i]);
        __syncthreads();
        return m;
    };

    if (threadIdx.x == 0 && B.started != nullptr) atomicAdd(B.started, 1u);   // see grn_series_pirk_kernel
    fill_exp2_table(tab);
    for (int i = tid; i < 3 * S * NPAD; i += THREADS) fline[i] = 0.0;
    __syncthreads();

    const int n = P.n;
    const int Kc = (B.mode == 0) ? P.K : 1;
    const double rtol = P.rtol, atol = P.atol;
    const bool gene_ok_[1] = {g < n};
    const bool gene_ok = gene_ok_[0];
    const bool writer = stage == 0;

    // my stage's row of A
    double arow[S];
#pragma unroll
    for (int j = 0; j < S; ++j) arow[j] = 0.0;
    double ci = 0.0;
#pragma unroll
    for (int i = 0; i < S; ++i)
        if (stage == i) {
            ci = T::c(i);
#pragma unroll
            for (int j = 0; j < S; ++j) arow[j] = T::A(i, j);
        }

    WT wts;
    wts.clear();
    double y = 0.0, lam = 0.0, vm = 0.0, Fp[S];
    int fpar = 0;
    double *ymine = yline + stage * NPAD + g;
    const double *yrow = yline + stage * NPAD;

    // f(Yin) of my stage and gene
    auto rhs = [&](double Yin, bool keep_sigma) {
        // (the line is rewritten only after a CTA-wide barrier: every right-hand side is
        // followed by a gather or a reduction before the next one)
        *ymine = Yin;
        stage_barrier();
        double z[1];
        wts.matvec(yrow, z);
        const double ly = -lam * Yin;
        double F;
        float sv;
        if constexpr (FAST) {
            double sr, se;
            logistic_fast_parts(z[0], tab, sr, se);
            F = fma(vm * sr, se, fma(vm, sr, ly));
            sv = (float)sr;
        } else {
            const double sg = logistic_libm(z[0]);
            F = fma(vm, sg, ly);
            sv = (float)sg;
        }
        if (keep_sigma) sline[stage * NPAD + g] = sv;
        return F;
    };
    // every stage's derivative of my gene (CTA-wide exchange, double-buffered)
    auto gather = [&](double F, double (&Fall)[S]) {
        double *fl = fline + fpar * (S * NPAD);
        fl[stage * NPAD + g] = F;
        __syncthreads();
#pragma unroll
        for (int j = 0; j < S; ++j) Fall[j] = fl[j * NPAD + g];
        fpar ^= 1;
    };
    auto bsum = [&](const double (&Fall)[S]) {
        double a = T::b(0) * Fall[0], b2 = T::b(1) * Fall[1];
#pragma unroll
        for (int j = 2; j < S; j += 2) {
            a = fma(T::b(j), Fall[j], a);
            if (j + 1 < S) b2 = fma(T::b(j + 1), Fall[j + 1], b2);
        }
        return a + b2;
    };

    while (true) {
        if (tid == 0) *slot = atomicAdd(B.work_counter, 1u);
        __syncthreads();
        unsigned w = *slot;
        __syncthreads();
        const unsigned work_count = B.resume ? __ldcg(B.long_count) : B.work_count;
        if (w >= work_count) break;
        int resume_ti = 0;
        if (B.resume) {
            const unsigned long long e = __ldcg(B.long_list + w);
            w = (unsigned)(e & 0xffffffffull);
            resume_ti = (int)(e >> 32);
        } else {
            w += B.work_begin;
            if (B.order != nullptr) w = B.order[w];
        }
        const int c = (int)w;
        const double *x = B.X + (size_t)c * P.dim;
        lam = 0.0; vm = 0.0; y = 0.0;
        if (gene_ok) { lam = x[g]; vm = x[n + g]; y = P.y0[g]; }
        wts.load(P, x, g, gene_ok_, FAST ? kSigScale : 1.0);
        const double *tl = P.times;
        const int nt = P.T;
        double t = tl[0], tend = (nt > 1) ? tl[1] : t;
        int st = 0;
        unsigned nrej = 0, natt = 0;
        int left = (nt < 2) ? 0 : nt - 1;
        bool rejected = false;
        double *Yc = B.Y + (size_t)c * P.T * n;
        double h_resume = 0.0;
        if (B.resume) {
            const double *Yl = Yc + (size_t)(P.T - 1) * n;
            if (gene_ok) y = __ldcg(Yl + g);
            t = __ldcg(B.cost_main + c);
            h_resume = __ldcg(B.pert_cost + c);
            left = nt - resume_ti;
            tend = tl[resume_ti];
            __syncthreads();           // every thread has its state before anyone may write that row
        } else if (gene_ok && writer) {
            Yc[g] = y;
        }

        // first derivative + starting step (Hairer/Norsett/Wanner II.4, max-norm)
        const double f0 = rhs(y, false);
        const double sc0 = fma(rtol, fabs(y), atol);
        const double d0 = cta_fmax(fabs(y) / sc0), d1 = cta_fmax(fabs(f0) / sc0);
        const double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        const double f1 = rhs(fma(h0, f0, y), false);
        const double d2 = cta_fmax(fabs(f1 - f0) / sc0) / h0;
        const double dm = fmax(d1, d2);
        const double h1 = (dm <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / dm, 1.0 / (M + 1));
        double h = fmin(100.0 * h0, h1);
        if (B.resume) h = h_resume;
        double hp = h, inv_hp = 1.0 / h;
#pragma unroll
        for (int j = 0; j < S; ++j) Fp[j] = f0;

        unsigned budget_left = P.max_steps;
        if (B.resume) {
            const unsigned used = B.traj_steps[(size_t)c * P.K];
            budget_left = used < budget_left ? budget_left - used : 1u;
        }
        while (left > 0) {
            // ------------------------------------------------ one attempt
            double hs = h;
            bool hit = false;
            if (t + 1.01 * hs >= tend) { hs = tend - t; hit = true; }
            double ha[S];
#pragma unroll
            for (int j = 0; j < S; ++j) ha[j] = hs * arow[j];

            // predictor: the previous step's collocation polynomial continued to t + c_i hs
            double Y, F, Fall[S];
            {
                const double xq = ci * (hs * inv_hp);
                const double x2 = xq * xq;
                double pa = 0.0, pb = 0.0;
#pragma unroll
                for (int j = 0; j < S; ++j) {
                    double qj = fma(x2, fma(T::q(j, 3), xq, T::q(j, 2)), fma(T::q(j, 1), xq, T::q(j, 0)));
                    if constexpr (S == 5) qj = fma(x2 * x2, T::q(j, 4), qj);
                    qj *= xq;
                    if (j & 1) pb = fma(qj, Fp[j], pb);
                    else pa = fma(qj, Fp[j], pa);
                }
                Y = fma(hp, pa + pb, y);
            }
            F = rhs(Y, false);
            double sb_old = 0.0, isc = 0.0;
#pragma unroll
            for (int k = 1; k <= M; ++k) {
                gather(F, Fall);
                double a = fma(ha[0], Fall[0], y), b2 = ha[1] * Fall[1];
#pragma unroll
                for (int j = 2; j < S; j += 2) {
                    a = fma(ha[j], Fall[j], a);
                    if (j + 1 < S) b2 = fma(ha[j + 1], Fall[j + 1], b2);
                }
                Y = a + b2;
                if (k == M) {
                    sb_old = bsum(Fall);
                    const double yl = fma(hs, sb_old, y);
                    isc = fast_rcp(fma(rtol, fmax(fabs(y), fabs(yl)), atol));
                }
                F = rhs(Y, k == M);
            }
            gather(F, Fall);               // its barrier also publishes the logistic values
            const double bs = bsum(Fall);
            const double yn = fma(hs, bs, y);
            const double e = fabs(hs * (bs - sb_old)) * isc;
            float ef = (e <= 3.0e38) ? (float)e : CUDART_INF_F;      // NaN -> reject
            {   // resolution limit: swing of my gene's logistic term over the stages
                float lo = 2.0f, hi = -1.0f;
#pragma unroll
                for (int j = 0; j < S; ++j) {
                    const float v = sline[j * NPAD + g];
                    lo = fminf(lo, v); hi = fmaxf(hi, v);
                }
                const float ratio = (hi - lo) * (1.0f / kPirkSwing);
                if (ratio > 1.0f && gene_ok) {
                    float lg, ex;
                    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(ratio));
                    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"((float)(2 * S) * lg));
                    ef = fmaxf(ef, ex);
                }
            }
            if (!gene_ok) ef = 0.0f;
            const float qf = cta_max(ef);
            const float qc = fminf(fmaxf(qf, 1e-30f), 3.0e38f);
            float lg, ex;
            asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(qc));
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"(kExpo * lg));
            const float fac0 = 0.9f * ex;
            const bool accept = qf <= 1.0f;
            ++natt;
            bool failed = false;
            if (accept) {
                t = hit ? tend : t + hs;
                y = yn;
                if (hs > 0.0) {
#pragma unroll
                    for (int j = 0; j < S; ++j) Fp[j] = Fall[j];
                    hp = hs;
                    inv_hp = fast_rcp(hs);
                }
                float fac = fminf(kPirkFacMax, fmaxf(0.2f, fac0));
                if (rejected) fac = fminf(fac, 1.0f);
                rejected = false;
                const double hnew = hs * (double)fac;
                h = hit ? fmax(hnew, h) : hnew;
                if (hit) {
                    if (gene_ok && writer) Yc[(size_t)(nt - left) * n + g] = y;
                    --left;
                    if (left > 0) tend = tl[nt - left];
                }
            } else {
                ++nrej;
                rejected = true;
                h = hs * (double)fmaxf(0.2f, fminf(fac0, 1.0f));
                if (!(fabs(h) > 1e-290) || t + h == t) { st |= 2; failed = true; }
            }
            --budget_left;
            if (!failed && left > 0 && budget_left == 0u) { st |= 1; failed = true; }
            if (failed) {
                for (int s = nt - left; s < nt; ++s)
                    if (gene_ok && writer) Yc[(size_t)s * n + g] = CUDART_NAN;
                left = 0;
            }
        }
        const unsigned nacc = natt - nrej;

        // ---------------------------------------------------------- epilogue
        __syncthreads();                   // the samples are visible to the whole CTA
        double part = 0.0;
        if (gene_ok && writer) {
            double m = -CUDART_INF;
            bool isnan_ = false;
            for (int s = 0; s < P.T; ++s) {
                const double v = Yc[s * n + g];
                isnan_ |= (v != v);
                m = fmax(m, v);
            }
            if (isnan_) m = CUDART_NAN;       // numpy.amax propagates NaN
            for (int s = 0; s < P.T; ++s) {
                const double d = P.data[s * n + g] - Yc[s * n + g] / m;
                part += d * d / P.sigma2[g];
            }
        }
        part = cta_sum(part);
        if (tid == 0) {
            B.cost_main[c] = part;
            B.traj_steps[(size_t)c * P.K] = natt + (B.resume ? B.traj_steps[(size_t)c * P.K] : 0u);
            atomicAdd(&B.counters[(size_t)c * 3 + 0], (unsigned long long)nacc);
            atomicAdd(&B.counters[(size_t)c * 3 + 1], (unsigned long long)nrej);
            atomicAdd(&B.counters[(size_t)c * 3 + 2], (unsigned long long)((M + 1) * S) * natt + 2ull);
            if (st) atomicOr(&B.status[c], st);
            if (B.sp_attempts != nullptr) atomicAdd(B.sp_attempts, (unsigned long long)natt);
        }
        if (B.mode == 0) {
            __threadfence();
            __syncthreads();
            if (tid == 0) *slot = (unsigned)atomicAdd(&B.done[c], 1);
            __syncthreads();
            const unsigned old = *slot;
            __syncthreads();
            if ((int)old == Kc - 1) {
                // last trajectory of candidate c: knock-down cost + total
                // (perturbation_functions.pyx:82-102), one gene per thread of stage 0
                __threadfence();
                double pc = 0.0;
                if (P.K > 1) {
                    const int np2 = n + 2;
                    const bool ok = gene_ok && writer;
                    double rmax = 0.0, pp = 0.0;
                    for (int i = 0; i < n; ++i) {
                        const double *Es = B.E + ((size_t)c * P.K + 1 + P.U + i) * n;
                        const double *Ecn = B.E + ((size_t)c * P.K + P.knock_ctrl[i]) * n;
                        if (ok) {
                            double rr = __ldcg(Es + g) / __ldcg(Ecn + g) - 1.0;
                            if (rr > 2.0) rr = 2.0;
                            rmax = fmax(rmax, fabs(rr));
                            const double p = P.pert[(size_t)i * np2 + g];
                            pp = fma(p, p, pp);
                        }
                    }
                    rmax = cta_fmax(rmax);
                    pp = cta_sum(pp);
                    double rp = 0.0, rr2 = 0.0;
                    // an infinite ratio: the limit of the residual (see grn_cost_kernel)
                    const bool blown = rmax > 1.7e308;
                    const double inv = blown ? 0.0 : (rmax > 0.0) ? 1.0 / rmax : 0.0;
                    for (int i = 0; i < n; ++i) {
                        const double *Es = B.E + ((size_t)c * P.K + 1 + P.U + i) * n;
                        const double *Ecn = B.E + ((size_t)c * P.K + P.knock_ctrl[i]) * n;
                        if (ok) {
                            double rr = __ldcg(Es + g) / __ldcg(Ecn + g) - 1.0;
                            if (rr > 2.0) rr = 2.0;
                            rr = blown ? (fabs(rr) > 1.7e308 ? copysign(1.0, rr) : rr * 0.0) : rr * inv;
                            const double p = P.pert[(size_t)i * np2 + g];
                            rp = fma(rr, p, rp);
                            rr2 = fma(rr, rr, rr2);
                        }
                    }
                    rp = cta_sum(rp);
                    rr2 = cta_sum(rr2);
                    pc = (rmax > 0.0) ? pp - rp * rp / rr2 : pp;
                }
                if (tid == 0) {
                    const double cm = __ldcg(B.cost_main + c);
                    B.pert_cost[c] = pc;
                    B.cost[c] = (P.K > 1) ? fma(P.pert_factor, pc, cm) : cm;
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace lassim
