// Stage-parallel time-series kernel for sm_100a: one trajectory per WARP,
// integrated with a parallel-iterated Gauss collocation pair.
//
// A small population waits for its longest time series (a limit cycle needs
// 16,000 steps of the 8(5,3) pair, 12 right-hand sides one after the other in
// each), so what counts there is the number of SEQUENTIAL right-hand sides, not
// their total.  The S-stage Gauss method (order 2S) solved by fixed-point
// iteration evaluates its S stages independently of each other:
//
//   Y_i^(0)   = collocation polynomial of the previous step, continued to t_n + c_i h
//   Y_i^(k+1) = y_n + h sum_j A_ij f(Y_j^(k)),          k = 0 .. M-1
//   y_(n+1)   = y_n + h sum_j b_j f(Y_j^(M))
//   error     = h sum_j b_j (f(Y_j^(M)) - f(Y_j^(M-1)))  (the last correction; for a time series
//               continued from a hand-over scaled by rho / (1 - rho), rho the ratio of the
//               last two corrections: the last correction is the error of the iterate BEFORE
//               the one that is kept - see the attempt's error measure below)
//
// (van der Houwen & Sommeijer's PIRK scheme with a continued-collocation
// predictor).  The S stages sit side by side in the lanes of the warp - lane
// = stage * LPS + l owns genes 2l and 2l+1 of that stage - so an attempt costs
// M + 1 sequential right-hand sides (6 for S = 5, 5 for S = 4) and, on an
// accuracy-limited trajectory, no more attempts than the 8(5,3) pair needs.
// Error control as everywhere: weight atol + rtol*max(|y_n|, |y_(n+1)|),
// max-norm, factor 0.9 err^(-1/(2S)) within [0.2, kPirkFacMax], no growth after a
// rejection.
//
// Resolution limit.  The last correction measures how far the iteration is from
// the collocation solution, not how far that solution is from the truth; the two
// go together while the solution is smooth on the scale of the step, and part
// company when a logistic term switches INSIDE a step (a regulator sum sweeping
// through zero faster than the step resolves: five nodes see a smooth curve where
// there is a near-jump, the iteration converges happily, the result is off by
// 1e-5).  The explicit pairs catch that with their low-order embedded estimates;
// here the attempt is also rejected when any gene's logistic term differs by more
// than kPirkSwing between two stages of the step, with the step cut in proportion.
// Regular steps of a limit cycle stay below 0.37 (99.9 %), the missed switches
// were at 0.8 and above (tools/study/pirk_study.py).
//
// Costs S times the lanes of the standard kernel, so the host uses it for the
// time series of small batches only.
//
// Two shared-memory exchanges per iteration: every lane needs f(Y_j) of its own
// genes from all stages (to form Y_i), then the whole line Y_i (for W Y_i).
#pragma once
#include "grn_kernels.cuh"
#include "pirk_tables.cuh"

namespace lassim {

#ifndef LASSIM_PIRK_FACMAX
#define LASSIM_PIRK_FACMAX 10.0f
#endif
#ifndef LASSIM_PIRK_SWING
#define LASSIM_PIRK_SWING 0.5f
#endif
#ifndef LASSIM_PIRK_SHARP_FLOOR
#define LASSIM_PIRK_SHARP_FLOOR 0.1f
#endif
// smallest factor the contraction rate may take off the error estimate (1 = the plain last
// correction); tools/study/pirk_estimator_study.py
constexpr float kPirkSharpFloor = LASSIM_PIRK_SHARP_FLOOR;
constexpr float kPirkFacMax = LASSIM_PIRK_FACMAX;
constexpr float kPirkSwing = LASSIM_PIRK_SWING;
// threads per CTA: 4 trajectories of one warp, 2 of two warps, 2 of three warps
__host__ __device__ constexpr int pirk_threads(int WARPS) { return WARPS == 3 ? 192 : 128; }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// stage lines in shared memory per warp: one per LPS lanes of the warp
__host__ __device__ constexpr int pirk_lines(int LPS) { return (32 + LPS - 1) / LPS; }
// doubles between the lines of consecutive stages.  Six lanes per stage: 12, so that the
// warp's 16-byte stores are contiguous (4 wavefronts instead of 8) while the stages'
// broadcast loads still fall on different banks; eight lanes per stage: NC + 2, a store
// is one 128-byte row per stage and the skew keeps the stages' loads apart.
__host__ __device__ constexpr int pirk_stride(int LPS, int NC) {
    return (LPS % 8 == 0) ? NC + 2 : 2 * LPS;
}
// WARPS > 1: doubles of one exchange slot of a warp group (solution + derivatives of every
// lane, error measure and step factor), S stages
__host__ __device__ constexpr int pirk_slot(int S) { return 32 * 2 * (S + 1) + 2; }
// per warp: derivative lines, state lines, logistic values (float2 per lane);
// WARPS > 1 adds, per group of warps, 2 (parity) x WARPS (role) exchange slots and a
// work-item slot
constexpr size_t pirk_smem_doubles(int S, int LPS, int NC, int warps) {
    return (size_t)kSmemHeader +
           (size_t)(pirk_threads(warps) / 32) *
               (2 * pirk_lines(LPS) * pirk_stride(LPS, NC) + pirk_lines(LPS) * LPS) +
           (warps > 1 ? (size_t)(pirk_threads(warps) / (32 * warps)) *
                            (2 * warps * pirk_slot(S) + 2) : 0);
}

// Step sizes attempted side by side, as multiples of the controller's proposal, largest
// first: two warps {1, 0.7} (as in the speculative explicit kernel), three {1.2, 1, 0.7}
// (CPU study of the two Th2 limit cycles: {1.1, 1, 0.7} 11,930 sequential attempts, {1.15, 1, 0.7}
// 11,808, {1.2, 1, 0.7} 11,670, {1.25, 1, 0.75} 11,780, four sizes {1, 0.85, 0.7, 0.5} 12,004).
__host__ __device__ constexpr double pirk_ratio(int WARPS, int role) {
    return WARPS == 2 ? (role == 0 ? 1.0 : 0.7)
         : WARPS == 3 ? (role == 0 ? 1.2 : role == 1 ? 1.0 : 0.7)
                      : 1.0;
}

// WARPS > 1: that many warps per trajectory, on different schedulers.  All carry the same
// state; each attempts the step from that state at its own multiple of the proposed size; the
// largest attempt that passes its error test is taken by all of them (a limit cycle rejects
// 17-25 % of its proposed steps, and a step 20 % above the proposal passes about half the
// time).  One exchange through shared memory and one named barrier per attempt.
template <int S, int LPS, int NC, bool FAST, int WARPS = 1>
__global__ void __launch_bounds__(pirk_threads(WARPS), WARPS == 3 ? 1 : 2)
grn_series_pirk_kernel(const DevProblem P, const DevBatch B) {
    static_assert(S * LPS <= 32 && 2 * LPS <= NC + 3, "stages x lanes must fit one warp");
    static_assert(S == 4 || S == 5, "predictor polynomial written out for 4 and 5 stages");
    constexpr int R = 2;
    constexpr int M = S;                   // corrections per attempt
    constexpr int STR = pirk_stride(LPS, NC);
    static_assert(STR >= NC && STR % 2 == 0, "a stage line holds NC doubles, 16-byte aligned");
    constexpr int LINES = pirk_lines(LPS);
    constexpr int PER_WARP = 2 * LINES * STR + LINES * LPS;
    constexpr float kExpo = -1.0f / (2 * S);
    using T = GaussTable<S>;
    extern __shared__ double smem[];

    const int lane = threadIdx.x & 31;
    const int stage = lane / LPS, l = lane - stage * LPS;
    const bool staged = stage < S;
    const int wid = threadIdx.x >> 5;
    double *tab = smem;
    double *fline = smem + kSmemHeader + wid * PER_WARP;
    double *yline = fline + LINES * STR;
    float2 *sline = reinterpret_cast<float2 *>(yline + LINES * STR);
    // warp group (WARPS > 1): role r attempts pirk_ratio(WARPS, r) times the proposed step
    constexpr bool DUAL = WARPS > 1;
    constexpr int THREADS = pirk_threads(WARPS);
    const int pairq = wid / WARPS, role = wid - pairq * WARPS;
    constexpr int SLOT = pirk_slot(S);
    double *xch = smem + kSmemHeader + (THREADS / 32) * PER_WARP + pairq * (2 * WARPS * SLOT + 2);
    unsigned *work_slot = reinterpret_cast<unsigned *>(xch + 2 * WARPS * SLOT);
    int par = 0;
    auto pair_barrier = [&]() {
        if constexpr (DUAL) {          // literal ids: the kernel reserves 3 barriers, not all 16
            if (pairq == 0) asm volatile("bar.sync 1, %0;" ::"n"(32 * WARPS) : "memory");
            else asm volatile("bar.sync 2, %0;" ::"n"(32 * WARPS) : "memory");
        }
    };
    // this CTA runs: the knock-down kernel behind a hand-over waits until all of them do
    // (start_gate_kernel)
    if (threadIdx.x == 0 && B.started != nullptr) atomicAdd(B.started, 1u);
    fill_exp2_table(tab);
    for (int i = lane; i < PER_WARP; i += 32) fline[i] = 0.0;
    __syncthreads();
    double *fmine = fline + stage * STR + 2 * l, *ymine = yline + stage * STR + 2 * l;
    const double *fcol = fline + 2 * l;    // + j*STR: stage j's derivative of my genes
    const double *yrow = yline + stage * STR;

    const int n = P.n;
    const int Kc = (B.mode == 0) ? P.K : 1;
    const double rtol = P.rtol, atol = P.atol;

    // my stage's row of A and node
    double arow[S], ci = 0.0;
#pragma unroll
    for (int j = 0; j < S; ++j) arow[j] = 0.0;
#pragma unroll
    for (int i = 0; i < S; ++i)
        if (stage == i) {
            ci = T::c(i);
#pragma unroll
            for (int j = 0; j < S; ++j) arow[j] = T::A(i, j);
        }

    bool gene_ok[R];
#pragma unroll
    for (int r = 0; r < R; ++r) gene_ok[r] = staged && (2 * l + r) < n;
    const bool writer = stage == 0 && role == 0;

    DenseW<R, NC> wts;
    double y[R], lam[R], vm[R], Fp[S][R];

    // f(Yin) of my stage and genes: publish the stage's line, mat-vec, logistic; the
    // Newton step of the logistic's reciprocal is folded into the last FMA.
    // keep_sigma: leave the logistic values where the swing test reads them.
    auto rhs = [&](const double (&Yin)[R], double (&Fout)[R], bool keep_sigma) {
        *reinterpret_cast<double2 *>(ymine) = make_double2(Yin[0], Yin[1]);
        __syncwarp();
        double z[R];
        wts.matvec(yrow, z);
        float sv[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double ly = -lam[r] * Yin[r];
            if constexpr (FAST) {
                double sr, se;
                logistic_fast_parts(z[r], tab, sr, se);
                Fout[r] = fma(vm[r] * sr, se, fma(vm[r], sr, ly));
                sv[r] = (float)sr;
            } else {
                const double sg = logistic_libm(z[r]);
                Fout[r] = fma(vm[r], sg, ly);
                sv[r] = (float)sg;
            }
        }
        if (keep_sigma) sline[stage * LPS + l] = make_float2(sv[0], sv[1]);
    };
    // every stage's derivative of my genes
    auto gather = [&](const double (&F)[R], double (&Fall)[S][R]) {
        *reinterpret_cast<double2 *>(fmine) = make_double2(F[0], F[1]);
        __syncwarp();
#pragma unroll
        for (int j = 0; j < S; ++j) {
            const double2 v = *reinterpret_cast<const double2 *>(fcol + j * STR);
            Fall[j][0] = v.x; Fall[j][1] = v.y;
        }
    };
    // sum_j b_j Fall[j][r] in two partial sums
    auto bsum = [&](const double (&Fall)[S][R], int r) {
        double a = T::b(0) * Fall[0][r], b2 = T::b(1) * Fall[1][r];
#pragma unroll
        for (int j = 2; j < S; j += 2) {
            a = fma(T::b(j), Fall[j][r], a);
            if (j + 1 < S) b2 = fma(T::b(j + 1), Fall[j + 1][r], b2);
        }
        return a + b2;
    };

    // First item of trajectory slot q of CTA b: q * gridDim.x + b (odd q: counted from the other
    // end, q * gridDim.x + gridDim.x - 1 - b), later ones from the queue behind those.  With the hand-over list ordered longest first (order_long_list_kernel) the
    // longest time series then sit on different SMs, each next to one from further down the
    // list; taken from the queue, the two slots of a CTA would get neighbours of the list - the
    // two limit cycles of a population on ONE SM, sharing schedulers to the end (+3 ms).  The
    // odd slots run backwards so that the longest entry shares its SM with the shortest of the
    // first wave, or, when the list is shorter than the wave, with nobody.
    constexpr unsigned kSlots = THREADS / (32 * WARPS);
    bool first_item = true;
    while (true) {
        unsigned w = 0;
        if (first_item) {
            w = (unsigned)pairq * gridDim.x + ((pairq & 1) ? gridDim.x - 1u - blockIdx.x : blockIdx.x);
            first_item = false;
        } else if constexpr (DUAL) {
            if (role == 0 && lane == 0) *work_slot = atomicAdd(B.work_counter, 1u);
            pair_barrier();
            w = *work_slot + kSlots * gridDim.x;
            pair_barrier();            // both have read it before the next item overwrites it
        } else {
            if (lane == 0) w = atomicAdd(B.work_counter, 1u);
            w = __shfl_sync(0xffffffffu, w, 0) + kSlots * gridDim.x;
        }
        // resume: the work items are the time series the kernel ahead of this one handed over
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
        const int c = (int)w;                      // time series: item w is candidate w
        const double *x = B.X + (size_t)c * P.dim;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int g = 2 * l + r;
            lam[r] = 0.0; vm[r] = 0.0; y[r] = 0.0;
            if (gene_ok[r]) { lam[r] = x[g]; vm[r] = x[n + g]; y[r] = P.y0[g]; }
        }
        wts.load(P, x, l, gene_ok, FAST ? kSigScale : 1.0);
        const double *tl = P.times;
        const int nt = P.T;
        double t = tl[0], tend = (nt > 1) ? tl[1] : t;
        int st = 0;
        unsigned nrej = 0, natt = 0;               // attempts; accepted = natt - nrej
        int left = (nt < 2) ? 0 : nt - 1;          // sample times still to reach
        bool rejected = false;
        double *Yc = B.Y + (size_t)c * P.T * n;
        double h_resume = 0.0;
        if (B.resume) {
            // state left by the hand-over: y in the last sample row, t and h in the cost slots
            const double *Yl = Yc + (size_t)(P.T - 1) * n;
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (gene_ok[r]) y[r] = __ldcg(Yl + 2 * l + r);
            t = __ldcg(B.cost_main + c);
            h_resume = __ldcg(B.pert_cost + c);
            left = nt - resume_ti;
            tend = tl[resume_ti];
            __syncwarp();              // every lane has its state before anyone may write that row
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (gene_ok[r] && writer) Yc[2 * l + r] = y[r];
        }

        // first derivative + starting step (Hairer/Norsett/Wanner II.4, max-norm)
        double f0[R], ys[R], f1[R];
        rhs(y, f0, false);
        double d0 = 0.0, d1 = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double sc = fma(rtol, fabs(y[r]), atol);
            d0 = fmax(d0, fabs(y[r]) / sc);
            d1 = fmax(d1, fabs(f0[r]) / sc);
        }
        d0 = warp_max(d0); d1 = warp_max(d1);
        const double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
#pragma unroll
        for (int r = 0; r < R; ++r) ys[r] = fma(h0, f0[r], y[r]);
        __syncwarp();
        rhs(ys, f1, false);
        double d2 = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r)
            d2 = fmax(d2, fabs(f1[r] - f0[r]) / fma(rtol, fabs(y[r]), atol));
        d2 = warp_max(d2) / h0;
        const double dm = fmax(d1, d2);
        const double h1 = (dm <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / dm, 1.0 / (M + 1));
        double h = fmin(100.0 * h0, h1);
        if (B.resume) h = h_resume;    // the explicit pair's last proposal: the controller adapts
        // no previous step yet: with every stage's derivative = f(y0) the continued
        // "polynomial" is the Euler line y0 + c_i h f(y0)
        double hp = h, inv_hp = 1.0 / h;
#pragma unroll
        for (int j = 0; j < S; ++j) { Fp[j][0] = f0[0]; Fp[j][1] = f0[1]; }
        __syncwarp();

        // attempts left in the budget, counted down in a register (comparing with the kernel
        // parameter would reload it from the constant bank at the end of every attempt)
        unsigned budget_left = P.max_steps;
        if (B.resume) {                // what the standard kernel used before the hand-over counts
            const unsigned used = B.traj_steps[(size_t)c * P.K];
            budget_left = used < budget_left ? budget_left - used : 1u;
        }
        asm volatile("" : "+r"(budget_left));
        while (left > 0) {
            // ------------------------------------------------ one attempt
            // every role's step (all warps know all of them), clipped to the next sample time
            double hs_of[WARPS];
            bool hit_of[WARPS];
#pragma unroll
            for (int q = 0; q < WARPS; ++q) {
                hs_of[q] = h * pirk_ratio(WARPS, q);
                hit_of[q] = false;
                if (t + 1.01 * hs_of[q] >= tend) { hs_of[q] = tend - t; hit_of[q] = true; }
            }
            double hs = hs_of[0];
            bool hit = hit_of[0];
#pragma unroll
            for (int q = 1; q < WARPS; ++q)
                if (role == q) { hs = hs_of[q]; hit = hit_of[q]; }
            double ha[S];
#pragma unroll
            for (int j = 0; j < S; ++j) ha[j] = hs * arow[j];

            // predictor: int_1^(1 + c_i hs/hp) of the previous step's polynomial,
            // Q_j(x) = x (q_j0 + q_j1 x + ...) in Estrin form
            double Y[R], F[R], Fall[S][R];
            {
                const double xq = ci * (hs * inv_hp);
                const double x2 = xq * xq;
                double pa[R] = {0.0, 0.0}, pb[R] = {0.0, 0.0};
#pragma unroll
                for (int j = 0; j < S; ++j) {
                    double qj = fma(x2, fma(T::q(j, 3), xq, T::q(j, 2)),
                                    fma(T::q(j, 1), xq, T::q(j, 0)));
                    if constexpr (S == 5) qj = fma(x2 * x2, T::q(j, 4), qj);
                    qj *= xq;
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        if (j & 1) pb[r] = fma(qj, Fp[j][r], pb[r]);
                        else pa[r] = fma(qj, Fp[j][r], pa[r]);
                    }
                }
#pragma unroll
                for (int r = 0; r < R; ++r) Y[r] = fma(hp, pa[r] + pb[r], y[r]);
            }
            rhs(Y, F, false);
            double sb_old[R] = {0.0, 0.0}, sb_older[R] = {0.0, 0.0}, isc[R] = {0.0, 0.0};
#pragma unroll
            for (int k = 1; k <= M; ++k) {
                gather(F, Fall);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    // two partial sums: half the dependent chain
                    double a = fma(ha[0], Fall[0][r], y[r]), b2 = ha[1] * Fall[1][r];
#pragma unroll
                    for (int j = 2; j < S; j += 2) {
                        a = fma(ha[j], Fall[j][r], a);
                        if (j + 1 < S) b2 = fma(ha[j + 1], Fall[j + 1][r], b2);
                    }
                    Y[r] = a + b2;
                }
                if (k == M - 1) {
                    // two corrections short of the final one: with the next it gives the rate at
                    // which the corrections shrink (off the dependent chain)
#pragma unroll
                    for (int r = 0; r < R; ++r) sb_older[r] = bsum(Fall, r);
                }
                if (k == M) {
                    // the solution one correction short of the final one: its distance to
                    // the final one is the error estimate; it also stands in for y_(n+1)
                    // in the error weight (they differ by the estimate itself), which
                    // keeps the reciprocal of the weight off the end of the attempt
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        sb_old[r] = bsum(Fall, r);
                        const double yl = fma(hs, sb_old[r], y[r]);
                        isc[r] = fast_rcp(fma(rtol, fmax(fabs(y[r]), fabs(yl)), atol));
                    }
                }
                rhs(Y, F, k == M);
            }
            gather(F, Fall);
            double yn[R];
            float ef = 0.0f, ef0 = 0.0f, es = 0.0f;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const double a = bsum(Fall, r);
                yn[r] = fma(hs, a, y[r]);
                const double e = fabs(hs * (a - sb_old[r])) * isc[r];
                const double e0 = fabs(hs * (sb_old[r] - sb_older[r])) * isc[r];
                ef = fmaxf(ef, (e <= 3.0e38) ? (float)e : CUDART_INF_F);   // NaN -> reject
                ef0 = fmaxf(ef0, (e0 <= 3.0e38) ? (float)e0 : CUDART_INF_F);
            }
            {   // resolution limit: swing of my genes' logistic terms over the stages
                // (written before the barrier inside the last gather)
                float lo0 = 2.0f, hi0 = -1.0f, lo1 = 2.0f, hi1 = -1.0f;
#pragma unroll
                for (int j = 0; j < S; ++j) {
                    const float2 v = sline[j * LPS + l];
                    lo0 = fminf(lo0, v.x); hi0 = fmaxf(hi0, v.x);
                    lo1 = fminf(lo1, v.y); hi1 = fmaxf(hi1, v.y);
                }
                const float ratio = fmaxf(hi0 - lo0, hi1 - lo1) * (1.0f / kPirkSwing);
                if (ratio > 1.0f && staged) {
                    // as an error (ratio^(2S)) the controller cuts the step by 0.9 / ratio
                    float lg, ex;
                    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(ratio));
                    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"((float)(2 * S) * lg));
                    es = ex;
                }
            }
            // non-negative floats order like their bit patterns
            float qf = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(ef)));
            if constexpr (kPirkSharpFloor < 1.0f) {
                // The last correction d_M = |y(M) - y(M-1)| is the error of y(M-1); with the rate
                // rho = d_M / d_(M-1) at which the corrections shrink, the error of y(M), which
                // is what is kept, is rho / (1 - rho) times that.  Applied within
                // [kPirkSharpFloor, 1], only while the iteration contracts (rho < 0.9), and only
                // to time series continued from a hand-over (the default path): integrated from
                // t0 by this kernel (modes on request), the initial transient of one th2_late
                // candidate came out 1e-6 off with it - there the rate of one step says little
                // about the next.  Limit cycles: 15 % fewer attempts.
                const float q0 = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(ef0)));
                if (B.resume && q0 < 3.0e38f && qf < 0.9f * q0) {
                    const float rho = __fdividef(qf, q0);
                    qf *= fminf(1.0f, fmaxf(kPirkSharpFloor, __fdividef(rho, 1.0f - rho)));
                }
            }
            qf = fmaxf(qf, __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(es))));

            const float qc = fminf(fmaxf(qf, 1e-30f), 3.0e38f);
            float lg, ex;
            asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(lg) : "f"(qc));
            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ex) : "f"(kExpo * lg));
            float fac0 = 0.9f * ex;
            bool fallback = false;         // the shorter attempt is the one that counts
            if constexpr (DUAL) {
                // every warp publishes its outcome, one barrier, then all make the same choice:
                // the largest attempt that passed (roles are ordered by step size)
                double *mine_slot = xch + (par * WARPS + role) * SLOT;
#pragma unroll
                for (int j = 0; j < S; ++j)
                    *reinterpret_cast<double2 *>(mine_slot + (j * 32 + lane) * 2) =
                        make_double2(Fall[j][0], Fall[j][1]);
                *reinterpret_cast<double2 *>(mine_slot + (S * 32 + lane) * 2) =
                    make_double2(yn[0], yn[1]);
                if (lane == 0)
                    *reinterpret_cast<float2 *>(mine_slot + SLOT - 2) = make_float2(qf, fac0);
                pair_barrier();
                int pick = WARPS - 1;          // nobody passed: the smallest attempt sets the next size
                float2 qo[WARPS];
#pragma unroll
                for (int q = WARPS - 1; q >= 0; --q) {
                    qo[q] = *reinterpret_cast<const float2 *>(xch + (par * WARPS + q) * SLOT + SLOT - 2);
                    if (qo[q].x <= 1.0f) pick = q;
                }
                float2 chosen = qo[0];
                hs = hs_of[0]; hit = hit_of[0];
#pragma unroll
                for (int q = 1; q < WARPS; ++q)
                    if (pick == q) { chosen = qo[q]; hs = hs_of[q]; hit = hit_of[q]; }
                if (pick != role && chosen.x <= 1.0f) {
                    const double *from = xch + (par * WARPS + pick) * SLOT;
#pragma unroll
                    for (int j = 0; j < S; ++j) {
                        const double2 v = *reinterpret_cast<const double2 *>(from + (j * 32 + lane) * 2);
                        Fall[j][0] = v.x; Fall[j][1] = v.y;
                    }
                    const double2 v = *reinterpret_cast<const double2 *>(from + (S * 32 + lane) * 2);
                    yn[0] = v.x; yn[1] = v.y;
                }
                qf = chosen.x;
                fac0 = chosen.y;
                fallback = pick > 0;           // a larger attempt failed: no growth next time
                par ^= 1;
            }
            const bool accept = qf <= 1.0f;
            ++natt;
            bool failed = false;
            if (accept) {
                t = hit ? tend : t + hs;
#pragma unroll
                for (int r = 0; r < R; ++r) y[r] = yn[r];
                if (hs > 0.0) {            // a repeated sample time keeps the last real step
#pragma unroll
                    for (int j = 0; j < S; ++j) { Fp[j][0] = Fall[j][0]; Fp[j][1] = Fall[j][1]; }
                    hp = hs;
                    inv_hp = fast_rcp(hs);
                }
                float fac = fminf(kPirkFacMax, fmaxf(0.2f, fac0));
                if (rejected || fallback) fac = fminf(fac, 1.0f);
                rejected = fallback;       // the proposed step did fail: no growth next time
                const double hnew = hs * (double)fac;
                h = hit ? fmax(hnew, h) : hnew;
                if (hit) {
                    double *Ys = Yc + (size_t)(nt - left) * n;
#pragma unroll
                    for (int r = 0; r < R; ++r)
                        if (gene_ok[r] && writer) Ys[2 * l + r] = y[r];
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
#pragma unroll
                    for (int r = 0; r < R; ++r)
                        if (gene_ok[r] && writer) Yc[(size_t)s * n + 2 * l + r] = CUDART_NAN;
                left = 0;
            }
        }
        const unsigned nacc = natt - nrej;

        // ---------------------------------------------------------- epilogue
        __syncwarp();
        double part = 0.0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int g = 2 * l + r;
            if (gene_ok[r] && writer) {
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
        }
        part = warp_sum(part);
        if (lane == 0 && role == 0) {
            B.cost_main[c] = part;
            B.traj_steps[(size_t)c * P.K] = natt + (B.resume ? B.traj_steps[(size_t)c * P.K] : 0u);
            atomicAdd(&B.counters[(size_t)c * 3 + 0], (unsigned long long)nacc);
            atomicAdd(&B.counters[(size_t)c * 3 + 1], (unsigned long long)nrej);
            atomicAdd(&B.counters[(size_t)c * 3 + 2],
                      (unsigned long long)((M + 1) * S * WARPS) * natt + 2ull);
            if (st) atomicOr(&B.status[c], st);
            if (B.sp_attempts != nullptr) atomicAdd(B.sp_attempts, (unsigned long long)natt);
        }
        if (B.mode == 0 && role == 0) {
            __threadfence();
            __syncwarp();
            unsigned old = 0;
            if (lane == 0) old = (unsigned)atomicAdd(&B.done[c], 1);
            old = __shfl_sync(0xffffffffu, old, 0);
            if ((int)old == Kc - 1) {
                // last trajectory of candidate c: knock-down cost + total
                // (perturbation_functions.pyx:82-102), one gene per lane
                __threadfence();
                double pc = 0.0;
                if (P.K > 1) {
                    const int np2 = n + 2;
                    const int g = lane;
                    const bool ok = g < n;
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
                    rmax = warp_max(rmax);
                    pp = warp_sum(pp);
                    double rp = 0.0, rr2 = 0.0;
                    // an infinite ratio (a control end state that decayed to 1e-190 under a knocked
                    // one of 1e120): the limit of the residual, in which that entry alone counts
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
                    rp = warp_sum(rp);
                    rr2 = warp_sum(rr2);
                    pc = (rmax > 0.0) ? pp - rp * rp / rr2 : pp;
                }
                if (lane == 0) {
                    const double cm = __ldcg(B.cost_main + c);
                    B.pert_cost[c] = pc;
                    B.cost[c] = (P.K > 1) ? fma(P.pert_factor, pc, cm) : cm;
                }
            }
        }
        __syncwarp();
    }
}

}  // namespace lassim
