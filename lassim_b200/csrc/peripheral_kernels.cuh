// Peripheral genes against a FIXED core (BASELINE config 5) without integrating the core
// again for every candidate.
//
// A peripheral gene does not act back on the core, so the core trajectory is the same for
// every candidate of every gene.  core_dense_kernel integrates it ONCE (one warp, lane =
// core gene, Dormand-Prince 8(5,3) with its 7th-order continuous extension - Hairer's
// dop853 / contd8, tables in dop853_dense_tables.cuh) and writes, per accepted step k, the
// interval [t_k, t_k + h_k] and the interpolant's coefficient vectors.  Sample times are
// step ends.  peripheral_march_kernel then gives every candidate ONE THREAD: the candidate's
// state is a scalar,
//     x' = -lambda x + vmax / (1 + exp(-w . y_core(t))),
// and because the interpolant is linear in its coefficient vectors, w . y_core(t) inside
// interval k is a polynomial in theta = (t - t_k) / h_k whose 8 scalar coefficients are 8 dot
// products per candidate and interval.  All threads walk the core's intervals together (the
// table is read with warp-uniform loads) and sub-step adaptively inside each with the
// scalar 8(5,3) pair: same tolerances, error weight, controller and jump guard as the cost
// kernels.  No shuffles, no shared-memory exchange, 13x fewer states than the joint system.
#pragma once
#include "dop853_dense_tables.cuh"
#include "grn_kernels.cuh"

namespace lassim {

// Layout of the core table (doubles): per interval kCoreHdr header entries
// (t_k, h_k, sample index reached at its end or -1, unused) followed by 8 vectors of n:
// y(t_k) and the 7 coefficient vectors F0..F6 of the interpolant.
constexpr int kCoreHdr = 4;
__host__ __device__ constexpr size_t core_interval_doubles(int n) { return kCoreHdr + 8 * (size_t)n; }

struct CoreTable {
    double *data;          // [capacity][core_interval_doubles(n)]
    int *count;            // number of intervals written; -1 - status on failure
    double *samples;       // [T][n] core states at the sample times
    int capacity;
};

// y_old + x*(F0 + (1-x)*(F1 + x*(F2 + (1-x)*(F3 + x*(F4 + (1-x)*(F5 + x*F6))))))
__device__ __forceinline__ double dense_eval(const double (&c)[8], double x) {
    const double u = 1.0 - x;
    double a = c[7] * x;
    a = (a + c[6]) * u;
    a = (a + c[5]) * x;
    a = (a + c[4]) * u;
    a = (a + c[3]) * x;
    a = (a + c[2]) * u;
    a = (a + c[1]) * x;
    return a + c[0];
}

// One warp integrates the core (n <= 32 genes, lane = gene, dense row of W in registers).
template <int NC>
__global__ void __launch_bounds__(32, 1)
core_dense_kernel(const DevProblem P, const double *__restrict__ core_x, const CoreTable out,
                  const double rtol, const double atol) {
    using D = Dop853Dense;
    extern __shared__ double smem[];
    double *tab = smem;
    double *ybuf = smem + kSmemHeader;             // NC + 2 doubles
    const int g = threadIdx.x, n = P.n;
    const bool ok_[1] = {g < n};
    const bool ok = ok_[0];
    fill_exp2_table(tab);
    for (int i = g; i < NC + 2; i += 32) ybuf[i] = 0.0;
    __syncwarp();

    DenseW<1, NC> wts;
    wts.load(P, core_x, g, ok_, kSigScale);
    double lam = 0.0, vm = 0.0, y = 0.0;
    if (ok) { lam = core_x[g]; vm = core_x[n + g]; y = P.y0[g]; }
    int slo = 0x7ff00000, shi = 0;                 // logistic swing over an attempt (jump guard)
    int slast = 0x3fe00000;                        // logistic value of the last evaluation
    auto rhs = [&](double yin) {
        if (g < NC) ybuf[g] = yin;
        __syncwarp();
        double z[1];
        wts.matvec(ybuf, z);
        __syncwarp();
        double sr, se;
        logistic_fast_parts(z[0], tab, sr, se);
        const double sg = fma(sr, se, sr);
        slast = __double2hiint(sg);
        slo = min(slo, slast); shi = max(shi, slast);
        return fma(vm, sg, -lam * yin);
    };
    auto wmax = [&](double v) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
        return v;
    };

    const int T = P.T;
    double t = P.times[0];
    if (ok) out.samples[g] = y;
    double K[D::kExtended];
    K[0] = rhs(y);
    int sprev = slast;                             // logistic value at y: where an attempt's swing starts
    // starting step (Hairer/Norsett/Wanner II.4, max-norm), as in the cost kernels
    double h;
    {
        const double sc = fma(rtol, fabs(y), atol);
        const double d0 = wmax(ok ? fabs(y) / sc : 0.0), d1 = wmax(ok ? fabs(K[0]) / sc : 0.0);
        const double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        const double f1 = rhs(fma(h0, K[0], y));
        const double d2 = wmax(ok ? fabs(f1 - K[0]) / sc : 0.0) / h0;
        const double dm = fmax(d1, d2);
        const double h1 = (dm <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / dm, 0.125);
        h = fmin(100.0 * h0, h1);
    }
    int k = 0, status = 0;
    bool rejected = false;
    unsigned attempts = 0;
    for (int ti = 1; ti < T && status == 0; ++ti) {
        const double tend = P.times[ti];
        while (status == 0) {
            if (attempts >= P.max_steps) { status = 1; break; }
            double hs = h;
            bool hit = false;
            if (t + 1.01 * hs >= tend) { hs = tend - t; hit = true; }
            slo = sprev; shi = sprev;
#pragma unroll
            for (int s = 1; s < D::kStages; ++s) {
                double acc = 0.0;
#pragma unroll
                for (int j = 0; j < s; ++j)
                    if (D::A(s, j) != 0.0) acc = fma(D::A(s, j), K[j], acc);
                K[s] = rhs(fma(hs, acc, y));
            }
            double yn;
            {
                double acc = 0.0;
#pragma unroll
                for (int j = 0; j < D::kStages; ++j)
                    if (D::B(j) != 0.0) acc = fma(D::B(j), K[j], acc);
                yn = fma(hs, acc, y);
            }
            K[D::kStages] = rhs(yn);
            const int snew = slast;
            double e5 = 0.0, e3 = 0.0;
#pragma unroll
            for (int j = 0; j <= D::kStages; ++j) {
                if (D::E5(j) != 0.0) e5 = fma(D::E5(j), K[j], e5);
                if (D::E3(j) != 0.0) e3 = fma(D::E3(j), K[j], e3);
            }
            const double isc = 1.0 / fma(rtol, fmax(fabs(y), fabs(yn)), atol);
            double q5 = ok ? fabs(hs * e5) * isc : 0.0, q3 = ok ? fabs(hs * e3) * isc : 0.0;
            const double sw = __hiloint2double(shi, 0) - __hiloint2double(min(slo, shi), 0);
            if (ok && sw > (double)P.swing) q5 = fmax(q5, fabs(hs) * sw * vm * isc);
            if (!(q5 == q5) || !(q3 == q3)) q5 = CUDART_INF;
            q5 = wmax(q5); q3 = wmax(q3);
            const double den = fma(q5, q5, 0.01 * q3 * q3);
            const double err = (q5 < 1e18 && q3 < 1e18) ? (den > 0.0 ? q5 * q5 / sqrt(den) : 0.0)
                                                        : CUDART_INF;
            const double fac0 = 0.9 * pow(fmin(fmax(err, 1e-30), 1e300), -0.125);
            ++attempts;
            if (err <= 1.0) {
                if (k >= out.capacity) { status = 4; break; }
                // the three extra stages and the coefficient vectors of the interpolant
#pragma unroll
                for (int s = D::kStages + 1; s < D::kExtended; ++s) {
                    double acc = 0.0;
#pragma unroll
                    for (int j = 0; j < s; ++j)
                        if (D::A(s, j) != 0.0) acc = fma(D::A(s, j), K[j], acc);
                    K[s] = rhs(fma(hs, acc, y));
                }
                double *row = out.data + (size_t)k * core_interval_doubles(n);
                if (g == 0) {
                    row[0] = t; row[1] = hs; row[2] = hit ? (double)ti : -1.0; row[3] = 0.0;
                }
                if (ok) {
                    const double dy = yn - y;
                    double *v = row + kCoreHdr + g;
                    v[0] = y;
                    v[n] = dy;
                    v[2 * n] = fma(hs, K[0], -dy);
                    v[3 * n] = 2.0 * dy - hs * (K[D::kStages] + K[0]);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        double acc = 0.0;
#pragma unroll
                        for (int j = 0; j < D::kExtended; ++j)
                            if (D::D(i, j) != 0.0) acc = fma(D::D(i, j), K[j], acc);
                        v[(4 + i) * n] = hs * acc;
                    }
                }
                ++k;
                t = hit ? tend : t + hs;
                y = yn;
                sprev = snew;
                K[0] = K[D::kStages];
                double fac = fmin(10.0, fmax(0.2, fac0));
                if (rejected) fac = fmin(fac, 1.0);
                rejected = false;
                const double hnew = hs * fac;
                h = hit ? fmax(hnew, h) : hnew;
                if (hit) {
                    if (ok) out.samples[(size_t)ti * n + g] = y;
                    break;
                }
            } else {
                rejected = true;
                h = hs * fmax(0.2, fmin(fac0, 1.0));
                if (!(fabs(h) > 1e-290) || t + h == t) status = 2;
            }
        }
    }
    if (g == 0) *out.count = status ? -1 - status : k;
}

struct PeriphBatch {
    int N;                         // candidates
    int n;                         // core genes
    int T;                         // sample times
    const double *X;               // [N][n + 2] = lambda_p, vmax_p, w_p(n)
    const int *pgene;              // [N] gene of each candidate
    const unsigned char *pmask;    // [genes][n] allowed regulators
    const double *pdata;           // [genes][T]
    const double *psigma2;         // [genes]
    const double *py0;             // [genes]
    const double *table;           // core table
    int intervals;
    double *Ys;                    // [T][N] scratch: the candidates' states at the sample times
    double *cost;                  // [N]
    int *status;                   // [N]
    unsigned int *steps;           // [N] attempts made
    unsigned long long *counters;  // [N*3] accepted, rejected, right-hand sides
    double rtol, atol;
    unsigned int max_steps;
    float swing;
};

// One thread per candidate; NC = core genes rounded up (weights in registers).
template <int NC>
__global__ void __launch_bounds__(128)
peripheral_march_kernel(const PeriphBatch B) {
    using D = Dop853Dense;
    __shared__ double tab[kTabDoubles];
    fill_exp2_table(tab);
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < B.N;
    const int n = B.n, T = B.T;
    const int ii = live ? i : 0;
    const int gene = B.pgene[ii];
    const double *xp = B.X + (size_t)ii * (n + 2);
    const double lam = xp[0], vm = xp[1];
    double w[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j)
        w[j] = (j < n && B.pmask[(size_t)gene * n + j]) ? xp[2 + j] * kSigScale : 0.0;
    double x = B.py0[gene];
    double xmax = x;
    bool bad = !(x == x);
    if (live) B.Ys[i] = x;
    const double rtol = B.rtol, atol = B.atol;

    double zc[8];                  // w . (y_k, F0 .. F6) of the current interval
    double inv_h = 0.0;
    int slo = 0x7ff00000, shi = 0, slast = 0x3fe00000, sprev = 0x3fe00000;
    // f at local time s of the current interval, state xs
    auto rhs = [&](double s, double xs) {
        const double z = dense_eval(zc, s * inv_h);
        double sr, se;
        logistic_fast_parts(z, tab, sr, se);
        const double sg = fma(sr, se, sr);
        slast = __double2hiint(sg);
        slo = min(slo, slast); shi = max(shi, slast);
        return fma(vm, sg, -lam * xs);
    };

    double h = 0.0, k1 = 0.0;
    unsigned nacc = 0, nrej = 0;
    int st = 0;
    bool rejected = false;
    const size_t stride = core_interval_doubles(n);
    for (int k = 0; k < B.intervals; ++k) {
        const double *row = B.table + (size_t)k * stride;
        const double hk = __ldg(row + 1);
        const int samp = (int)__ldg(row + 2);
#pragma unroll
        for (int m = 0; m < 8; ++m) {
            double a0 = 0.0, a1 = 0.0;
            const double *v = row + kCoreHdr + (size_t)m * n;
#pragma unroll
            for (int j = 0; j < NC; j += 2) {
                if (j < n) a0 = fma(w[j], __ldg(v + j), a0);
                if (j + 1 < n) a1 = fma(w[j + 1], __ldg(v + j + 1), a1);
            }
            zc[m] = a0 + a1;
        }
        inv_h = hk > 0.0 ? 1.0 / hk : 0.0;
        double s = 0.0;
        if (k == 0) {
            // first derivative and starting step (scalar form of the cost kernels' rule)
            k1 = rhs(0.0, x);
            sprev = slast;
            const double sc = fma(rtol, fabs(x), atol);
            const double d0 = fabs(x) / sc, d1 = fabs(k1) / sc;
            const double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
            const double f1 = rhs(fmin(h0, hk), fma(h0, k1, x));
            const double d2 = fabs(f1 - k1) / sc / h0;
            const double dm = fmax(d1, d2);
            const double h1 = (dm <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / dm, 0.125);
            h = fmin(100.0 * h0, h1);
        }
        // sub-steps inside the interval; the warp stays together until all its lanes are through
        bool inside = live && st == 0 && hk > 0.0;
        // (the budget is also checked here: an attempt that ends its interval leaves the loop
        // below before the check there)
        if (inside && nacc + nrej >= B.max_steps) { st |= 1; inside = false; }
        while (__any_sync(0xffffffffu, inside)) {
            double hs = h;
            bool hit = false;
            if (s + 1.01 * hs >= hk) { hs = hk - s; hit = true; }
            slo = sprev; shi = sprev;
            double K[D::kStages + 1];
            K[0] = k1;
#pragma unroll
            for (int q = 1; q < D::kStages; ++q) {
                double acc = 0.0;
#pragma unroll
                for (int j = 0; j < q; ++j)
                    if (D::A(q, j) != 0.0) acc = fma(D::A(q, j), K[j], acc);
                K[q] = rhs(fma(D::C(q), hs, s), fma(hs, acc, x));
            }
            double xn;
            {
                double acc = 0.0;
#pragma unroll
                for (int j = 0; j < D::kStages; ++j)
                    if (D::B(j) != 0.0) acc = fma(D::B(j), K[j], acc);
                xn = fma(hs, acc, x);
            }
            K[D::kStages] = rhs(hit ? hk : s + hs, xn);
            const int snew = slast;
            double e5 = 0.0, e3 = 0.0;
#pragma unroll
            for (int j = 0; j <= D::kStages; ++j) {
                if (D::E5(j) != 0.0) e5 = fma(D::E5(j), K[j], e5);
                if (D::E3(j) != 0.0) e3 = fma(D::E3(j), K[j], e3);
            }
            const double isc = fast_rcp(fma(rtol, fmax(fabs(x), fabs(xn)), atol));
            double q5 = fabs(hs * e5) * isc;
            const double q3 = fabs(hs * e3) * isc;
            const double sw = __hiloint2double(shi, 0) - __hiloint2double(min(slo, shi), 0);
            if (sw > (double)B.swing) q5 = fmax(q5, fabs(hs) * sw * vm * isc);
            // Hairer's 8(5,3) measure err5^2 / sqrt(err5^2 + 0.01 err3^2), in single precision
            // like the cost kernels (a step-size decision)
            const float n0 = (q5 <= 3.0e38) ? (float)q5 : CUDART_INF_F;     // NaN -> inf
            const float n1 = (q3 <= 3.0e38) ? (float)q3 : CUDART_INF_F;
            float err = CUDART_INF_F;
            if (n0 < 1.0e18f && n1 < 1.0e18f) {
                const float den = fmaf(n0, n0, 0.01f * n1 * n1);
                err = den > 0.0f ? n0 * n0 * rsqrtf(den) : 0.0f;
            }
            const float fac0 = 0.9f * exp2f(-0.125f * __log2f(fminf(fmaxf(err, 1e-30f), 3.0e38f)));
            if (inside) {
                if (err <= 1.0f) {
                    ++nacc;
                    s = hit ? hk : s + hs;
                    x = xn;
                    sprev = snew;
                    k1 = K[D::kStages];
                    float fac = fminf(LASSIM_FACMAX, fmaxf(0.2f, fac0));
                    if (rejected) fac = fminf(fac, 1.0f);
                    rejected = false;
                    const double hnew = hs * (double)fac;
                    h = hit ? fmax(hnew, h) : hnew;
                    if (hit) inside = false;
                } else {
                    ++nrej;
                    rejected = true;
                    h = hs * (double)fmaxf(0.2f, fminf(fac0, 1.0f));
                    if (!(fabs(h) > 1e-290) || s + h == s) { st |= 2; inside = false; }
                }
                if (inside && nacc + nrej >= B.max_steps) { st |= 1; inside = false; }
            }
        }
        if (samp >= 0 && live) {
            const double v = st ? CUDART_NAN : x;
            B.Ys[(size_t)samp * B.N + i] = v;
            bad |= !(v == v);
            xmax = fmax(xmax, v);
        }
    }
    if (!live) return;
    // time-series cost of the peripheral state (core_problem.py:89-90 applied to one gene)
    const double m = bad ? CUDART_NAN : xmax;          // numpy.amax propagates NaN
    const double s2 = B.psigma2[gene];
    double cost = 0.0;
    for (int tsi = 0; tsi < T; ++tsi) {
        const double d = B.pdata[(size_t)gene * T + tsi] - B.Ys[(size_t)tsi * B.N + i] / m;
        cost += d * d / s2;
    }
    B.cost[i] = cost;
    B.status[i] = st;
    B.steps[i] = nacc + nrej;
    B.counters[(size_t)i * 3 + 0] = nacc;
    B.counters[(size_t)i * 3 + 1] = nrej;
    B.counters[(size_t)i * 3 + 2] = 13ull * (nacc + nrej) + 2ull;
}

}  // namespace lassim
