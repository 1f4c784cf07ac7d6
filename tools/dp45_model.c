/*
 * DEVELOPMENT TOOL, not product code and not the oracle.
 *
 * Scalar C model of the step controller the CUDA kernel uses (Dormand-Prince
 * 5(4), LSODA-style error weights ewt = rtol*|y| + atol, max-norm).  It exists
 * so that tolerances, step counts and parity against the golden vectors can be
 * studied in the GPU-less build container.  Nothing in lassim_b200 links or
 * loads it; tests use it only to cross-check the device counters.
 *
 * gcc -O2 -shared -fPIC tools/dp45_model.c -o tools/_dp45_model.so -lm
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define MAXN 256

typedef struct {
    int n;
    const double *lam, *vmax, *W; /* W dense n*n row-major */
    long nrhs;
} sys_t;

static void rhs(sys_t *s, const double *y, double *f) {
    int n = s->n;
    for (int i = 0; i < n; ++i) {
        double acc = 0.0;
        const double *w = s->W + (size_t)i * n;
        for (int j = 0; j < n; ++j) acc += w[j] * y[j];
        f[i] = -s->lam[i] * y[i] + s->vmax[i] / (1.0 + exp(-acc));
    }
    s->nrhs++;
}

static const double A21 = 1.0 / 5.0;
static const double A31 = 3.0 / 40.0, A32 = 9.0 / 40.0;
static const double A41 = 44.0 / 45.0, A42 = -56.0 / 15.0, A43 = 32.0 / 9.0;
static const double A51 = 19372.0 / 6561.0, A52 = -25360.0 / 2187.0,
                    A53 = 64448.0 / 6561.0, A54 = -212.0 / 729.0;
static const double A61 = 9017.0 / 3168.0, A62 = -355.0 / 33.0,
                    A63 = 46732.0 / 5247.0, A64 = 49.0 / 176.0,
                    A65 = -5103.0 / 18656.0;
static const double B1 = 35.0 / 384.0, B3 = 500.0 / 1113.0, B4 = 125.0 / 192.0,
                    B5 = -2187.0 / 6784.0, B6 = 11.0 / 84.0;
static const double E1 = 71.0 / 57600.0, E3 = -71.0 / 16695.0,
                    E4 = 71.0 / 1920.0, E5 = -17253.0 / 339200.0,
                    E6 = 22.0 / 525.0, E7 = -1.0 / 40.0;

/* Integrate from ts[0] to ts[nt-1], sampling at every ts[i]; Y is nt*n.
 * Returns status: 0 ok, 1 step budget exhausted, 2 non-finite / h underflow.
 * counters[0]=accepted, [1]=rejected, [2]=rhs evals. */
int dp45_integrate(int n, const double *lam, const double *vmax,
                   const double *W, const double *y0, int nt, const double *ts,
                   double rtol, double atol, long max_steps, double beta,
                   double *Y, long *counters) {
    sys_t s = {n, lam, vmax, W, 0};
    double y[MAXN], yn[MAXN], ys[MAXN];
    double k1[MAXN], k2[MAXN], k3[MAXN], k4[MAXN], k5[MAXN], k6[MAXN], k7[MAXN];
    long nacc = 0, nrej = 0;
    int status = 0;
    memcpy(y, y0, sizeof(double) * n);
    memcpy(Y, y0, sizeof(double) * n);
    double t = ts[0];
    rhs(&s, y, k1);
    /* initial step (Hairer, max-norm) */
    double d0 = 0, d1 = 0;
    for (int i = 0; i < n; ++i) {
        double sc = atol + rtol * fabs(y[i]);
        d0 = fmax(d0, fabs(y[i]) / sc);
        d1 = fmax(d1, fabs(k1[i]) / sc);
    }
    double h0 = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
    for (int i = 0; i < n; ++i) ys[i] = y[i] + h0 * k1[i];
    rhs(&s, ys, k2);
    double d2 = 0;
    for (int i = 0; i < n; ++i) {
        double sc = atol + rtol * fabs(y[i]);
        d2 = fmax(d2, fabs(k2[i] - k1[i]) / sc);
    }
    d2 /= h0;
    double dm = fmax(d1, d2);
    double h1 = (dm <= 1e-15) ? fmax(1e-6, h0 * 1e-3) : pow(0.01 / dm, 0.2);
    double h = fmin(100 * h0, h1);
    double errold = 1e-4;
    int rejected = 0;

    for (int it = 1; it < nt && status == 0; ++it) {
        double tend = ts[it];
        while (1) {
            if (nacc + nrej >= max_steps) { status = 1; break; }
            int hit = 0;
            double hs = h;
            if (t + 1.01 * hs >= tend) { hs = tend - t; hit = 1; }
            for (int i = 0; i < n; ++i) ys[i] = y[i] + hs * (A21 * k1[i]);
            rhs(&s, ys, k2);
            for (int i = 0; i < n; ++i) ys[i] = y[i] + hs * (A31 * k1[i] + A32 * k2[i]);
            rhs(&s, ys, k3);
            for (int i = 0; i < n; ++i) ys[i] = y[i] + hs * (A41 * k1[i] + A42 * k2[i] + A43 * k3[i]);
            rhs(&s, ys, k4);
            for (int i = 0; i < n; ++i) ys[i] = y[i] + hs * (A51 * k1[i] + A52 * k2[i] + A53 * k3[i] + A54 * k4[i]);
            rhs(&s, ys, k5);
            for (int i = 0; i < n; ++i) ys[i] = y[i] + hs * (A61 * k1[i] + A62 * k2[i] + A63 * k3[i] + A64 * k4[i] + A65 * k5[i]);
            rhs(&s, ys, k6);
            for (int i = 0; i < n; ++i) yn[i] = y[i] + hs * (B1 * k1[i] + B3 * k3[i] + B4 * k4[i] + B5 * k5[i] + B6 * k6[i]);
            rhs(&s, yn, k7);
            double err = 0;
            int bad = 0;
            for (int i = 0; i < n; ++i) {
                double e = hs * (E1 * k1[i] + E3 * k3[i] + E4 * k4[i] + E5 * k5[i] + E6 * k6[i] + E7 * k7[i]);
                double sc = atol + rtol * fmax(fabs(y[i]), fabs(yn[i]));
                double q = fabs(e) / sc;
                if (!(q == q) || isinf(q)) bad = 1;
                err = fmax(err, q);
            }
            if (bad) {
                nrej++;
                h = hs * 0.2;
                rejected = 1;
                if (!(fabs(h) > 1e-300) || t + h == t) { status = 2; break; }
                continue;
            }
            if (err <= 1.0) {
                nacc++;
                t = hit ? tend : t + hs;
                memcpy(y, yn, sizeof(double) * n);
                memcpy(k1, k7, sizeof(double) * n);
                double fac;
                if (err < 1e-30) fac = 10.0;
                else {
                    fac = 0.9 * pow(err, -(0.2 - 0.75 * beta)) * pow(errold, beta);
                    fac = fmin(10.0, fmax(0.2, fac));
                }
                if (rejected) fac = fmin(fac, 1.0);
                rejected = 0;
                errold = fmax(err, 1e-4);
                double hnew = hs * fac;
                /* a step clipped to land on a sample time must not shrink h */
                h = hit ? fmax(hnew, h) : hnew;
                if (hit) break;
            } else {
                nrej++;
                double fac = fmax(0.2, 0.9 * pow(err, -0.2));
                h = hs * fac;
                rejected = 1;
                if (t + h == t) { status = 2; break; }
            }
        }
        memcpy(Y + (size_t)it * n, y, sizeof(double) * n);
    }
    counters[0] = nacc; counters[1] = nrej; counters[2] = s.nrhs;
    return status;
}
