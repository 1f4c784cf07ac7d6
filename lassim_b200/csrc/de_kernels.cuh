// Differential evolution on the device: trial-vector generation, greedy
// selection and ring migration for I islands of NP individuals, so that a whole
// evolve round (propose -> grn_cost_kernel -> select, `gen` times) runs without
// a host round trip.  Semantics are those of lassim_b200/pygmo/algorithm.py
// (pagmo 1.x `de`: variants 1-10, generation-synchronous, out-of-box components
// re-drawn uniformly, a trial wins only if strictly better, NaN never wins).
// One thread per individual; the work is tiny next to the cost evaluation.
#pragma once
#include <cuda_runtime.h>
#include <curand_kernel.h>
#include <math_constants.h>

namespace lassim {

struct DeParams {
    int islands, np, dim, variant;
    double f, cr;
    unsigned long long seed;
    unsigned long long generation;   // advances the random stream
    // position of this GPU's islands in the whole archipelago (local island j is island
    // gid_base + j * gid_stride): keys the random streams, so that islands on different
    // GPUs draw different numbers from the same seed
    int gid_base = 0, gid_stride = 1;
    // individuals [ind_begin, ind_end) of the islands * np this launch works on (a lane of
    // the asynchronous evolve handles a group of consecutive islands)
    int ind_begin = 0, ind_end = 0;
};

// cost with "worse than anything" for NaN
__device__ __forceinline__ double de_key(double v) { return (v != v) ? CUDART_INF : v; }

// trial[I*NP*D] from X[I*NP*D], F[I*NP]
__global__ void de_trial_kernel(const DeParams p, const double *__restrict__ X,
                                const double *__restrict__ F, const double *__restrict__ lb,
                                const double *__restrict__ ub, double *__restrict__ trial) {
    const int gi = p.ind_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (gi >= p.ind_end) return;
    const int isl = gi / p.np, i = gi - isl * p.np;
    const int NP = p.np, D = p.dim;
    const double *Xi = X + (size_t)isl * NP * D;
    const double *Fi = F + (size_t)isl * NP;
    // one Philox subsequence per individual of the whole archipelago; a generation draws at
    // most 5 partners (+ rare repeats), one start index, D crossover uniforms and D redraws,
    // two 32-bit numbers per double: the stream advances by 4 D + 64 per generation
    curandStatePhilox4_32_10_t rng;
    const unsigned long long gisl = (unsigned long long)(p.gid_base + isl * p.gid_stride);
    curand_init(p.seed, gisl * (unsigned long long)NP + (unsigned long long)i,
                p.generation * (unsigned long long)(4 * D + 64), &rng);

    int best = 0;
    double fb = de_key(Fi[0]);
    for (int j = 1; j < NP; ++j) {
        const double v = de_key(Fi[j]);
        if (v < fb) { fb = v; best = j; }
    }
    // five mutually distinct partners, all different from i
    int r[5];
    for (int a = 0; a < 5; ++a) {
        bool again;
        do {
            r[a] = (int)(curand(&rng) % (unsigned)(NP - 1));
            r[a] += (r[a] >= i);
            again = false;
            for (int b = 0; b < a; ++b) again |= (r[b] == r[a]);
        } while (again);
    }
    const double *xi = Xi + (size_t)i * D, *xb = Xi + (size_t)best * D;
    const double *x1 = Xi + (size_t)r[0] * D, *x2 = Xi + (size_t)r[1] * D;
    const double *x3 = Xi + (size_t)r[2] * D, *x4 = Xi + (size_t)r[3] * D;
    const double *x5 = Xi + (size_t)r[4] * D;
    const int kind = (p.variant - 1) % 5;
    const double Fw = p.f;
    auto mutant = [&](int j) -> double {
        switch (kind) {
            case 0: return xb[j] + Fw * (x2[j] - x3[j]);                       // best/1
            case 1: return x1[j] + Fw * (x2[j] - x3[j]);                       // rand/1
            case 2: return xi[j] + Fw * (xb[j] - xi[j]) + Fw * (x1[j] - x2[j]);   // rand-to-best/1
            case 3: return xb[j] + Fw * (x1[j] + x2[j] - x3[j] - x4[j]);       // best/2
            default: return x5[j] + Fw * (x1[j] + x2[j] - x3[j] - x4[j]);      // rand/2
        }
    };
    double *ti = trial + (size_t)gi * D;
    for (int j = 0; j < D; ++j) ti[j] = xi[j];
    auto place = [&](int j) {
        double v = mutant(j);
        if (!(v >= lb[j] && v <= ub[j]))
            v = lb[j] + (ub[j] - lb[j]) * curand_uniform_double(&rng);
        ti[j] = v;
    };
    if (p.variant <= 5) {
        // exponential crossover: a run of consecutive components from a random start,
        // continued while rand < CR (at least one component)
        int j = (int)(curand(&rng) % (unsigned)D), len = 0;
        do {
            place(j);
            j = (j + 1) % D;
            ++len;
        } while (len < D && curand_uniform_double(&rng) < p.cr);
    } else {
        // binomial crossover: each component with probability CR, one forced
        const int forced = (int)(curand(&rng) % (unsigned)D);
        for (int j = 0; j < D; ++j)
            if (j == forced || curand_uniform_double(&rng) < p.cr) place(j);
    }
}

// failed integrations cost +inf (what CoreProblem.objfun_batch does on the host)
__global__ void de_mask_status_kernel(int count, const int *__restrict__ status,
                                      double *__restrict__ cost) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count && status[i] != 0) cost[i] = CUDART_INF;
}

// Ft[0 .. ind_end - ind_begin): costs of the trials of individuals [ind_begin, ind_end)
__global__ void de_select_kernel(const DeParams p, double *__restrict__ X, double *__restrict__ F,
                                 const double *__restrict__ trial, const double *__restrict__ Ft) {
    const int li = blockIdx.x * blockDim.x + threadIdx.x;
    const int gi = p.ind_begin + li;
    if (gi >= p.ind_end) return;
    const double ft = Ft[li];
    if (ft < de_key(F[gi])) {    // false for a NaN trial: NaN never wins, and never stays
        F[gi] = ft;
        const double *t = trial + (size_t)gi * p.dim;
        double *x = X + (size_t)gi * p.dim;
        for (int j = 0; j < p.dim; ++j) x[j] = t[j];
    }
}

// best[I*(D+1)] rows (x | f) of every island
__global__ void de_best_kernel(const DeParams p, const double *__restrict__ X,
                               const double *__restrict__ F, double *__restrict__ best) {
    const int isl = blockIdx.x;
    if (threadIdx.x != 0 || isl >= p.islands) return;
    const double *Fi = F + (size_t)isl * p.np;
    int b = 0;
    double fb = de_key(Fi[0]);
    for (int j = 1; j < p.np; ++j) {
        const double v = de_key(Fi[j]);
        if (v < fb) { fb = v; b = j; }
    }
    const double *x = X + ((size_t)isl * p.np + b) * p.dim;
    double *out = best + (size_t)isl * (p.dim + 1);
    for (int j = 0; j < p.dim; ++j) out[j] = x[j];
    out[p.dim] = Fi[b];
}

// Ring migration from a snapshot of the islands' best individuals (rows x | f,
// `n_table` islands in the ring; this device's island j is table row
// first_island + j * stride - islands are dealt round-robin to ranks): each island
// receives its two
// neighbours' bests in turn (left, then right; for two islands the one other),
// each replacing the island's current worst if strictly better.
__global__ void de_migrate_ring_kernel(const DeParams p, double *__restrict__ X,
                                       double *__restrict__ F, const double *__restrict__ table,
                                       int n_table, int first_island, int stride) {
    const int isl = blockIdx.x;
    if (threadIdx.x != 0 || isl >= p.islands || n_table < 2) return;
    const int me = first_island + isl * stride;
    const int nb[2] = {(me + n_table - 1) % n_table, (me + 1) % n_table};
    const int count = (n_table == 2) ? 1 : 2;
    double *Fi = F + (size_t)isl * p.np;
    for (int a = 0; a < count; ++a) {
        const int src = (n_table == 2) ? nb[1] : nb[a];
        const double *row = table + (size_t)src * (p.dim + 1);
        const double fin = row[p.dim];
        int w = 0;
        double fw = de_key(Fi[0]);
        for (int j = 1; j < p.np; ++j) {
            const double v = de_key(Fi[j]);
            if (v > fw) { fw = v; w = j; }
        }
        if (fin < fw) {
            Fi[w] = fin;
            double *x = X + ((size_t)isl * p.np + w) * p.dim;
            for (int j = 0; j < p.dim; ++j) x[j] = row[j];
        }
    }
}

}  // namespace lassim
