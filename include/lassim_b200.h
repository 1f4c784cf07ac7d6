/*
 * lassim_b200 - C ABI of the B200 cost-evaluation engine for LASSIM core
 * problems.  This is the drop-in boundary: plain pointers and sizes, FP64,
 * row-major, caller-owned buffers, no exceptions, no torch types.
 *
 * Every entry point returns 0 on success and a negative LASSIM_E_* code on
 * failure; lassim_last_error() gives the message of the last failure on that
 * handle (or of the last failed lassim_create when called with NULL).
 *
 * A handle belongs to one (process, device).  It is not thread-safe.  Calls
 * block until their results are in the caller's buffers unless the name ends
 * in _device (those enqueue on the given stream and return).
 *
 * Reference interfaces replaced (paths in gmariotti/lassim):
 *   source/core/functions/common_functions.pxd:17-23        ode plugin
 *   source/core/functions/perturbation_functions.pxd:10-14  knock-down plugin
 *   source/core/core_problem.py:75-92, 134-155              _objfun_impl
 *   source/core/core_problem.py:189-245                     factory / build
 */
#ifndef LASSIM_B200_H
#define LASSIM_B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct lassim_handle lassim_handle;

enum {
    LASSIM_OK = 0,
    LASSIM_E_INVALID = -1,   /* bad argument (NULL, size, shape)            */
    LASSIM_E_CUDA = -2,      /* a CUDA runtime call failed                  */
    LASSIM_E_STATE = -3,     /* call order (no problem set, ...)            */
    LASSIM_E_UNSUPPORTED = -4 /* problem size outside what the kernels cover */
};

/* Per-candidate status written by the evaluation (bit field, 0 = ok). */
enum {
    LASSIM_ST_OK = 0,
    LASSIM_ST_STEP_BUDGET = 1, /* a trajectory used max_steps step attempts */
    LASSIM_ST_NONFINITE = 2    /* step size underflow / non-finite state    */
};

/* Library version, "major.minor.patch". */
const char *lassim_version(void);

/* Create / destroy an engine on CUDA device `device`. */
int lassim_create(int device, lassim_handle **out);
int lassim_destroy(lassim_handle *h);
const char *lassim_last_error(const lassim_handle *h);

/*
 * Problem constants; replaces CoreProblemFactory.new_instance + build
 * (source/core/core_problem.py:168-187, 219-245).
 *   n        number of transcription factors (state size)
 *   T        number of sample times, times[0] is the start (state = y0)
 *   times[T], y0[n], data[T*n] (row = sample time), sigma[n]
 *   mask[n*n] row-major, non-zero where gene `row` is regulated by gene `col`
 *            (the reference's vector_map_mask, source/customs/
 *            core_functions.py:38-60); the True entries in row-major order are
 *            the order of the k values in a decision vector
 *   pert     NULL, or [n*(n+2)]: n x n knock-down matrix followed by t_start,
 *            t_end per row (perturbation_functions.pyx:46-49)
 *   pert_factor  weight of the knock-down cost in the total
 * Decision vector layout: x = [lambda(n) | vmax(n) | k(nnz)], dim = 2n + nnz
 * (common_functions.pyx:80-82).
 */
int lassim_set_problem(lassim_handle *h, int n, int T, const double *times,
                       const double *y0, const double *data,
                       const double *sigma, const unsigned char *mask,
                       const double *pert, double pert_factor);

/* Replace only the regulation mask (each edge-pruning iteration rebuilds it,
 * source/customs/core_functions.py:101-120).  dim becomes 2n + popcount. */
int lassim_set_mask(lassim_handle *h, const unsigned char *mask);

/* Integrator controls.  Defaults mirror scipy.integrate.odeint as the
 * reference calls it (common_functions.pyx:86-90): rtol = atol = 1.49012e-8,
 * error weight rtol*|y| + atol, max-norm.  max_steps bounds the step attempts
 * of one trajectory (default 20,000,000; the reference's mxstep=1e8 would let
 * one candidate hold a GPU for minutes). */
int lassim_set_tolerances(lassim_handle *h, double rtol, double atol,
                          long long max_steps);

/* Integrator: LASSIM_METHOD_DOP853 = Dormand-Prince 8(5,3) (default: about 5x
 * fewer steps than the 5(4) pair at these tolerances and closer to the
 * converged answer) or LASSIM_METHOD_DP45 = Dormand-Prince 5(4).  Both use the
 * same error weights, norm and tolerances.  (The reference has one integrator,
 * LSODA - variable order up to 12 - common_functions.pyx:86-90.) */
enum { LASSIM_METHOD_DP45 = 0, LASSIM_METHOD_DOP853 = 1 };
int lassim_set_method(lassim_handle *h, int method);

/* Jump guard of the explicit pairs.  Their embedded error estimates cannot see a logistic
 * term that switches early in a step (the stages there carry no weight in the solution or
 * in its embedded companions; with the 8(5,3) pair that is the first quarter of the step),
 * and a knocked-down gene that grows like exp(|lambda'| t) drives the regulator sums of its
 * targets through zero within 1e-50 time units: a discontinuity as far as any step is
 * concerned.  A step in which some logistic term swings by more than `swing` between two
 * stages is therefore charged hs * vmax * swing as its error: it shrinks until the switch
 * is resolved or, for a true discontinuity, located to the tolerance (what LSODA's
 * controller does through rejections).  Default 0.5; 2 or more switches the guard off
 * (then 3 % of uniform Th2 candidates end a knock-down solve with a state off by 1e-5 to
 * 1e-2, as SciPy's own DOP853 does on 18 %).  LASSIM_JUMP_GUARD in the environment sets the
 * initial value.  (No reference counterpart.) */
int lassim_set_jump_guard(lassim_handle *h, double swing);

/* Kernel for the time-series trajectories of problems of at most 16 genes.  A
 * small batch waits for its longest time series (a limit cycle needs 15,000+
 * step attempts where most candidates need 600), so these modes trade lanes for
 * fewer SEQUENTIAL right-hand sides on such trajectories:
 *   0  the standard kernel (16 lanes per trajectory, the pair lassim_set_method names)
 *   1  speculative attempts: one warp per time series; every attempt is made at the
 *      proposed step size and, by the other half of the warp, at 0.7 of it; when
 *      the full step fails its error test and the shorter one passes, the shorter
 *      one is accepted instead of starting over
 *   3  stage-parallel: one warp per time series; a 5-stage Gauss collocation pair
 *      (4-stage above 12 genes) solved by fixed-point iteration with its stages
 *      side by side in the warp, 6 (5) sequential right-hand sides per attempt
 *      instead of 12, same tolerances, weights and norm; steps are also limited so
 *      that no logistic term swings by more than 0.5 inside one (csrc/pirk_kernel.cuh)
 *   4  the same with a second warp attempting 0.7 of every step (as in 1)
 *   6  the same with three warps attempting 1.2, 1 and 0.7 of the proposed step
 *      (the largest that passes its error test is taken)
 *   5  hand-over: every time series starts in the standard kernel; those still
 *      going after a cap of attempts (lassim_set_series_cap) stop at the step they are
 *      on and continue in kernel 6 (kernel 4 above 8192 candidates), launched behind it
 *      on the same stream, longest first (a one-CTA kernel in between orders the
 *      hand-over list by the attempts each entry still has ahead of it), one trajectory
 *      per CTA while SMs last.  The many short time series stay cheap, the few long ones
 *      get the fast kernel almost to themselves.  In kernel 6 / 4 a time series continued
 *      this way has its error estimate (the last correction of the iteration) scaled by
 *      the rate at which the corrections shrink, rho / (1 - rho) within [0.1, 1]: the
 *      last correction is the error of the iterate before the one that is kept
 *   2  automatic (default): 5 for batches of at most 16,384 candidates, else 0
 * Results are deterministic for a given mode and stay inside the same error control.
 * In the automatic mode a candidate's cost does not depend on the batch it is evaluated
 * in, as long as the batch has at most 4096 candidates: alone (a batch of one,
 * _objfun_impl) or inside a population it takes the same path and gets the same bits.
 * Larger batches are scheduled for throughput (the cap follows the batch size) and agree
 * at the level of the integration tolerance.  The knock-down trajectories always run on
 * the standard kernel, concurrently on a second stream; behind a hand-over (batches of 64
 * and more) that kernel starts when the capped kernel is through and the CTAs of the
 * stage-parallel kernel are in place (a one-thread gate kernel waits for them).
 * (No reference counterpart.)
 * LASSIM_SPECULATE in the environment sets the initial mode. */
int lassim_set_speculation(lassim_handle *h, int mode);

/* Attempts a time series makes in the standard kernel before mode 5 hands it over;
 * 0 (default) = automatic: 768 for batches of at most 4096 candidates (four times that
 * for the 5(4) pair), three per four candidates up to 8000 beyond.  LASSIM_SERIES_CAP in
 * the environment sets the initial value.  (No reference counterpart.) */
int lassim_set_series_cap(lassim_handle *h, unsigned int attempts);

/* Current decision-vector length (2n + nnz), trajectories per candidate
 * (1 + unique time pairs + n when a knock-down matrix is set, else 1). */
int lassim_dim(const lassim_handle *h);
int lassim_trajectories_per_candidate(const lassim_handle *h);

/*
 * Evaluate N candidates.  X[N*dim] in HOST memory; outputs in host memory.
 * cost[N] = pert_factor * pert_cost + cost_main  (core_problem.py:155).
 * Optional (may be NULL): cost_main[N], pert_cost[N], status[N],
 * counters[N*3] = {accepted steps, rejected steps, RHS evaluations} summed
 * over the candidate's trajectories.
 * Replaces N calls of CoreWithPerturbationsProblem._objfun_impl
 * (core_problem.py:134-155) / CoreProblem._objfun_impl (:75-92).
 * pert_cost is the residual of min_a |a r - p|^2 (perturbation_functions.pyx:93-102)
 * computed as p.p - (r.p)^2 / (r.r) with r scaled by its largest entry.  One deviation:
 * where an end-state ratio overflows to -inf (a control state that decayed to 1e-190 under
 * a knocked one of 1e120; one row in a thousand of a uniform Th2 population) the toolbox's
 * numpy.linalg.lstsq raises LinAlgError; here the residual takes its limit for that
 * entry -> -inf, p.p - p_k^2.  A NaN ratio gives a NaN cost, as there.
 */
int lassim_eval(lassim_handle *h, int N, const double *X, double *cost,
                double *cost_main, double *pert_cost, int *status,
                unsigned long long *counters);

/* Same, with X and every non-NULL output in DEVICE memory, enqueued on
 * `stream` (a cudaStream_t passed as void*).  NULL selects the handle's own
 * non-blocking stream, NOT the legacy default stream: to run on the default
 * stream pass cudaStreamLegacy or cudaStreamPerThread explicitly.  Returns
 * after the launch; synchronise the stream before reading. */
int lassim_eval_device(lassim_handle *h, int N, const double *dX, double *dcost,
                       double *dcost_main, double *dpert_cost, int *dstatus,
                       unsigned long long *dcounters, void *stream);

/*
 * Time-series trajectories at the sample times: Y[N*T*n] (host), Y[c][0] = y0.
 * Replaces the ode plugin odeint1e8_lassim (common_functions.pyx:55-90) for a
 * batch; status[N] optional.
 */
int lassim_trajectories(lassim_handle *h, int N, const double *X, double *Y,
                        int *status);

/*
 * Peripheral genes fitted against a FIXED core (BASELINE.json config 5).  The
 * reference has no implementation of this (its README lists it as future
 * work, README.md:77-80), so the definition is by analogy with the core
 * problem: gene p obeys
 *     dx_p/dt = -lambda_p x_p + vmax_p / (1 + exp(-sum_j w_pj x_j(t)))
 * where x_j(t) is the core trajectory produced by `core_x` (a decision vector
 * of the problem currently set, which must have no more than 32 genes), from
 * (y0_core, y0_p) over the core's sample times, and the cost is the time-series
 * cost (core_problem.py:89-90) of the peripheral state alone:
 *     sum_t (data_p[t] - x_p(t)/max_t x_p)^2 / sigma_p^2.
 * The gene does not act back on the core, so lassim_set_peripheral integrates the
 * core ONCE (8(5,3) pair with its 7th-order continuous extension, a hundred times
 * tighter than the candidates' tolerances; an error here means core_x has no
 * finite trajectory) and every candidate is a scalar equation driven by that
 * trajectory: one thread per candidate, the 8(5,3) pair whatever
 * lassim_set_method says, same tolerances, error weight and controller as
 * everywhere.  lassim_set_peripheral_joint(h, 1) integrates the
 * joint (n+1)-state system per candidate in the cost kernel instead (core of at
 * most 31 genes; 13x the work for the Th2 core; kept as a cross-check).
 *   genes    number of peripheral genes
 *   data[genes*T], sigma[genes], y0[genes]
 *   mask[genes*n] non-zero where core gene j may regulate gene p (NULL = all)
 * A candidate is xp = [lambda_p, vmax_p, w_p(n)] (entries outside the mask are
 * ignored); gene[i] says which gene candidate i belongs to, so one call can
 * evaluate the populations of many genes.  Host buffers.
 */
int lassim_set_peripheral(lassim_handle *h, const double *core_x, int genes,
                          const double *data, const double *sigma, const double *y0,
                          const unsigned char *mask);
int lassim_set_peripheral_joint(lassim_handle *h, int joint);
int lassim_eval_peripheral(lassim_handle *h, int N, const int *gene, const double *Xp,
                           double *cost, int *status);
/* Joint trajectories Y[N*T*(n+1)] of the same candidates (last column = x_p). */
int lassim_peripheral_trajectories(lassim_handle *h, int N, const int *gene,
                                   const double *Xp, double *Y);

/*
 * Differential evolution on the device.  PyGMO's DE calls the objective once per
 * individual from C++ into Python (the archipelago of
 * source/core/base_optimization.py:159-203 drives it); here the populations of
 * all islands of this GPU live in device memory and one evolve round runs
 * `generations` x (trial vectors -> evaluation -> greedy selection) on the
 * handle's stream without a host round trip.  Semantics as pagmo 1.x `de`:
 * variants 1-10, generation-synchronous, out-of-box components re-drawn
 * uniformly, a trial wins only if strictly better, failed integrations cost +inf,
 * NaN never wins.
 *   lassim_de_init   X0[islands*individuals*dim] initial populations (host),
 *                    lb/ub[dim], individuals >= 6, f and cr in [0,1]
 *   lassim_de_evolve enqueues; lassim_de_get / _best synchronise
 *   lassim_de_best   table[islands*(dim+1)]: rows (x | f) of every island's best
 *   lassim_de_migrate_ring  PyGMO's default policy on a ring: each of this GPU's
 *                    islands (island j is row first_island + j*stride of a table of
 *                    n_table islands gathered by the caller across GPUs; islands
 *                    are dealt round-robin to ranks) takes its two neighbours'
 *                    bests, each replacing the island's worst if strictly better
 *   lassim_de_set_individual  seed one individual (e.g. a known solution)
 *   lassim_de_set_island_ids  where this GPU's islands sit in the whole archipelago
 *                    (local island j is island first_island + j * stride): keys the
 *                    random streams so that islands on different GPUs draw different
 *                    numbers from one seed.  Default 0, 1.
 *   lassim_de_set_lanes  islands do not wait for each other inside an evolve round
 *                    (PyGMO evolves each island in a worker of its own and joins at the
 *                    end, base_optimization.py:173-174): the islands are split into
 *                    `lanes` groups, each advancing through its generations on its own
 *                    stream with its own scratch; a slow candidate holds up its own
 *                    island's next generation only.  0 = automatic (one island per lane,
 *                    at most 16 lanes while the islands hold at most 1024 individuals
 *                    together, at most 8 up to 2048, 4 up to 3072; lockstep beyond, where
 *                    one batch per generation fills the machine by itself and lanes would
 *                    take SMs from each other); 1 = all islands in lockstep.  Up to 4096
 *                    individuals the result does not depend on the number of lanes.
 *                    LASSIM_DE_LANES in the environment sets the initial value.
 */
int lassim_de_init(lassim_handle *h, int islands, int individuals, int variant, double f,
                   double cr, unsigned long long seed, const double *lb, const double *ub,
                   const double *X0);
int lassim_de_set_individual(lassim_handle *h, int island, int index, const double *x);
int lassim_de_set_island_ids(lassim_handle *h, int first_island, int stride);
int lassim_de_set_lanes(lassim_handle *h, int lanes);
int lassim_de_evolve(lassim_handle *h, int generations);
int lassim_de_best(lassim_handle *h, double *table);
int lassim_de_migrate_ring(lassim_handle *h, const double *table, int n_table, int first_island,
                           int stride);
int lassim_de_get(lassim_handle *h, double *X, double *F);
long long lassim_de_evaluations(const lassim_handle *h);

/*
 * End states of the knock-down part of the LAST lassim_eval / _eval_device
 * call: E[N*K*n], K = lassim_trajectories_per_candidate(); slot 0 is unused
 * (time series), 1..U the controls in order of first appearance, 1+U+i the
 * knocked system of row i (perturbation_functions.pyx:64-85).  For tests and
 * diagnostics.
 */
int lassim_last_end_states(lassim_handle *h, int N, double *E);

/* Step attempts per trajectory of the LAST evaluation: steps[N*K] (after
 * lassim_trajectories or lassim_eval_peripheral only slot 0 of each candidate is
 * non-zero). */
int lassim_last_traj_steps(lassim_handle *h, int N, unsigned int *steps);

/*
 * Evaluate N candidates WITHOUT allocating anything (no counterpart in the
 * reference; added for hosts that cannot call cudaMalloc when they evaluate).
 * Same result as lassim_eval_device's cost: dX[N*dim] and dcost[N] are DEVICE
 * buffers the caller owns, the work is enqueued on `stream` (NULL = the
 * handle's own stream) and the call returns after the launch.
 *
 * Who needs it: a process whose GPU memory already belongs to another
 * framework's caching allocator (a library-side cudaMalloc would fail or
 * fragment that pool), a latency-critical loop that cannot take the
 * device-wide synchronisation cudaMalloc/cudaFree imply, a job held to a hard
 * memory budget (MIG slice, MPS limit).  lassim_eval and lassim_eval_device
 * grow their scratch with cudaMalloc whenever a batch is larger than any
 * before; this entry point never calls cudaMalloc, cudaFree or any allocating
 * routine.
 *
 * Its scratch is a fixed 8 MiB array inside the library image (each GPU that
 * loads the library holds those 8 MiB until the process exits).  A batch needs
 * about 8*N*(T*n + K*n) + 4*N*K + 56*N bytes (K = trajectories per candidate);
 * for the Th2 example that is ~2.2 KB per candidate, i.e. up to ~3,700
 * candidates per call.  Larger batches are rejected with LASSIM_E_INVALID and
 * a message that names both sizes: split them or use lassim_eval_device.
 *
 * The first call on a device also looks up the array's device address; make
 * that first call before starting any stream capture.  The problem must have
 * been set (lassim_set_problem allocates; do it during set-up).  All calls in
 * a process share the one array per GPU and it keeps the previous call's
 * contents (the entry point clears the part that must start at zero), so two
 * calls whose kernels overlap on the same GPU - different streams or different
 * handles - race on it: order them.  lassim_last_end_states / _traj_steps do
 * not see this path's scratch.
 */
int lassim_eval_device_noalloc(lassim_handle *h, int N, const double *dX,
                               double *dcost, void *stream);

/* Algorithmic FP64 flops of one step attempt of one trajectory with the current
 * problem and integrator (roofline numerator): 12*nnz + 90*n for
 * the 5(4) pair, 24*nnz + 212*n for the 8(5,3) pair.  Name of the kernel the
 * current problem dispatches to.  (No reference counterpart.) */
long long lassim_flops_per_attempt(const lassim_handle *h);
const char *lassim_kernel_name(lassim_handle *h);

/* The same two for the kernel that integrates the TIME SERIES of a batch of N
 * candidates under the current lassim_set_speculation mode: equal to the above
 * when the standard kernel does; twice the flops for speculative attempts (both
 * halves of the warp work); for the stage-parallel Gauss pair with S stages and
 * S corrections, (S+1)*S right-hand sides of 2*nnz + 5*n flops plus the stage
 * sums, predictor and estimate: 60*nnz + 474*n for S = 5, times the warps that
 * work on a trajectory (2 in mode 4, 3 in modes 5 and 6).  (No reference
 * counterpart.) */
long long lassim_series_flops_per_attempt(const lassim_handle *h, int N);
const char *lassim_series_kernel_name(lassim_handle *h, int N);

/* Attempts made by the stage-parallel kernel in the last evaluation, summed over
 * its trajectories (with mode 5 a time series makes its first attempts in the
 * standard kernel; lassim_last_traj_steps reports the sum of both).  Blocks until
 * the evaluation has finished.  (No reference counterpart.) */
int lassim_last_stage_parallel_attempts(lassim_handle *h, unsigned long long *attempts);

/* Device time of the stage-parallel kernel of the last evaluation in milliseconds
 * (CUDA events on the launching stream around that one launch); LASSIM_E_STATE if
 * the last evaluation did not use it.  Blocks until it has finished.  (No reference
 * counterpart.) */
int lassim_last_stage_parallel_ms(lassim_handle *h, float *ms);

/* Device time from the start of the last evaluation to the launch of its stage-parallel
 * kernel in milliseconds: with the hand-over (mode 5) the phase in which every time series
 * makes its first attempts in the standard kernel.  LASSIM_E_STATE as above.  (No reference
 * counterpart.) */
int lassim_last_capped_phase_ms(lassim_handle *h, float *ms);

/* Number of kernels this handle has launched so far (an evaluation is one launch, two
 * concurrent ones with a speculative time-series kernel, five with the hand-over: capped
 * kernel, list order, stage-parallel kernel, gate, knock-down kernel). */
long long lassim_launch_count(const lassim_handle *h);

/* The handle's own stream (a cudaStream_t as void*): what NULL selects in the _device
 * entry points and what lassim_de_evolve forks from and joins; record events on it to
 * time device-resident work, or make other streams wait for it. */
void *lassim_stream(const lassim_handle *h);

/* Device time of the last evaluation kernel in milliseconds (CUDA events on
 * the launching stream); blocks until that kernel has finished. */
int lassim_last_kernel_ms(lassim_handle *h, float *ms);

/* FP64 FMA throughput probe: runs a register-resident DFMA kernel for about
 * `iters` dependent FMAs per thread and returns achieved TFLOP/s (2 flops per
 * FMA).  Used as the roofline denominator (no reference counterpart). */
int lassim_fp64_peak_probe(lassim_handle *h, int iters, double *tflops);

#ifdef __cplusplus
}
#endif
#endif /* LASSIM_B200_H */
