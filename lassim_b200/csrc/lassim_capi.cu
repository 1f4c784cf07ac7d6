// Host side of liblassim_b200.so: the C ABI declared in include/lassim_b200.h.
// Owns the device copies of the problem constants, the scratch buffers and the
// launch logic for the kernels in grn_kernels.cuh.  There is no CPU fallback:
// every evaluation entry point launches a CUDA kernel or fails.
#include "../../include/lassim_b200.h"
#include "grn_kernels.cuh"
#include "pirk_kernel.cuh"
#include "pirk_wide_kernel.cuh"
#include "peripheral_kernels.cuh"
#include "de_kernels.cuh"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <tuple>
#include <vector>

using namespace lassim;

namespace {

std::string g_create_error;

template <typename T>
struct DevBuf {
    T *p = nullptr;
    size_t cap = 0;   // elements
    cudaError_t reserve(size_t count) {
        if (count <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, count * sizeof(T));
        if (e == cudaSuccess) cap = count;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

// Hand-over of long time series (series_policy below).  Up to kBatchIndependentMax
// candidates per call every choice that touches the numbers - which kernel integrates a
// time series, after how many attempts it is handed over, how many step sizes the
// stage-parallel kernel tries per attempt - is the same for every batch size, so a
// candidate's cost does not depend on the batch it is evaluated in.
constexpr int kBatchIndependentMax = 4096;
constexpr unsigned kHandoverCap = 768;       // attempts in the standard kernel (8(5,3) pair)

}  // namespace

// Everything ONE evaluation in flight writes to besides the caller's buffers, with the
// auxiliary stream its knock-down trajectories run on.  The handle owns one for the
// evaluation entry points and one per lane of the on-device differential evolution
// (islands evolve concurrently, each lane on its own stream with its own scratch).
struct Scratch {
    DevBuf<double> cost, cost_main, pert_cost, Y, E;
    DevBuf<unsigned char> zero;   // work counters | done | status | counters
    DevBuf<unsigned int> steps;
    DevBuf<unsigned long long> long_list;   // hand-over list of long time series
    cudaStream_t aux = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaError_t init(int prio) {
        cudaError_t e = cudaStreamCreateWithPriority(&aux, cudaStreamNonBlocking, prio);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming);
        return e;
    }
    void release() {
        cost.release(); cost_main.release(); pert_cost.release(); Y.release(); E.release();
        zero.release(); steps.release(); long_list.release();
        if (ev_fork) cudaEventDestroy(ev_fork);
        if (ev_join) cudaEventDestroy(ev_join);
        if (aux) cudaStreamDestroy(aux);
        ev_fork = ev_join = nullptr; aux = nullptr;
    }
};

// One lane of the on-device differential evolution: a group of consecutive islands that
// advances through its generations on its own stream.
struct DeLane {
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_done = nullptr;
    Scratch scratch;
    int first = 0, count = 0;     // local islands [first, first + count)
};

struct lassim_handle {
    int device = 0;
    int sm_count = 0;
    int prio_lo = 0, prio_hi = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t ev_sp0 = nullptr, ev_sp1 = nullptr;   // around the stage-parallel kernel
    bool sp_timed = false;
    int speculate = 2;        // lassim_set_speculation mode (2 = automatic)
    unsigned series_cap = 0;      // hand-over cap; 0 = automatic (series_cap_for)
    int handover_warps = 0;       // warps per trajectory of the kernel long time series move to; 0 = automatic
    int no_if = 0;                // LASSIM_NO_IF=1 (diagnostic): no integrating factor for growing genes
    bool no_order = false;        // LASSIM_NO_ORDER=1 (A/B): the hand-over list in the order of arrival
    int knock_after_cap = -1;     // LASSIM_KNOCK_AFTER_CAP (A/B): 1 = the knock-down kernel starts when the
                                  // stage-parallel kernel's CTAs are placed, 0 = at once, -1 = by batch size
    int knock_per_sm = 0;         // LASSIM_KNOCK_PER_SM=k (A/B): CTAs per SM of the knock-down kernel next to a hand-over
    float jump_guard = 0.5f;      // lassim_set_jump_guard
    bool timed = false;
    std::string err, kernel_name, series_kernel_name;
    long long launches = 0;
    // resident CTAs per SM of every (kernel, CTA size, dynamic shared memory) launched so far:
    // the attribute call and the occupancy query are made once per shape, not per launch
    std::map<std::tuple<const void *, int, size_t>, int> per_sm_cache;

    // problem (host copies)
    bool has_problem = false;
    int n = 0, T = 0, K = 1, U = 0, nnz = 0, dim = 0, max_deg = 0;
    int method = 1;   // kMethodDP45 = 0 / kMethodDOP853 = 1 (default)
    std::vector<double> pert_h;
    double pert_factor = 0.0;
    double rtol = 1.49012e-8, atol = 1.49012e-8;
    long long max_steps = 20000000;

    // problem (device)
    DevBuf<double> d_times, d_y0, d_data, d_sigma2, d_pert, d_traj_tt;
    DevBuf<int> d_kidx, d_row_ptr, d_col_idx, d_traj_prefix, d_knock_ctrl;

    // peripheral genes fitted against a fixed core (lassim_set_peripheral)
    int p_genes = 0;
    DevBuf<double> d_core_x, d_pdata, d_psigma2, d_py0;
    DevBuf<unsigned char> d_pmask;
    DevBuf<int> d_pgene;
    // ... the core integrated once (peripheral_kernels.cuh): its table of intervals with the
    // interpolant's coefficients, valid for the tolerances it was built with
    bool periph_joint = false;          // LASSIM_PERIPHERAL_JOINT=1: the joint (n+1)-state path
    DevBuf<double> d_core_table, d_core_samples, d_pYs;
    DevBuf<int> d_core_count, d_pstatus;
    DevBuf<unsigned int> d_psteps;
    DevBuf<unsigned long long> d_pcounters;
    int core_intervals = -1;            // -1: not built
    double core_rtol = 0.0, core_atol = 0.0;
    float core_swing = 0.0f;
    bool last_periph_march = false;

    // scratch of the evaluation entry points
    DevBuf<double> d_X, d_probe;
    Scratch main;
    int last_N = 0;
    bool last_had_E = false;

    // on-device differential evolution (lassim_de_*)
    bool de_ready = false, de_evaluated = false;
    DeParams de{};
    long long de_evaluations = 0;
    DevBuf<double> de_X, de_F, de_trial, de_best, de_table, de_lb, de_ub;
    int de_lanes_wanted = 0;             // 0 = automatic (lassim_de_set_lanes)
    std::vector<DeLane> de_lanes;
    cudaEvent_t ev_de_fork = nullptr;
};

namespace {

// Makes the handle's device current for the duration of a call and puts the
// caller's device back afterwards (a host application may be driving several).
struct DeviceGuard {
    int prev = -1;
    cudaError_t err;
    explicit DeviceGuard(int device) {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != device) err = cudaSetDevice(device);
        else if (err == cudaSuccess) prev = -1;      // already current: nothing to restore
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

int fail(lassim_handle *h, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (h) h->err = buf; else g_create_error = buf;
    return code;
}

#define CU(call)                                                               \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess)                                                 \
            return fail(h, LASSIM_E_CUDA, "%s failed: %s", #call,              \
                        cudaGetErrorString(e_));                               \
    } while (0)

template <typename T>
int upload(lassim_handle *h, DevBuf<T> &buf, const T *src, size_t count) {
    CU(buf.reserve(count ? count : 1));
    if (count)
        CU(cudaMemcpyAsync(buf.p, src, count * sizeof(T), cudaMemcpyHostToDevice,
                           h->stream));
    return LASSIM_OK;
}

int upload_mask(lassim_handle *h, const unsigned char *mask) {
    const int n = h->n;
    std::vector<int> kidx((size_t)n * n), row_ptr(n + 1, 0), col_idx;
    int cnt = 0, max_deg = 0;
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) {
            const bool on = mask[(size_t)i * n + j] != 0;
            kidx[(size_t)i * n + j] = on ? cnt++ : -1;
            if (on) col_idx.push_back(j);
        }
        row_ptr[i + 1] = cnt;
        if (row_ptr[i + 1] - row_ptr[i] > max_deg) max_deg = row_ptr[i + 1] - row_ptr[i];
    }
    h->nnz = cnt;
    h->dim = 2 * n + cnt;
    h->max_deg = max_deg;
    int rc;
    if (n <= 32) {   // the dense kernels read the n*n index map
        if ((rc = upload(h, h->d_kidx, kidx.data(), kidx.size()))) return rc;
    }
    if ((rc = upload(h, h->d_row_ptr, row_ptr.data(), row_ptr.size()))) return rc;
    if ((rc = upload(h, h->d_col_idx, col_idx.data(), col_idx.size()))) return rc;
    CU(cudaStreamSynchronize(h->stream));   // the vectors are locals
    return LASSIM_OK;
}

struct ZeroLayout {
    size_t off_done, off_status, off_counters, bytes;
};

ZeroLayout zero_layout(int N) {
    ZeroLayout z;
    z.off_done = 32;       // [0,16): work counters and hand-over count, [16,24): stage-parallel attempts,
                           // [24,28): stage-parallel CTAs that have started (start_gate_kernel)
    z.off_status = z.off_done + sizeof(int) * (size_t)N;
    z.off_counters = (z.off_status + sizeof(int) * (size_t)N + 15) & ~(size_t)15;
    z.bytes = z.off_counters + sizeof(unsigned long long) * 3 * (size_t)N;
    return z;
}

typedef void (*kernel_fn)(const DevProblem, const DevBatch);

struct KernelChoice {
    kernel_fn fn;
    int L, R;          // lanes per group, genes per lane
    const char *name;
    kernel_fn fn_shadow = nullptr;   // warp-per-trajectory speculative variant, if built
    kernel_fn fn_pirk = nullptr;     // warp-per-trajectory stage-parallel variant, if built
    size_t pirk_smem = 0;
    kernel_fn fn_pirk2 = nullptr;    // the same with a second warp attempting a shorter step
    size_t pirk2_smem = 0;
    kernel_fn fn_pirk3 = nullptr;    // three warps, three step sizes
    size_t pirk3_smem = 0;
    kernel_fn fn_capped = nullptr;   // fn with the hand-over cap compiled in
    kernel_fn fn_wide = nullptr;     // CTA-per-trajectory stage-parallel kernel (17 .. 64 genes)
    int wide_threads = 0;
    size_t wide_smem = 0;
};

// dynamic shared memory of one CTA: header (2^(j/256) table, reduction scratch)
// + the groups' double-buffered exchange lines
size_t smem_bytes(const KernelChoice &kc) {
    const int groups_per_cta = threads_for(kc.L) / kc.L;
    return ((size_t)kSmemHeader + (size_t)groups_per_cta * 2 * (kc.L * kc.R + 2)) * sizeof(double);
}

template <int L, int NC>
kernel_fn dense_kernel(int method, bool fast) {
    if (method == kMethodDOP853) return grn_cost_kernel<L, 1, DenseW<1, NC>, kMethodDOP853, true>;
    if (!fast) return grn_cost_kernel<L, 1, DenseW<1, NC>, kMethodDP45, false>;
    return grn_cost_kernel<L, 1, DenseW<1, NC>, kMethodDP45, true>;
}

// the standard kernel with the attempt cap compiled in (time series of a hand-over evaluation)
template <int NC>
kernel_fn handover_kernel(int method) {
    if (method == kMethodDOP853)
        return grn_cost_kernel<16, 1, DenseW<1, NC>, kMethodDOP853, true, false, false, true>;
    return grn_cost_kernel<16, 1, DenseW<1, NC>, kMethodDP45, true, false, false, true>;
}

template <int NC>
kernel_fn shadow_kernel(int method) {
    if (method == kMethodDOP853)
        return grn_cost_kernel<16, 1, DenseW<1, NC>, kMethodDOP853, true, false, true>;
    return grn_cost_kernel<16, 1, DenseW<1, NC>, kMethodDP45, true, false, true>;
}

// stage-parallel time-series kernel: 5 Gauss stages x 6 lanes for n <= 12, 4 x 8 up to 16
void pirk_kernel(int n, KernelChoice *out) {
    if (n <= 12) {
        out->fn_pirk = grn_series_pirk_kernel<5, 6, 12, true, 1>;
        out->pirk_smem = pirk_smem_doubles(5, 6, 12, 1) * sizeof(double);
        out->fn_pirk2 = grn_series_pirk_kernel<5, 6, 12, true, 2>;
        out->pirk2_smem = pirk_smem_doubles(5, 6, 12, 2) * sizeof(double);
        out->fn_pirk3 = grn_series_pirk_kernel<5, 6, 12, true, 3>;
        out->pirk3_smem = pirk_smem_doubles(5, 6, 12, 3) * sizeof(double);
    } else {
        out->fn_pirk = grn_series_pirk_kernel<4, 8, 16, true, 1>;
        out->pirk_smem = pirk_smem_doubles(4, 8, 16, 1) * sizeof(double);
        out->fn_pirk2 = grn_series_pirk_kernel<4, 8, 16, true, 2>;
        out->pirk2_smem = pirk_smem_doubles(4, 8, 16, 2) * sizeof(double);
        out->fn_pirk3 = grn_series_pirk_kernel<4, 8, 16, true, 3>;
        out->pirk3_smem = pirk_smem_doubles(4, 8, 16, 3) * sizeof(double);
    }
}

// the 17 .. 64-gene kernels with the hand-over cap compiled in, and the stage-parallel kernel
// their long time series continue in (5 Gauss stages x 32 or 64 lanes)
template <int NC>
void wide_dense(int method, KernelChoice *out) {
    out->fn_capped = method == kMethodDOP853
        ? grn_cost_kernel<32, 1, DenseW<1, NC>, kMethodDOP853, true, false, false, true>
        : grn_cost_kernel<32, 1, DenseW<1, NC>, kMethodDP45, true, false, false, true>;
    out->fn_wide = grn_series_pirk_wide_kernel<5, 1, DenseW<1, NC>, true>;
    out->wide_threads = pirk_wide_threads(5, 1);
    out->wide_smem = pirk_wide_smem_doubles(5, 1) * sizeof(double);
}

void wide_sparse(int method, int max_deg, KernelChoice *out) {
    if (max_deg <= 6) {
        out->fn_capped = method == kMethodDOP853
            ? grn_cost_kernel<64, 1, EllW<1, 6>, kMethodDOP853, true, false, false, true>
            : grn_cost_kernel<64, 1, EllW<1, 6>, kMethodDP45, true, false, false, true>;
        out->fn_wide = grn_series_pirk_wide_kernel<5, 2, EllW<1, 6>, true>;
    } else if (max_deg <= 8) {
        out->fn_capped = method == kMethodDOP853
            ? grn_cost_kernel<64, 1, EllW<1, 8>, kMethodDOP853, true, false, false, true>
            : grn_cost_kernel<64, 1, EllW<1, 8>, kMethodDP45, true, false, false, true>;
        out->fn_wide = grn_series_pirk_wide_kernel<5, 2, EllW<1, 8>, true>;
    } else {
        out->fn_capped = method == kMethodDOP853
            ? grn_cost_kernel<64, 1, CsrW<1>, kMethodDOP853, true, false, false, true>
            : grn_cost_kernel<64, 1, CsrW<1>, kMethodDP45, true, false, false, true>;
        out->fn_wide = grn_series_pirk_wide_kernel<5, 2, CsrW<1>, true>;
    }
    out->wide_threads = pirk_wide_threads(5, 2);
    out->wide_smem = pirk_wide_smem_doubles(5, 2) * sizeof(double);
}

template <int L, int R>
kernel_fn sparse_kernel(int method, int max_deg) {
    // the two shapes of configs 3 and 4 (in-degree 5) also exist with 6 ELL slots per row:
    // a quarter fewer loads and FMAs in the mat-vec than with 8, a dozen registers less
    if constexpr ((L == 64 && R == 1) || (L == 128 && R == 2)) {
        if (max_deg <= 6) {
            if (method == kMethodDOP853) return grn_cost_kernel<L, R, EllW<R, 6>, kMethodDOP853, true>;
            return grn_cost_kernel<L, R, EllW<R, 6>, kMethodDP45, true>;
        }
    }
    if (max_deg <= 8) {
        if (method == kMethodDOP853) return grn_cost_kernel<L, R, EllW<R, 8>, kMethodDOP853, true>;
        return grn_cost_kernel<L, R, EllW<R, 8>, kMethodDP45, true>;
    }
    if (method == kMethodDOP853) return grn_cost_kernel<L, R, CsrW<R>, kMethodDOP853, true>;
    return grn_cost_kernel<L, R, CsrW<R>, kMethodDP45, true>;
}

// Which kernel serves a problem.  n <= 16: 16-lane groups, dense rows; n <= 32:
// one warp, dense rows; beyond: ELL rows (in-degree <= 8) or CSR, a 64-thread CTA with one
// gene per lane up to n = 64, then two genes per lane in CTAs of 64, 128, 256, 384 and 512
// threads (n <= 128, 256, 512, 768, 1024).
// LASSIM_GROUP=4 selects the 4-lane x 3-gene throughput shape for n <= 12 and
// LASSIM_EXACT_MATH=1 libm exp + IEEE division instead of the fast logistic
// (both diagnostics, DP45 only).
bool choose_kernel(const lassim_handle *h, KernelChoice *out) {
    const int n = h->n, method = h->method;
    const char *env = getenv("LASSIM_GROUP");
    const int want = env ? atoi(env) : 0;
    const char *ex = getenv("LASSIM_EXACT_MATH");
    const bool fast = !(ex && atoi(ex) != 0);
    const int nc = (n + 3) & ~3;
    const char *e8 = getenv("LASSIM_ELL8");               // A/B: 8 ELL slots also for in-degree <= 6
    const int deg = (e8 && atoi(e8) != 0 && h->max_deg < 7) ? 7 : h->max_deg;
    if (n <= 12 && want == 4) {
        if (method == kMethodDP45)
            *out = {grn_cost_kernel<4, 3, DenseW<3, 12>, kMethodDP45, true>, 4, 3, "dense 4x3"};
        else
            *out = {grn_cost_kernel<4, 3, DenseW<3, 12>, kMethodDOP853, true>, 4, 3, "dense 4x3"};
        return true;
    }
    if (n <= 16) {
        kernel_fn fn = nc <= 4 ? dense_kernel<16, 4>(method, fast)
                     : nc <= 8 ? dense_kernel<16, 8>(method, fast)
                     : nc <= 12 ? dense_kernel<16, 12>(method, fast)
                                : dense_kernel<16, 16>(method, fast);
        *out = {fn, 16, 1, "dense 16x1"};
        if (fast) {
            out->fn_shadow = nc <= 4 ? shadow_kernel<4>(method)
                           : nc <= 8 ? shadow_kernel<8>(method)
                           : nc <= 12 ? shadow_kernel<12>(method)
                                      : shadow_kernel<16>(method);
            pirk_kernel(n, out);
            out->fn_capped = nc <= 4 ? handover_kernel<4>(method)
                           : nc <= 8 ? handover_kernel<8>(method)
                           : nc <= 12 ? handover_kernel<12>(method)
                                      : handover_kernel<16>(method);
        }
        return true;
    }
    if (n <= 32) {
        kernel_fn fn = nc <= 24 ? dense_kernel<32, 24>(method, fast)
                                : dense_kernel<32, 32>(method, fast);
        *out = {fn, 32, 1, "dense 32x1"};
        if (fast) { if (nc <= 24) wide_dense<24>(method, out); else wide_dense<32>(method, out); }
        return true;
    }
    // 33..64 genes: one gene per lane, a CTA of 64 threads per trajectory.  With two genes per
    // lane in one warp the kernel is issue-bound on its own FP64 instructions and spills (50
    // genes: 2.95 us per attempt alone against 1.87; config 3, which waits for one time series of
    // 24,835 attempts, goes from 114 to 74 ms).  Beyond 64 genes the batches are throughput-bound
    // and two genes per lane win (200 genes, config 4: 1.09 s against 1.64 s with 256 x 1).
    // LASSIM_SPARSE_R2=1 brings the one-warp shape back (A/B).
    const char *r2 = getenv("LASSIM_SPARSE_R2");
    if (n <= 64) {
        if (r2 && atoi(r2) != 0) *out = {sparse_kernel<32, 2>(method, deg), 32, 2, "sparse 32x2"};
        else {
            *out = {sparse_kernel<64, 1>(method, deg), 64, 1, "sparse 64x1"};
            wide_sparse(method, deg, out);
        }
        return true;
    }
    const char *wide = getenv("LASSIM_SPARSE_WIDE");      // A/B: the coarser steps 128 / 512 only
    const bool fine = !(wide && atoi(wide) != 0);
    // CTA sizes in steps of 2 (1.5 at the top): idle warps of a CTA-wide group still pay every
    // barrier (70 genes: 128 x 2 -> 64 x 2 took a batch from 187 to 99 ms; 300 genes: 512 x 2 ->
    // 256 x 2 from 894 to 514 ms; 600 genes: 512 x 2 -> 384 x 2 from 750 to 618 ms; the
    // intermediate 96 and 192 gained 2-3 % and were left out)
    if (n <= 128 && fine) { *out = {sparse_kernel<64, 2>(method, deg), 64, 2, "sparse 64x2"}; return true; }
    if (n <= 256) { *out = {sparse_kernel<128, 2>(method, deg), 128, 2, "sparse 128x2"}; return true; }
    if (n <= 512 && fine) { *out = {sparse_kernel<256, 2>(method, deg), 256, 2, "sparse 256x2"}; return true; }
    if (n <= 768 && fine) { *out = {sparse_kernel<384, 2>(method, deg), 384, 2, "sparse 384x2"}; return true; }
    if (n <= 1024) { *out = {sparse_kernel<512, 2>(method, deg), 512, 2, "sparse 512x2"}; return true; }
    return false;
}

// Launches the kernel(s) of one evaluation on stream `s`.  `counters` points at two
// zeroed work-queue counters.  Allocates nothing.
//
// Speculative time-series kernel (one warp per trajectory, see grn_kernels.cuh) for
// batches small enough that the longest trajectory, not throughput, sets the time:
// the time series go to the speculative kernel on the caller's stream (they are the
// long ones), the other trajectories run concurrently on the auxiliary stream, and
// `s` waits for both.
// Which kernel integrates the time series of a batch of N candidates:
// 0 the standard one, 1 speculative, 3 stage-parallel (lassim_set_speculation).
// Automatic (2): hand-over (5) wherever the stage-parallel kernel exists (n <= 16) up to
// 16,384 candidates, the standard kernel alone beyond (there the batch outlasts its longest
// time series and the stage-parallel kernel's lanes are better spent on throughput).
// Measured on Th2 populations (tools/gpu_handover_sweep.py; ms per evaluation):
//        N   standard   hand-over (cap)
//      128     30.3       17.7  (500)    18.5 (768)
//     1024     34.9       21.4  (768)
//     4096     41.2       31.9  (3072)   33-34 (768..1500)
//    16384     78.0       68.3  (8000)
//    65536    177.7      203.0  (8000)
int series_mode(const lassim_handle *h, const KernelChoice &kc, int N, bool periph) {
    if (periph) return 0;
    const bool automatic = h->speculate == 2;
    // (beyond 16 genes the hand-over is on request only: a batch of 50-gene candidates is bound by
    // the total work of its knock-down trajectories, not by its longest time series - config 3:
    // 71 ms without, 97-107 ms with the hand-over, whose CTAs then share every SM with the
    // knock-down kernel; alone, the longest time series takes 35 ms instead of 46)
    if (kc.fn_pirk3 != nullptr && (h->speculate == 5 || (automatic && N <= 16384))) return 5;
    if (kc.fn_wide != nullptr && h->speculate == 5) return 5;
    if (kc.fn_wide != nullptr && (h->speculate == 3 || h->speculate == 4 || h->speculate == 6)) return 3;
    if (kc.fn_pirk3 != nullptr && h->speculate == 6) return 6;
    if (kc.fn_pirk2 != nullptr && h->speculate == 4) return 4;
    if (kc.fn_pirk != nullptr && h->speculate == 3) return 3;
    if (kc.fn_shadow != nullptr && h->speculate == 1) return 1;
    return 0;
}

// Warps per handed-over trajectory: three step sizes per attempt; two for the crowded
// stage-parallel kernel of batches above 8192 candidates (16,384: 60.6 against 62.4 ms).
int handover_warps_for(const lassim_handle *h, int N) {
    if (h->handover_warps == 2 || h->handover_warps == 3) return h->handover_warps;
    return N <= 8192 ? 3 : 2;
}

// Attempts a time series gets in the standard kernel before it is handed over: enough that
// the many short ones finish there (their hand-over would crowd the stage-parallel kernel),
// few enough that the long ones lose little time at the slower pace.  One number for every
// batch of at most kBatchIndependentMax candidates (a candidate evaluated alone and the same
// candidate in a population then take the same path and get the same bits); beyond, where
// throughput counts, three attempts per four candidates up to 8000.  The 5(4) pair takes
// about four attempts where the 8(5,3) pair takes one.
unsigned series_cap_for(const lassim_handle *h, int N) {
    if (h->series_cap != 0u) return h->series_cap;
    long long c = kHandoverCap;
    if (N > kBatchIndependentMax) { c = 3LL * N / 4; c = c > 8000 ? 8000 : c; }
    return (unsigned)(h->method == kMethodDP45 ? 4 * c : c);
}

int launch_kernels(lassim_handle *h, const KernelChoice &kc, const DevProblem &P,
                   const DevBatch &B, int N, int traj_per_candidate, bool periph,
                   cudaStream_t s, unsigned int *counters, const Scratch &sc, bool time_it) {
    // time-series kernel of the batch (series_mode above)
    const int smode = series_mode(h, kc, N, periph);
    const bool wide = kc.fn_wide != nullptr;
    const kernel_fn series_fn = wide ? (smode >= 3 ? kc.fn_wide : nullptr)
                              : smode == 6 ? kc.fn_pirk3 : smode >= 4 ? kc.fn_pirk2 : smode == 3 ? kc.fn_pirk
                              : smode == 1 ? kc.fn_shadow : nullptr;
    const bool speculate = series_fn != nullptr;
    auto launch = [&](kernel_fn fn, int threads, size_t smem, int lanes_per_item, unsigned begin,
                      unsigned count, unsigned int *counter, cudaStream_t st,
                      unsigned series_cap = 0u, int resume = 0, int cap_per_sm = 0,
                      unsigned *blocks_out = nullptr, bool spread = false) -> cudaError_t {
        if (count == 0) return cudaSuccess;
        const int items_per_cta = threads / lanes_per_item;
        cudaError_t e;
        int per_sm = 0;
        const auto key = std::make_tuple((const void *)fn, threads, smem);
        const auto hit = h->per_sm_cache.find(key);
        if (hit != h->per_sm_cache.end()) {
            per_sm = hit->second;
        } else {
            e = cudaFuncSetAttribute((const void *)fn,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, smem);
            if (e != cudaSuccess) return e;
            if (per_sm < 1) per_sm = 1;
            h->per_sm_cache[key] = per_sm;
        }
        long long blocks = ((long long)count + items_per_cta - 1) / items_per_cta;
        const long long resident = (long long)h->sm_count * per_sm;
        // stage-parallel kernels (several trajectories per CTA, one CTA per SM): one trajectory
        // per CTA for as long as there are SMs - two limit cycles packed into one CTA share its
        // schedulers to the end (2 candidates: 26.3 ms against 19.1 for the slower one alone)
        if (spread && blocks < resident) blocks = count < resident ? (long long)count : resident;
        if (blocks > resident) blocks = resident;
        if (cap_per_sm > 0 && blocks > (long long)h->sm_count * cap_per_sm)
            blocks = (long long)h->sm_count * cap_per_sm;
        if (blocks < 1) blocks = 1;
        DevBatch Bl = B;
        Bl.work_begin = begin; Bl.work_count = count; Bl.work_counter = counter;
        Bl.series_cap = series_cap; Bl.resume = resume; Bl.long_count = counters + 2;
        Bl.no_if = h->no_if;
        Bl.sp_attempts = reinterpret_cast<unsigned long long *>(counters + 4);
        h->launches += 1;
        if (blocks_out != nullptr) { *blocks_out = (unsigned)blocks; Bl.started = counters + 6; }
        fn<<<(unsigned)blocks, threads, smem, st>>>(P, Bl);
        return cudaGetLastError();
    };
    const int threads = threads_for(kc.L);
    const size_t smem = smem_bytes(kc);
    const unsigned total = (unsigned)N * (unsigned)traj_per_candidate;
    if (time_it) h->sp_timed = false;
    if (speculate) {
        const unsigned rest = total - (unsigned)N;
        // Hand-over: the knock-down kernel starts when the capped kernel is through AND every CTA
        // of the stage-parallel kernel behind it has begun to run (a one-thread gate kernel on
        // the second stream waits for their count).  Started together with the capped kernel it
        // doubles the time of the capped phase (its CTAs fill every SM beside one capped CTA:
        // 2.95 against 1.48 ms for 1024 Th2 candidates) and then sits where the stage-parallel
        // CTAs want to go; started behind the capped kernel by the event alone it gets there
        // first (592 CTAs against 148 that need three slots each) and the long time series wait
        // for it (+2 ms).  This way the time series have the machine while they are many, the
        // stage-parallel CTAs get their SMs, and the knock-down CTAs fill what is left: 20.3 ->
        // 19.4 ms (1024), 24.2 -> 21.3 (2048), 31.6 -> 27.3 (4096), 65.8 -> 64.3 (16,384).
        // Batches below 64 keep the early start (8 candidates: 1.3 against 1.9 ms - nothing is
        // handed over there and the two kernels would only run one after the other).
        const bool late_fork = smode == 5 && (h->knock_after_cap < 0 ? N >= 64 : h->knock_after_cap != 0);
        if (rest > 0 && !late_fork) {
            CU(cudaEventRecord(sc.ev_fork, s));
            CU(cudaStreamWaitEvent(sc.aux, sc.ev_fork, 0));
        }
        if (smode == 5) {
            // every time series starts in the standard kernel; the ones still going after
            // the cap of attempts continue in the multi-warp stage-parallel kernel behind it
            CU(launch(kc.fn_capped, threads, smem, kc.L, 0u, (unsigned)N, counters, s, series_cap_for(h, N), 0));
            if (time_it) CU(cudaEventRecord(h->ev_sp0, s));
            if (rest > 0 && late_fork) {
                CU(cudaEventRecord(sc.ev_fork, s));
                CU(cudaStreamWaitEvent(sc.aux, sc.ev_fork, 0));
            }
            if (!h->no_order) {
                // longest first (order_long_list_kernel): one small CTA between the two
                order_long_list_kernel<<<1, 1024, 0, s>>>(B.long_list, counters + 2, B.cost_main,
                                                          B.pert_cost, P.times, P.T);
                CU(cudaGetLastError());
                h->launches += 1;
            }
            unsigned sp_blocks = 0;
            unsigned *gate = (rest > 0 && late_fork) ? &sp_blocks : nullptr;
            if (wide)
                CU(launch(kc.fn_wide, kc.wide_threads, kc.wide_smem, kc.wide_threads, 0u, (unsigned)N,
                          counters + 3, s, 0u, 1, 0, gate));
            else if (handover_warps_for(h, N) == 3)
                CU(launch(kc.fn_pirk3, pirk_threads(3), kc.pirk3_smem, 96, 0u, (unsigned)N, counters + 3,
                          s, 0u, 1, 0, gate, true));
            else
                CU(launch(kc.fn_pirk2, pirk_threads(2), kc.pirk2_smem, 64, 0u, (unsigned)N, counters + 3,
                          s, 0u, 1, 0, gate, true));
            if (gate != nullptr) {
                start_gate_kernel<<<1, 1, 0, sc.aux>>>(counters + 6, sp_blocks);
                CU(cudaGetLastError());
                h->launches += 1;
            }
        } else if (smode == 1) {
            CU(launch(series_fn, threads, smem, 32, 0u, (unsigned)N, counters, s));
        } else {
            if (time_it) CU(cudaEventRecord(h->ev_sp0, s));
            if (wide)
                CU(launch(series_fn, kc.wide_threads, kc.wide_smem, kc.wide_threads, 0u, (unsigned)N, counters, s));
            else if (smode == 6)
                CU(launch(series_fn, pirk_threads(3), kc.pirk3_smem, 96, 0u, (unsigned)N, counters, s,
                          0u, 0, 0, nullptr, true));
            else if (smode == 4)
                CU(launch(series_fn, pirk_threads(2), kc.pirk2_smem, 64, 0u, (unsigned)N, counters, s,
                          0u, 0, 0, nullptr, true));
            else
                CU(launch(series_fn, pirk_threads(1), kc.pirk_smem, 32, 0u, (unsigned)N, counters, s,
                          0u, 0, 0, nullptr, true));
        }
        if (smode != 1 && time_it) {
            CU(cudaEventRecord(h->ev_sp1, s));
            h->sp_timed = true;
        }
        if (rest > 0) {
            CU(launch(kc.fn, threads, smem, kc.L, (unsigned)N, rest, counters + 1, sc.aux, 0u, 0,
                      smode == 5 ? h->knock_per_sm : 0));
            CU(cudaEventRecord(sc.ev_join, sc.aux));
            CU(cudaStreamWaitEvent(s, sc.ev_join, 0));
        }
    } else {
        CU(launch(kc.fn, threads, smem, kc.L, 0u, total, counters, s));
    }
    return LASSIM_OK;
}

// Peripheral problems integrate n + 1 states (fixed core + one gene) with the
// dense warp-slice kernels; NC covers the n core columns.
template <int L, int NC>
kernel_fn peripheral_kernel(int method) {
    if (method == kMethodDP45)
        return grn_cost_kernel<L, 1, DenseW<1, NC>, kMethodDP45, true, true>;
    return grn_cost_kernel<L, 1, DenseW<1, NC>, kMethodDOP853, true, true>;
}

bool choose_peripheral_kernel(const lassim_handle *h, KernelChoice *out) {
    const int nc = h->n, np = nc + 1, method = h->method;
    const int cols = (nc + 3) & ~3;
    if (np <= 16) {
        kernel_fn fn = cols <= 4 ? peripheral_kernel<16, 4>(method)
                     : cols <= 8 ? peripheral_kernel<16, 8>(method)
                     : cols <= 12 ? peripheral_kernel<16, 12>(method)
                                  : peripheral_kernel<16, 16>(method);
        *out = {fn, 16, 1, "peripheral dense 16x1"};
        return true;
    }
    if (np <= 32) {
        kernel_fn fn = cols <= 24 ? peripheral_kernel<32, 24>(method)
                                  : peripheral_kernel<32, 32>(method);
        *out = {fn, 32, 1, "peripheral dense 32x1"};
        return true;
    }
    return false;
}

// `sc` = the scratch this evaluation writes to; `main` evaluations (the entry points) are
// the ones the handle's timing events and lassim_last_* diagnostics refer to.
int launch_eval(lassim_handle *h, Scratch &sc, int N, int mode, const double *dX, double *dY,
                cudaStream_t s, const int *d_pgene = nullptr) {
    const bool is_main = &sc == &h->main;
    const bool periph = d_pgene != nullptr;
    KernelChoice kc;
    if (periph) {
        if (!choose_peripheral_kernel(h, &kc))
            return fail(h, LASSIM_E_UNSUPPORTED,
                        "peripheral genes need a core of at most 31 genes (core has %d)", h->n);
    } else if (!choose_kernel(h, &kc)) {
        return fail(h, LASSIM_E_UNSUPPORTED,
                    "state size n=%d is outside the kernels built (n <= 1024)", h->n);
    }
    const int K = periph ? 1 : h->K, n = periph ? h->n + 1 : h->n;
    const ZeroLayout z = zero_layout(N);
    CU(sc.zero.reserve(z.bytes));
    CU(sc.cost.reserve(N));
    CU(sc.cost_main.reserve(N));
    CU(sc.pert_cost.reserve(N));
    CU(sc.steps.reserve((size_t)N * K));
    CU(sc.long_list.reserve((size_t)N));
    if (!dY) {
        CU(sc.Y.reserve((size_t)N * h->T * n));
        dY = sc.Y.p;
    }
    if (K > 1 && mode == 0) CU(sc.E.reserve((size_t)N * K * n));
    CU(cudaMemsetAsync(sc.zero.p, 0, z.bytes, s));
    CU(cudaMemsetAsync(sc.steps.p, 0, sizeof(unsigned) * (size_t)N * K, s));

    DevProblem P;
    P.n = n; P.T = h->T; P.K = K; P.U = h->U; P.dim = h->dim; P.nnz = h->nnz;
    P.times = h->d_times.p; P.y0 = h->d_y0.p; P.data = h->d_data.p;
    P.sigma2 = h->d_sigma2.p; P.kidx = h->d_kidx.p;
    P.row_ptr = h->d_row_ptr.p; P.col_idx = h->d_col_idx.p;
    P.pert = (K > 1) ? h->d_pert.p : nullptr;
    P.traj_prefix = h->d_traj_prefix.p; P.traj_tt = h->d_traj_tt.p;
    P.knock_ctrl = h->d_knock_ctrl.p;
    P.pert_factor = h->pert_factor; P.rtol = h->rtol; P.atol = h->atol;
    P.max_steps = (unsigned int)(h->max_steps > 4000000000LL ? 4000000000LL : h->max_steps);
    P.swing = h->jump_guard;
    P.core_x = h->d_core_x.p; P.pmask = h->d_pmask.p; P.pdata = h->d_pdata.p;
    P.psigma2 = h->d_psigma2.p; P.py0 = h->d_py0.p;
    if (periph) { P.U = 0; P.dim = h->n + 2; P.pert = nullptr; }

    DevBatch B;
    B.N = N; B.mode = mode; B.X = dX; B.pgene = d_pgene;
    B.work_begin = 0; B.work_count = 0;     // set per launch below
    B.cost = sc.cost.p; B.cost_main = sc.cost_main.p; B.pert_cost = sc.pert_cost.p;
    B.work_counter = reinterpret_cast<unsigned int *>(sc.zero.p);
    B.done = reinterpret_cast<int *>(sc.zero.p + z.off_done);
    B.status = reinterpret_cast<int *>(sc.zero.p + z.off_status);
    B.counters = reinterpret_cast<unsigned long long *>(sc.zero.p + z.off_counters);
    B.Y = dY; B.E = sc.E.p; B.traj_steps = sc.steps.p; B.order = nullptr;
    B.long_list = sc.long_list.p;

    if (is_main) CU(cudaEventRecord(h->ev0, s));
    int rc = launch_kernels(h, kc, P, B, N, mode == 0 ? K : 1, periph, s,
                            reinterpret_cast<unsigned int *>(sc.zero.p), sc, is_main);
    if (rc) return rc;
    if (is_main) {
        CU(cudaEventRecord(h->ev1, s));
        h->last_periph_march = false;
        h->timed = true;
        h->last_N = N;
        h->last_had_E = (K > 1 && mode == 0);
    }
    return LASSIM_OK;
}

// ---- peripheral genes: the core integrated once, one thread per candidate ------------------
typedef void (*core_fn)(const DevProblem, const double *, const CoreTable, const double, const double);
typedef void (*march_fn)(const PeriphBatch);

template <int NC>
void periph_kernels_nc(core_fn *core, march_fn *march) {
    *core = core_dense_kernel<NC>;
    *march = peripheral_march_kernel<NC>;
}

bool periph_kernels(int n, core_fn *core, march_fn *march, int *nc_out) {
    const int nc = (n + 3) & ~3;
    if (n > 32) return false;
    if (nc <= 4) periph_kernels_nc<4>(core, march);
    else if (nc <= 8) periph_kernels_nc<8>(core, march);
    else if (nc <= 12) periph_kernels_nc<12>(core, march);
    else if (nc <= 16) periph_kernels_nc<16>(core, march);
    else if (nc <= 24) periph_kernels_nc<24>(core, march);
    else periph_kernels_nc<32>(core, march);
    *nc_out = nc <= 4 ? 4 : nc <= 8 ? 8 : nc <= 12 ? 12 : nc <= 16 ? 16 : nc <= 24 ? 24 : 32;
    return true;
}

// The core is one trajectory: it is integrated a hundred times tighter than the candidates
// (its error is then no part of theirs) and costs a few milliseconds per sweep.
constexpr double kCoreTolFactor = 0.01;

// Integrates the core of the peripheral set-up into h->d_core_table (blocks).  Re-run when
// the tolerances change.
int build_core_table(lassim_handle *h) {
    core_fn core; march_fn march; int nc = 0;
    if (!periph_kernels(h->n, &core, &march, &nc))
        return fail(h, LASSIM_E_UNSUPPORTED,
                    "peripheral genes need a core of at most 32 genes (core has %d)", h->n);
    const int n = h->n, T = h->T;
    DevProblem P{};
    P.n = n; P.T = T; P.K = 1; P.U = 0; P.dim = h->dim; P.nnz = h->nnz;
    P.times = h->d_times.p; P.y0 = h->d_y0.p; P.kidx = h->d_kidx.p;
    // the step budget of lassim_set_tolerances is per CANDIDATE; the one core solve gets the
    // default budget at least (it is bounded by the table's capacity anyway)
    const long long core_budget = h->max_steps > 20000000LL ? h->max_steps : 20000000LL;
    P.max_steps = (unsigned int)(core_budget > 4000000000LL ? 4000000000LL : core_budget);
    P.swing = h->jump_guard;
    const size_t smem = ((size_t)kSmemHeader + nc + 2) * sizeof(double);
    CU(h->d_core_count.reserve(1));
    CU(h->d_core_samples.reserve((size_t)T * n));
    int capacity = 1 << 14;
    for (int attempt = 0; attempt < 4; ++attempt, capacity *= 8) {
        CU(h->d_core_table.reserve((size_t)capacity * core_interval_doubles(n)));
        CoreTable out{h->d_core_table.p, h->d_core_count.p, h->d_core_samples.p, capacity};
        core<<<1, 32, smem, h->stream>>>(P, h->d_core_x.p, out, h->rtol * kCoreTolFactor,
                                         h->atol * kCoreTolFactor);
        CU(cudaGetLastError());
        h->launches += 1;
        int count = 0;
        CU(cudaMemcpyAsync(&count, h->d_core_count.p, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
        CU(cudaStreamSynchronize(h->stream));
        if (count >= 0) {
            h->core_intervals = count;
            h->core_rtol = h->rtol; h->core_atol = h->atol; h->core_swing = h->jump_guard;
            return LASSIM_OK;
        }
        const int status = -1 - count;
        if (status != 4)
            return fail(h, LASSIM_E_INVALID,
                        "the fixed core could not be integrated (%s): peripheral genes need a core "
                        "solution with a finite trajectory",
                        status == 1 ? "step budget used up" : "non-finite state or step-size underflow");
    }
    return fail(h, LASSIM_E_UNSUPPORTED, "the core trajectory needs more than %d steps", capacity / 8);
}

// Enqueues the march over the core's intervals for N candidates already on the device.
int launch_periph_march(lassim_handle *h, int N, const double *dX, const int *d_gene, cudaStream_t s) {
    if (h->core_intervals < 0 || h->core_rtol != h->rtol || h->core_atol != h->atol ||
        h->core_swing != h->jump_guard) {
        int rc = build_core_table(h);
        if (rc) return rc;
    }
    core_fn core; march_fn march; int nc = 0;
    periph_kernels(h->n, &core, &march, &nc);
    CU(h->main.cost.reserve(N));
    CU(h->d_pYs.reserve((size_t)h->T * N));
    CU(h->d_pstatus.reserve(N));
    CU(h->d_psteps.reserve(N));
    CU(h->d_pcounters.reserve((size_t)3 * N));
    PeriphBatch B{};
    B.N = N; B.n = h->n; B.T = h->T; B.X = dX; B.pgene = d_gene;
    B.pmask = h->d_pmask.p; B.pdata = h->d_pdata.p; B.psigma2 = h->d_psigma2.p; B.py0 = h->d_py0.p;
    B.table = h->d_core_table.p; B.intervals = h->core_intervals;
    B.Ys = h->d_pYs.p; B.cost = h->main.cost.p; B.status = h->d_pstatus.p; B.steps = h->d_psteps.p;
    B.counters = h->d_pcounters.p;
    B.rtol = h->rtol; B.atol = h->atol;
    B.max_steps = (unsigned int)(h->max_steps > 4000000000LL ? 4000000000LL : h->max_steps);
    B.swing = h->jump_guard;
    CU(cudaEventRecord(h->ev0, s));
    march<<<(N + 127) / 128, 128, 0, s>>>(B);
    CU(cudaGetLastError());
    CU(cudaEventRecord(h->ev1, s));
    h->launches += 1;
    h->timed = true;
    h->sp_timed = false;
    h->last_N = N;
    h->last_had_E = false;
    h->last_periph_march = true;
    return LASSIM_OK;
}

void release_de_lanes(lassim_handle *h) {
    for (DeLane &lane : h->de_lanes) {
        lane.scratch.release();
        if (lane.ev_done) cudaEventDestroy(lane.ev_done);
        if (lane.stream) cudaStreamDestroy(lane.stream);
    }
    h->de_lanes.clear();
}

}  // namespace

extern "C" {

const char *lassim_version(void) { return "0.2.0"; }

const char *lassim_last_error(const lassim_handle *h) {
    return h ? h->err.c_str() : g_create_error.c_str();
}

int lassim_create(int device, lassim_handle **out) {
    lassim_handle *h = nullptr;
    if (!out) return fail(nullptr, LASSIM_E_INVALID, "out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, LASSIM_E_CUDA, "no CUDA device: %s",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    if (device < 0 || device >= count)
        return fail(nullptr, LASSIM_E_INVALID, "device %d out of range [0,%d)", device, count);
    DeviceGuard guard_(device);
    if (guard_.err != cudaSuccess)
        return fail(nullptr, LASSIM_E_CUDA, "cudaSetDevice: %s", cudaGetErrorString(guard_.err));
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess)
        return fail(nullptr, LASSIM_E_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    if (prop.major != 10)
        return fail(nullptr, LASSIM_E_UNSUPPORTED,
                    "device %d is sm_%d%d; this library is built for sm_100a only",
                    device, prop.major, prop.minor);
    h = new lassim_handle();
    h->device = device;
    h->sm_count = prop.multiProcessorCount;
    // the auxiliary stream carries the short trajectories of a speculative evaluation:
    // lowest priority, so that the time series' blocks are placed first
    cudaDeviceGetStreamPriorityRange(&h->prio_lo, &h->prio_hi);
    if (cudaStreamCreateWithPriority(&h->stream, cudaStreamNonBlocking, h->prio_hi) != cudaSuccess ||
        h->main.init(h->prio_lo) != cudaSuccess ||
        cudaEventCreate(&h->ev0) != cudaSuccess || cudaEventCreate(&h->ev1) != cudaSuccess ||
        cudaEventCreate(&h->ev_sp0) != cudaSuccess || cudaEventCreate(&h->ev_sp1) != cudaSuccess ||
        cudaEventCreateWithFlags(&h->ev_de_fork, cudaEventDisableTiming) != cudaSuccess) {
        delete h;
        return fail(nullptr, LASSIM_E_CUDA, "stream/event creation failed");
    }
    if (const char *sp = getenv("LASSIM_SPECULATE")) h->speculate = atoi(sp);
    if (const char *sp = getenv("LASSIM_SERIES_CAP")) h->series_cap = (unsigned)atoi(sp);
    if (const char *sp = getenv("LASSIM_HANDOVER_WARPS")) h->handover_warps = atoi(sp);
    if (const char *sp = getenv("LASSIM_DE_LANES")) h->de_lanes_wanted = atoi(sp);
    if (const char *sp = getenv("LASSIM_NO_IF")) h->no_if = atoi(sp);
    if (const char *sp = getenv("LASSIM_NO_ORDER")) h->no_order = atoi(sp) != 0;
    if (const char *sp = getenv("LASSIM_KNOCK_AFTER_CAP")) h->knock_after_cap = atoi(sp);
    if (const char *sp = getenv("LASSIM_KNOCK_PER_SM")) h->knock_per_sm = atoi(sp);
    if (const char *sp = getenv("LASSIM_PERIPHERAL_JOINT")) h->periph_joint = atoi(sp) != 0;
    if (const char *sp = getenv("LASSIM_JUMP_GUARD")) h->jump_guard = (float)atof(sp);
    double tab[kExpTable];
    for (int j = 0; j < kExpTable; ++j) tab[j] = exp2((double)j / (double)kExpTable);
    if (cudaMemcpyToSymbol(g_exp2_table, tab, sizeof(tab)) != cudaSuccess) {
        lassim_destroy(h);
        return fail(nullptr, LASSIM_E_CUDA, "upload of the exp2 table failed");
    }
    *out = h;
    return LASSIM_OK;
}

int lassim_destroy(lassim_handle *h) {
    if (!h) return LASSIM_OK;
    DeviceGuard guard_(h->device);
    cudaDeviceSynchronize();
    h->d_times.release(); h->d_y0.release(); h->d_data.release(); h->d_sigma2.release();
    h->d_pert.release(); h->d_traj_tt.release(); h->d_kidx.release();
    h->d_row_ptr.release(); h->d_col_idx.release();
    h->d_core_x.release(); h->d_pdata.release(); h->d_psigma2.release(); h->d_py0.release();
    h->d_pmask.release(); h->d_pgene.release();
    h->d_core_table.release(); h->d_core_samples.release(); h->d_pYs.release();
    h->d_core_count.release(); h->d_pstatus.release(); h->d_psteps.release(); h->d_pcounters.release();
    h->de_X.release(); h->de_F.release(); h->de_trial.release(); h->de_best.release();
    h->de_table.release(); h->de_lb.release(); h->de_ub.release();
    h->d_traj_prefix.release(); h->d_knock_ctrl.release();
    h->d_X.release(); h->d_probe.release();
    h->main.release();
    release_de_lanes(h);
    cudaEventDestroy(h->ev0); cudaEventDestroy(h->ev1);
    cudaEventDestroy(h->ev_sp0); cudaEventDestroy(h->ev_sp1);
    if (h->ev_de_fork) cudaEventDestroy(h->ev_de_fork);
    cudaStreamDestroy(h->stream);
    delete h;
    return LASSIM_OK;
}

int lassim_set_problem(lassim_handle *h, int n, int T, const double *times,
                       const double *y0, const double *data, const double *sigma,
                       const unsigned char *mask, const double *pert,
                       double pert_factor) {
    if (!h) return LASSIM_E_INVALID;
    if (n < 1 || T < 1 || !times || !y0 || !data || !sigma || !mask)
        return fail(h, LASSIM_E_INVALID, "set_problem: n=%d T=%d or a NULL array", n, T);
    if (n > 1024)
        return fail(h, LASSIM_E_UNSUPPORTED,
                    "state size n=%d is outside the kernels built (n <= 1024)", n);
    for (int i = 1; i < T; ++i)
        if (!(times[i] >= times[i - 1]))
            return fail(h, LASSIM_E_UNSUPPORTED,
                        "sample times must be non-decreasing (times[%d]=%g < times[%d]=%g)",
                        i, times[i], i - 1, times[i - 1]);
    if (pert)
        for (int i = 0; i < n; ++i)
            if (!(pert[(size_t)i * (n + 2) + n + 1] >= pert[(size_t)i * (n + 2) + n]))
                return fail(h, LASSIM_E_UNSUPPORTED,
                            "perturbation row %d has t_end < t_start", i);
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    h->has_problem = false;
    h->p_genes = 0;
    h->core_intervals = -1;
    h->de_ready = false;
    h->n = n; h->T = T; h->pert_factor = pert_factor;
    std::vector<double> sigma2(n);
    for (int i = 0; i < n; ++i) sigma2[i] = sigma[i] * sigma[i];
    int rc;
    if ((rc = upload(h, h->d_times, times, (size_t)T))) return rc;
    if ((rc = upload(h, h->d_y0, y0, (size_t)n))) return rc;
    if ((rc = upload(h, h->d_data, data, (size_t)T * n))) return rc;
    if ((rc = upload(h, h->d_sigma2, sigma2.data(), (size_t)n))) return rc;

    // trajectory table: 0 = time series; 1..U = controls in order of first
    // appearance of their (t_start, t_end) pair, solved with the lambdas as
    // scaled by the rows before that appearance; 1+U+i = row i knocked
    // (perturbation_functions.pyx:64-85).
    std::vector<int> prefix(1, 0), knock_ctrl;
    std::vector<double> tt(2, 0.0);
    h->U = 0; h->K = 1;
    h->pert_h.clear();
    if (pert) {
        const int np2 = n + 2;
        h->pert_h.assign(pert, pert + (size_t)n * np2);
        std::vector<std::pair<double, double>> pairs;
        std::vector<int> first_row;
        knock_ctrl.resize(n);
        for (int i = 0; i < n; ++i) {
            const std::pair<double, double> p(pert[(size_t)i * np2 + n],
                                              pert[(size_t)i * np2 + n + 1]);
            int u = -1;
            for (size_t q = 0; q < pairs.size(); ++q)
                if (pairs[q] == p) { u = (int)q; break; }
            if (u < 0) { u = (int)pairs.size(); pairs.push_back(p); first_row.push_back(i); }
            knock_ctrl[i] = 1 + u;
        }
        h->U = (int)pairs.size();
        h->K = 1 + h->U + n;
        prefix.resize(h->K); tt.resize(2 * (size_t)h->K);
        for (int u = 0; u < h->U; ++u) {
            prefix[1 + u] = first_row[u];
            tt[2 * (1 + u)] = pairs[u].first; tt[2 * (1 + u) + 1] = pairs[u].second;
        }
        for (int i = 0; i < n; ++i) {
            const int k = 1 + h->U + i;
            prefix[k] = i + 1;
            tt[2 * k] = pert[(size_t)i * np2 + n]; tt[2 * k + 1] = pert[(size_t)i * np2 + n + 1];
        }
        if ((rc = upload(h, h->d_pert, pert, (size_t)n * np2))) return rc;
        if ((rc = upload(h, h->d_knock_ctrl, knock_ctrl.data(), (size_t)n))) return rc;
    }
    if ((rc = upload(h, h->d_traj_prefix, prefix.data(), prefix.size()))) return rc;
    if ((rc = upload(h, h->d_traj_tt, tt.data(), tt.size()))) return rc;
    if ((rc = upload_mask(h, mask))) return rc;   // synchronises: locals are done
    h->has_problem = true;
    return LASSIM_OK;
}

int lassim_set_mask(lassim_handle *h, const unsigned char *mask) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->has_problem) return fail(h, LASSIM_E_STATE, "set_mask before set_problem");
    if (!mask) return fail(h, LASSIM_E_INVALID, "mask is NULL");
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    // the peripheral set-up and the device populations were laid out for the old mask
    // (core_x is indexed through kidx, the populations have the old dimension)
    h->p_genes = 0;
    h->core_intervals = -1;
    h->de_ready = false;
    return upload_mask(h, mask);
}

int lassim_set_tolerances(lassim_handle *h, double rtol, double atol, long long max_steps) {
    if (!h) return LASSIM_E_INVALID;
    if (!(rtol > 0.0) || !(atol >= 0.0) || max_steps < 1)
        return fail(h, LASSIM_E_INVALID, "tolerances must be positive (rtol=%g atol=%g max_steps=%lld)",
                    rtol, atol, max_steps);
    h->rtol = rtol; h->atol = atol; h->max_steps = max_steps;
    return LASSIM_OK;
}

int lassim_set_series_cap(lassim_handle *h, unsigned int attempts) {
    if (!h) return LASSIM_E_INVALID;
    h->series_cap = attempts;
    return LASSIM_OK;
}

int lassim_set_jump_guard(lassim_handle *h, double swing) {
    if (!h) return LASSIM_E_INVALID;
    if (!(swing > 0.0))
        return fail(h, LASSIM_E_INVALID, "set_jump_guard: swing=%g (in (0, 1]; 2 or more = off)", swing);
    h->jump_guard = (float)swing;
    return LASSIM_OK;
}

int lassim_set_speculation(lassim_handle *h, int mode) {
    if (!h) return LASSIM_E_INVALID;
    if (mode < 0 || mode > 6)
        return fail(h, LASSIM_E_INVALID,
                    "time-series kernel mode %d (0 standard, 1 speculative, 2 auto, 3 stage-parallel, "
                    "4 stage-parallel with two step sizes, 5 hand-over, 6 stage-parallel with three "
                    "step sizes)", mode);
    h->speculate = mode;
    return LASSIM_OK;
}

int lassim_set_method(lassim_handle *h, int method) {
    if (!h) return LASSIM_E_INVALID;
    if (method != kMethodDP45 && method != kMethodDOP853)
        return fail(h, LASSIM_E_INVALID, "unknown integrator %d (0 = DP5(4), 1 = DP8(5,3))", method);
    h->method = method;
    return LASSIM_OK;
}

long long lassim_flops_per_attempt(const lassim_handle *h) {
    if (!h || !h->has_problem) return -1;
    // F_rhs = 2*nnz + 5*n per RHS evaluation.  5(4): 6 RHS + 42 + 3
    // FMAs/ops per gene = 12*nnz + 90*n.  8(5,3): 12 RHS + 74 tableau FMAs (50 in A,
    // 8 each in b, e5, e3) + 4 = 24*nnz + 212*n.
    const long long nnz = h->nnz, n = h->n;
    return h->method == kMethodDOP853 ? 24 * nnz + 212 * n : 12 * nnz + 90 * n;
}

long long lassim_series_flops_per_attempt(const lassim_handle *h, int N) {
    if (!h || !h->has_problem) return -1;
    KernelChoice kc;
    if (!choose_kernel(h, &kc)) return -1;
    const long long nnz = h->nnz, n = h->n;
    const int smode = series_mode(h, kc, N, false);
    switch (smode) {
        case 3:
        case 4:
        case 5:
        case 6: {
            // S stages, M = S corrections: (M+1)*S RHS evaluations of 2*nnz + 5*n flops,
            // M*S stage sums and the predictor (2*S flops per gene each), two weighted
            // sums for solution and estimate (2*S each), 4 for the error measure
            const long long S = (h->n <= 12 || h->n > 16) ? 5 : 4, M = S;
            const long long one = (M + 1) * S * (2 * nnz + 5 * n) + (M * S + S + 2) * 2 * S * n + 4 * n;
            // every warp of a trajectory's group works (one step size beyond 16 genes)
            const long long warps = kc.fn_wide != nullptr ? 1
                                  : smode == 6 ? 3 : smode == 5 ? handover_warps_for(h, N) : smode == 4 ? 2 : 1;
            return warps * one;
        }
        case 1: return 2 * lassim_flops_per_attempt(h);     // both halves of the warp
        default: return lassim_flops_per_attempt(h);
    }
}

int lassim_last_stage_parallel_attempts(lassim_handle *h, unsigned long long *attempts) {
    if (!h || !attempts) return LASSIM_E_INVALID;
    if (!h->timed || !h->main.zero.p) return fail(h, LASSIM_E_STATE, "no evaluation has run yet");
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(attempts, h->main.zero.p + 16, sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return LASSIM_OK;
}

const char *lassim_series_kernel_name(lassim_handle *h, int N) {
    if (!h || !h->has_problem) return "";
    KernelChoice kc;
    if (!choose_kernel(h, &kc)) return "";
    const int smode = series_mode(h, kc, N, false);
    switch (smode) {
        case 3:
        case 4:
        case 5:
        case 6: {
            if (kc.fn_wide != nullptr) {
                h->series_kernel_name = std::string(h->n <= 32 ? "grn_series_pirk_wide_kernel<5 Gauss stages x 32 lanes>"
                                                               : "grn_series_pirk_wide_kernel<5 Gauss stages x 64 lanes>") +
                                        (smode == 5 ? " behind " + std::string(lassim_kernel_name(h)) : "");
                break;
            }
            const int warps = smode == 6 ? 3 : smode == 5 ? handover_warps_for(h, N) : smode == 4 ? 2 : 1;
            h->series_kernel_name = std::string(h->n <= 12 ? "grn_series_pirk_kernel<5 Gauss stages x 6 lanes"
                                                            : "grn_series_pirk_kernel<4 Gauss stages x 8 lanes") +
                                    (warps == 3 ? ", three step sizes>" : warps == 2 ? ", two step sizes>" : ">") +
                                    (smode == 5 ? " behind " + std::string(lassim_kernel_name(h)) : "");
            break;
        }
        case 1:
            h->series_kernel_name = std::string("grn_cost_kernel<") + kc.name + ", speculative, " +
                                    (h->method == kMethodDOP853 ? "dop853>" : "dp45>");
            break;
        default:
            h->series_kernel_name = lassim_kernel_name(h);
    }
    return h->series_kernel_name.c_str();
}

const char *lassim_kernel_name(lassim_handle *h) {
    if (!h || !h->has_problem) return "";
    KernelChoice kc;
    if (!choose_kernel(h, &kc)) return "";
    h->kernel_name = std::string("grn_cost_kernel<") + kc.name +
                     (h->method == kMethodDOP853 ? ", dop853>" : ", dp45>");
    return h->kernel_name.c_str();
}

int lassim_dim(const lassim_handle *h) { return (h && h->has_problem) ? h->dim : -1; }

int lassim_trajectories_per_candidate(const lassim_handle *h) {
    return (h && h->has_problem) ? h->K : -1;
}

int lassim_eval_device(lassim_handle *h, int N, const double *dX, double *dcost,
                       double *dcost_main, double *dpert_cost, int *dstatus,
                       unsigned long long *dcounters, void *stream) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->has_problem) return fail(h, LASSIM_E_STATE, "eval before set_problem");
    if (N < 0 || (N > 0 && (!dX || !dcost)))
        return fail(h, LASSIM_E_INVALID, "eval: N=%d, X or cost is NULL", N);
    if (N == 0) return LASSIM_OK;
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
    int rc = launch_eval(h, h->main, N, 0, dX, nullptr, s);
    if (rc) return rc;
    const ZeroLayout z = zero_layout(N);
    const size_t nb = sizeof(double) * (size_t)N;
    CU(cudaMemcpyAsync(dcost, h->main.cost.p, nb, cudaMemcpyDeviceToDevice, s));
    if (dcost_main) CU(cudaMemcpyAsync(dcost_main, h->main.cost_main.p, nb, cudaMemcpyDeviceToDevice, s));
    if (dpert_cost) CU(cudaMemcpyAsync(dpert_cost, h->main.pert_cost.p, nb, cudaMemcpyDeviceToDevice, s));
    if (dstatus) CU(cudaMemcpyAsync(dstatus, h->main.zero.p + z.off_status, sizeof(int) * (size_t)N,
                                    cudaMemcpyDeviceToDevice, s));
    if (dcounters) CU(cudaMemcpyAsync(dcounters, h->main.zero.p + z.off_counters,
                                      sizeof(unsigned long long) * 3 * (size_t)N,
                                      cudaMemcpyDeviceToDevice, s));
    return LASSIM_OK;
}

int lassim_eval(lassim_handle *h, int N, const double *X, double *cost,
                double *cost_main, double *pert_cost, int *status,
                unsigned long long *counters) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->has_problem) return fail(h, LASSIM_E_STATE, "eval before set_problem");
    if (N < 0 || (N > 0 && (!X || !cost)))
        return fail(h, LASSIM_E_INVALID, "eval: N=%d, X or cost is NULL", N);
    if (N == 0) return LASSIM_OK;
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    cudaStream_t s = h->stream;
    CU(h->d_X.reserve((size_t)N * h->dim));
    CU(cudaMemcpyAsync(h->d_X.p, X, sizeof(double) * (size_t)N * h->dim,
                       cudaMemcpyHostToDevice, s));
    int rc = launch_eval(h, h->main, N, 0, h->d_X.p, nullptr, s);
    if (rc) return rc;
    const ZeroLayout z = zero_layout(N);
    const size_t nb = sizeof(double) * (size_t)N;
    CU(cudaMemcpyAsync(cost, h->main.cost.p, nb, cudaMemcpyDeviceToHost, s));
    if (cost_main) CU(cudaMemcpyAsync(cost_main, h->main.cost_main.p, nb, cudaMemcpyDeviceToHost, s));
    if (pert_cost) CU(cudaMemcpyAsync(pert_cost, h->main.pert_cost.p, nb, cudaMemcpyDeviceToHost, s));
    if (status) CU(cudaMemcpyAsync(status, h->main.zero.p + z.off_status, sizeof(int) * (size_t)N,
                                   cudaMemcpyDeviceToHost, s));
    if (counters) CU(cudaMemcpyAsync(counters, h->main.zero.p + z.off_counters,
                                     sizeof(unsigned long long) * 3 * (size_t)N,
                                     cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    return LASSIM_OK;
}

int lassim_trajectories(lassim_handle *h, int N, const double *X, double *Y, int *status) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->has_problem) return fail(h, LASSIM_E_STATE, "trajectories before set_problem");
    if (N < 0 || (N > 0 && (!X || !Y)))
        return fail(h, LASSIM_E_INVALID, "trajectories: N=%d, X or Y is NULL", N);
    if (N == 0) return LASSIM_OK;
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    cudaStream_t s = h->stream;
    CU(h->d_X.reserve((size_t)N * h->dim));
    CU(cudaMemcpyAsync(h->d_X.p, X, sizeof(double) * (size_t)N * h->dim,
                       cudaMemcpyHostToDevice, s));
    int rc = launch_eval(h, h->main, N, 1, h->d_X.p, nullptr, s);
    if (rc) return rc;
    const ZeroLayout z = zero_layout(N);
    CU(cudaMemcpyAsync(Y, h->main.Y.p, sizeof(double) * (size_t)N * h->T * h->n,
                       cudaMemcpyDeviceToHost, s));
    if (status) CU(cudaMemcpyAsync(status, h->main.zero.p + z.off_status, sizeof(int) * (size_t)N,
                                   cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    return LASSIM_OK;
}

int lassim_set_peripheral(lassim_handle *h, const double *core_x, int genes,
                          const double *data, const double *sigma, const double *y0,
                          const unsigned char *mask) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->has_problem) return fail(h, LASSIM_E_STATE, "set_peripheral before set_problem");
    if (!core_x || genes < 1 || !data || !sigma || !y0)
        return fail(h, LASSIM_E_INVALID, "set_peripheral: genes=%d or a NULL array", genes);
    if (h->n > 32 || (h->periph_joint && h->n > 31))
        return fail(h, LASSIM_E_UNSUPPORTED,
                    "peripheral genes need a core of at most %d genes (core has %d)",
                    h->periph_joint ? 31 : 32, h->n);
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    const int n = h->n, T = h->T;
    h->core_intervals = -1;
    std::vector<double> sigma2(genes);
    for (int g = 0; g < genes; ++g) sigma2[g] = sigma[g] * sigma[g];
    std::vector<unsigned char> all;
    if (!mask) { all.assign((size_t)genes * n, 1); mask = all.data(); }
    int rc;
    if ((rc = upload(h, h->d_core_x, core_x, (size_t)h->dim))) return rc;
    if ((rc = upload(h, h->d_pdata, data, (size_t)genes * T))) return rc;
    if ((rc = upload(h, h->d_psigma2, sigma2.data(), (size_t)genes))) return rc;
    if ((rc = upload(h, h->d_py0, y0, (size_t)genes))) return rc;
    if ((rc = upload(h, h->d_pmask, mask, (size_t)genes * n))) return rc;
    CU(cudaStreamSynchronize(h->stream));
    h->p_genes = genes;
    // the core is integrated here, once for the sweep (a core that cannot be integrated is
    // reported now, not by every evaluation)
    if (!h->periph_joint && (rc = build_core_table(h))) { h->p_genes = 0; return rc; }
    return LASSIM_OK;
}

int lassim_set_peripheral_joint(lassim_handle *h, int joint) {
    if (!h) return LASSIM_E_INVALID;
    if (joint && h->has_problem && h->n > 31)
        return fail(h, LASSIM_E_UNSUPPORTED,
                    "the joint peripheral path needs a core of at most 31 genes (core has %d)", h->n);
    h->periph_joint = joint != 0;
    return LASSIM_OK;
}

int lassim_eval_peripheral(lassim_handle *h, int N, const int *gene, const double *Xp,
                           double *cost, int *status) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->has_problem || h->p_genes < 1)
        return fail(h, LASSIM_E_STATE, "eval_peripheral before set_problem + set_peripheral");
    if (N < 0 || (N > 0 && (!gene || !Xp || !cost)))
        return fail(h, LASSIM_E_INVALID, "eval_peripheral: N=%d or a NULL array", N);
    if (N == 0) return LASSIM_OK;
    for (int i = 0; i < N; ++i)
        if (gene[i] < 0 || gene[i] >= h->p_genes)
            return fail(h, LASSIM_E_INVALID, "eval_peripheral: gene[%d]=%d outside [0,%d)",
                        i, gene[i], h->p_genes);
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    cudaStream_t s = h->stream;
    const int pdim = h->n + 2;
    CU(h->d_X.reserve((size_t)N * pdim));
    CU(h->d_pgene.reserve((size_t)N));
    CU(cudaMemcpyAsync(h->d_X.p, Xp, sizeof(double) * (size_t)N * pdim, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(h->d_pgene.p, gene, sizeof(int) * (size_t)N, cudaMemcpyHostToDevice, s));
    int rc;
    if (!h->periph_joint) {
        // the core integrated once, one thread per candidate (peripheral_kernels.cuh)
        if ((rc = launch_periph_march(h, N, h->d_X.p, h->d_pgene.p, s))) return rc;
        CU(cudaMemcpyAsync(cost, h->main.cost.p, sizeof(double) * (size_t)N, cudaMemcpyDeviceToHost, s));
        if (status) CU(cudaMemcpyAsync(status, h->d_pstatus.p, sizeof(int) * (size_t)N,
                                       cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        return LASSIM_OK;
    }
    // the joint (n+1)-state system per candidate in the cost kernel (cross-check)
    h->last_periph_march = false;
    if ((rc = launch_eval(h, h->main, N, 0, h->d_X.p, nullptr, s, h->d_pgene.p))) return rc;
    const ZeroLayout z = zero_layout(N);
    CU(cudaMemcpyAsync(cost, h->main.cost.p, sizeof(double) * (size_t)N, cudaMemcpyDeviceToHost, s));
    if (status) CU(cudaMemcpyAsync(status, h->main.zero.p + z.off_status, sizeof(int) * (size_t)N,
                                   cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    return LASSIM_OK;
}

int lassim_peripheral_trajectories(lassim_handle *h, int N, const int *gene, const double *Xp,
                                   double *Y) {
    if (!h) return LASSIM_E_INVALID;
    if (N > 0 && !Y) return fail(h, LASSIM_E_INVALID, "peripheral_trajectories: Y is NULL");
    std::vector<double> cost((size_t)(N > 0 ? N : 0));
    int rc = lassim_eval_peripheral(h, N, gene, Xp, cost.data(), nullptr);
    if (rc || N == 0) return rc;
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    if (h->last_periph_march) {
        // core columns from the one core solve, last column from the candidates' samples
        const int n = h->n, T = h->T;
        std::vector<double> core((size_t)T * n), ys((size_t)T * N);
        CU(cudaMemcpy(core.data(), h->d_core_samples.p, sizeof(double) * core.size(), cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(ys.data(), h->d_pYs.p, sizeof(double) * ys.size(), cudaMemcpyDeviceToHost));
        for (int c = 0; c < N; ++c)
            for (int t = 0; t < T; ++t) {
                double *row = Y + ((size_t)c * T + t) * (n + 1);
                for (int j = 0; j < n; ++j) row[j] = core[(size_t)t * n + j];
                row[n] = ys[(size_t)t * N + c];
            }
        return LASSIM_OK;
    }
    CU(cudaMemcpy(Y, h->main.Y.p, sizeof(double) * (size_t)N * h->T * (h->n + 1),
                  cudaMemcpyDeviceToHost));
    return LASSIM_OK;
}

int lassim_last_end_states(lassim_handle *h, int N, double *E) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->last_had_E || N != h->last_N || !E)
        return fail(h, LASSIM_E_STATE, "no knock-down end states for N=%d (last N=%d)", N, h->last_N);
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(E, h->main.E.p, sizeof(double) * (size_t)N * h->K * h->n, cudaMemcpyDeviceToHost));
    return LASSIM_OK;
}

int lassim_last_traj_steps(lassim_handle *h, int N, unsigned int *steps) {
    if (!h) return LASSIM_E_INVALID;
    if (N != h->last_N || !steps)
        return fail(h, LASSIM_E_STATE, "no step counts for N=%d (last N=%d)", N, h->last_N);
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    CU(cudaDeviceSynchronize());
    if (h->last_periph_march) {
        // one solve per candidate: its attempts in slot 0, the other K - 1 slots zero
        std::vector<unsigned> one((size_t)N);
        CU(cudaMemcpy(one.data(), h->d_psteps.p, sizeof(unsigned) * (size_t)N, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < (size_t)N * h->K; ++i) steps[i] = 0u;
        for (int i = 0; i < N; ++i) steps[(size_t)i * h->K] = one[i];
        return LASSIM_OK;
    }
    CU(cudaMemcpy(steps, h->main.steps.p, sizeof(unsigned) * (size_t)N * h->K, cudaMemcpyDeviceToHost));
    return LASSIM_OK;
}

long long lassim_launch_count(const lassim_handle *h) { return h ? h->launches : -1; }

void *lassim_stream(const lassim_handle *h) { return h ? (void *)h->stream : nullptr; }

int lassim_last_kernel_ms(lassim_handle *h, float *ms) {
    if (!h || !ms) return LASSIM_E_INVALID;
    if (!h->timed) return fail(h, LASSIM_E_STATE, "no kernel has been launched yet");
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    CU(cudaEventSynchronize(h->ev1));
    CU(cudaEventElapsedTime(ms, h->ev0, h->ev1));
    return LASSIM_OK;
}

int lassim_last_stage_parallel_ms(lassim_handle *h, float *ms) {
    if (!h || !ms) return LASSIM_E_INVALID;
    if (!h->sp_timed)
        return fail(h, LASSIM_E_STATE, "the last evaluation did not launch the stage-parallel kernel");
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    CU(cudaEventSynchronize(h->ev_sp1));
    CU(cudaEventElapsedTime(ms, h->ev_sp0, h->ev_sp1));
    return LASSIM_OK;
}

int lassim_last_capped_phase_ms(lassim_handle *h, float *ms) {
    if (!h || !ms) return LASSIM_E_INVALID;
    if (!h->sp_timed || !h->timed)
        return fail(h, LASSIM_E_STATE, "the last evaluation did not launch the stage-parallel kernel");
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    CU(cudaEventSynchronize(h->ev_sp0));
    CU(cudaEventElapsedTime(ms, h->ev0, h->ev_sp0));
    return LASSIM_OK;
}

int lassim_fp64_peak_probe(lassim_handle *h, int iters, double *tflops) {
    if (!h || !tflops || iters < 1) return LASSIM_E_INVALID;
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    const int blocks = h->sm_count * 8, threads = 256;
    CU(h->d_probe.reserve((size_t)blocks * threads));
    cudaStream_t s = h->stream;
    fp64_probe_kernel<<<blocks, threads, 0, s>>>(h->d_probe.p, iters / 8 + 1, 0.5);  // warm-up
    CU(cudaGetLastError());
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        CU(cudaEventRecord(h->ev0, s));
        fp64_probe_kernel<<<blocks, threads, 0, s>>>(h->d_probe.p, iters, 0.5);
        CU(cudaGetLastError());
        CU(cudaEventRecord(h->ev1, s));
        CU(cudaEventSynchronize(h->ev1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
        if (ms < best) best = ms;
        h->launches += 1;
    }
    h->timed = false;
    const double flops = 2.0 * 8.0 * (double)iters * (double)blocks * threads;
    *tflops = flops / (best * 1e-3) / 1e12;
    return LASSIM_OK;
}

}  // extern "C"


// ---------------------------------------------------------------------------
// Differential evolution on the device (de_kernels.cuh): the populations of all
// islands of this GPU stay in HBM; one evolve round is `gen` x (trial kernel ->
// evaluation kernel(s) -> selection kernel) on the handle's stream, no host
// round trip.  Replaces the per-individual objfun callbacks of PyGMO's DE
// (source/core/base_optimization.py:173-174 drives them).
// ---------------------------------------------------------------------------
namespace {

int de_evaluate_population(lassim_handle *h) {
    const int count = h->de.islands * h->de.np;
    cudaStream_t s = h->stream;
    int rc = launch_eval(h, h->main, count, 0, h->de_X.p, nullptr, s);
    if (rc) return rc;
    const ZeroLayout z = zero_layout(count);
    de_mask_status_kernel<<<(count + 255) / 256, 256, 0, s>>>(
        count, reinterpret_cast<const int *>(h->main.zero.p + z.off_status), h->main.cost.p);
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(h->de_F.p, h->main.cost.p, sizeof(double) * (size_t)count,
                       cudaMemcpyDeviceToDevice, s));
    h->launches += 1;
    h->de_evaluations += count;
    h->de_evaluated = true;
    return LASSIM_OK;
}

// Lanes of the asynchronous evolve: `lanes` groups of consecutive islands, each with its own
// stream and evaluation scratch.  A generation of a small archipelago is over in the time of
// its slowest time series, and lanes fill that wait; the fuller the machine is by one batch
// per generation, the less there is to fill, and lanes then only take SMs from each other's
// long time series.  Automatic (Th2, 10 generations, tools/gpu_lane_policy.py; ms in lockstep /
// with the lanes chosen here): up to 1024 individuals in all one island per lane up to 16
// (8 x 128: 205 / 169, 16 x 64: 169 / 133), up to 2048 at most 8 lanes (16 x 128: 234 / 211,
// 32 x 64: 218 / 204), up to 3072 at most 4 (24 x 128: 269 / 260), lockstep beyond (32 x 128:
// 300 against 389 in 16 lanes; 8 x 1024: 809 against 1338 in 8).
int de_make_lanes(lassim_handle *h) {
    release_de_lanes(h);
    const int islands = h->de.islands;
    const long long total = (long long)islands * h->de.np;
    const int most = total <= 1024 ? 16 : total <= 2048 ? 8 : total <= 3072 ? 4 : 1;
    int lanes = h->de_lanes_wanted > 0 ? h->de_lanes_wanted : (islands < most ? islands : most);
    if (lanes > islands) lanes = islands;
    h->de_lanes.resize(lanes);
    for (int l = 0; l < lanes; ++l) {
        DeLane &lane = h->de_lanes[l];
        lane.first = (int)((long long)islands * l / lanes);
        lane.count = (int)((long long)islands * (l + 1) / lanes) - lane.first;
        CU(cudaStreamCreateWithPriority(&lane.stream, cudaStreamNonBlocking, h->prio_hi));
        CU(cudaEventCreateWithFlags(&lane.ev_done, cudaEventDisableTiming));
        CU(lane.scratch.init(h->prio_lo));
    }
    return LASSIM_OK;
}

}  // namespace

extern "C" int lassim_de_init(lassim_handle *h, int islands, int individuals, int variant,
                              double f, double cr, unsigned long long seed, const double *lb,
                              const double *ub, const double *X0) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->has_problem) return fail(h, LASSIM_E_STATE, "de_init before set_problem");
    if (islands < 1 || individuals < 6 || variant < 1 || variant > 10 || !lb || !ub || !X0 ||
        !(f >= 0.0 && f <= 1.0) || !(cr >= 0.0 && cr <= 1.0))
        return fail(h, LASSIM_E_INVALID,
                    "de_init: islands=%d individuals=%d (>= 6) variant=%d (1..10) f=%g cr=%g "
                    "(both in [0,1]) or a NULL array", islands, individuals, variant, f, cr);
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    const size_t dim = (size_t)h->dim, count = (size_t)islands * individuals;
    CU(cudaStreamSynchronize(h->stream));        // nothing of a previous run is in flight
    CU(h->de_X.reserve(count * dim));
    CU(h->de_trial.reserve(count * dim));
    CU(h->de_F.reserve(count));
    CU(h->de_best.reserve((size_t)islands * (dim + 1)));
    // room for the migration table of an archipelago of up to 64 islands (this handle's and the
    // other ranks'): the first lassim_de_migrate_ring then allocates nothing - a cudaMalloc
    // there would wait for whatever the device is running, a collective in flight included
    CU(h->de_table.reserve((size_t)(islands > 64 ? islands : 64) * (dim + 1)));
    int rc;
    if ((rc = upload(h, h->de_lb, lb, dim))) return rc;
    if ((rc = upload(h, h->de_ub, ub, dim))) return rc;
    CU(cudaMemcpyAsync(h->de_X.p, X0, sizeof(double) * count * dim, cudaMemcpyHostToDevice,
                       h->stream));
    CU(cudaStreamSynchronize(h->stream));
    h->de = DeParams{islands, individuals, (int)dim, variant, f, cr, seed, 0ull};
    h->de.gid_base = 0; h->de.gid_stride = 1;
    h->de.ind_begin = 0; h->de.ind_end = (int)count;
    if ((rc = de_make_lanes(h))) return rc;
    h->de_ready = true;
    h->de_evaluated = false;
    h->de_evaluations = 0;
    return LASSIM_OK;
}

extern "C" int lassim_de_set_island_ids(lassim_handle *h, int first_island, int stride) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->de_ready) return fail(h, LASSIM_E_STATE, "de_set_island_ids before de_init");
    if (first_island < 0 || stride < 1)
        return fail(h, LASSIM_E_INVALID, "de_set_island_ids: first island %d, stride %d",
                    first_island, stride);
    h->de.gid_base = first_island;
    h->de.gid_stride = stride;
    return LASSIM_OK;
}

extern "C" int lassim_de_set_lanes(lassim_handle *h, int lanes) {
    if (!h) return LASSIM_E_INVALID;
    if (lanes < 0) return fail(h, LASSIM_E_INVALID, "de_set_lanes: lanes=%d", lanes);
    h->de_lanes_wanted = lanes;
    if (h->de_ready) {
        DeviceGuard guard_(h->device);
        CU(guard_.err);
        CU(cudaDeviceSynchronize());
        return de_make_lanes(h);
    }
    return LASSIM_OK;
}

extern "C" int lassim_de_set_individual(lassim_handle *h, int island, int index, const double *x) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->de_ready) return fail(h, LASSIM_E_STATE, "de_set_individual before de_init");
    if (island < 0 || island >= h->de.islands || index < 0 || index >= h->de.np || !x)
        return fail(h, LASSIM_E_INVALID, "de_set_individual: island %d index %d", island, index);
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    CU(cudaMemcpyAsync(h->de_X.p + ((size_t)island * h->de.np + index) * h->de.dim, x,
                       sizeof(double) * h->de.dim, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    h->de_evaluated = false;          // the whole population is re-evaluated next time
    return LASSIM_OK;
}

// One evolve round.  Islands do not wait for each other inside a round (PyGMO evolves every
// island in a worker of its own and joins them at the end,
// source/core/base_optimization.py:173-174): each lane - a group of islands - runs its
// `generations` x (trial vectors -> evaluation -> selection) on its own stream with its own
// scratch, so one island's slow candidate holds up that island's next generation only while
// the other islands' kernels fill the machine.  The lanes fork from the handle's stream and
// join it again at the end (migration and the getters follow on that stream).  A
// candidate's cost does not depend on the batch it is evaluated in, so the lanes change
// the schedule, not the result.
extern "C" int lassim_de_evolve(lassim_handle *h, int generations) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->de_ready) return fail(h, LASSIM_E_STATE, "de_evolve before de_init");
    if (h->de.dim != h->dim)
        return fail(h, LASSIM_E_STATE, "the problem dimension changed since de_init (%d -> %d)",
                    h->de.dim, h->dim);
    if (generations < 0) return fail(h, LASSIM_E_INVALID, "de_evolve: generations=%d", generations);
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    int rc;
    if (!h->de_evaluated && (rc = de_evaluate_population(h))) return rc;
    if (generations == 0) return LASSIM_OK;
    CU(cudaEventRecord(h->ev_de_fork, h->stream));
    for (DeLane &lane : h->de_lanes) CU(cudaStreamWaitEvent(lane.stream, h->ev_de_fork, 0));
    const int np = h->de.np, dim = h->de.dim;
    for (int g = 0; g < generations; ++g) {
        for (DeLane &lane : h->de_lanes) {
            cudaStream_t s = lane.stream;
            const int count = lane.count * np;
            DeParams p = h->de;
            p.ind_begin = lane.first * np;
            p.ind_end = p.ind_begin + count;
            const int blocks = (count + 127) / 128;
            de_trial_kernel<<<blocks, 128, 0, s>>>(p, h->de_X.p, h->de_F.p, h->de_lb.p,
                                                    h->de_ub.p, h->de_trial.p);
            CU(cudaGetLastError());
            const double *trial = h->de_trial.p + (size_t)p.ind_begin * dim;
            if ((rc = launch_eval(h, lane.scratch, count, 0, trial, nullptr, s))) return rc;
            const ZeroLayout z = zero_layout(count);
            de_mask_status_kernel<<<(count + 255) / 256, 256, 0, s>>>(
                count, reinterpret_cast<const int *>(lane.scratch.zero.p + z.off_status),
                lane.scratch.cost.p);
            de_select_kernel<<<blocks, 128, 0, s>>>(p, h->de_X.p, h->de_F.p, h->de_trial.p,
                                                     lane.scratch.cost.p);
            CU(cudaGetLastError());
            h->launches += 3;
            h->de_evaluations += count;
        }
        h->de.generation += 1;
    }
    for (DeLane &lane : h->de_lanes) {
        CU(cudaEventRecord(lane.ev_done, lane.stream));
        CU(cudaStreamWaitEvent(h->stream, lane.ev_done, 0));
    }
    return LASSIM_OK;       // enqueued; lassim_de_get / lassim_de_best synchronise
}

extern "C" int lassim_de_best(lassim_handle *h, double *table) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->de_ready || !table) return fail(h, LASSIM_E_STATE, "de_best before de_init / NULL");
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    int rc;
    if (!h->de_evaluated && (rc = de_evaluate_population(h))) return rc;
    de_best_kernel<<<h->de.islands, 32, 0, h->stream>>>(h->de, h->de_X.p, h->de_F.p, h->de_best.p);
    CU(cudaGetLastError());
    h->launches += 1;
    CU(cudaMemcpyAsync(table, h->de_best.p, sizeof(double) * (size_t)h->de.islands * (h->de.dim + 1),
                       cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return LASSIM_OK;
}

extern "C" int lassim_de_migrate_ring(lassim_handle *h, const double *table, int n_table,
                                      int first_island, int stride) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->de_ready || !table) return fail(h, LASSIM_E_STATE, "de_migrate before de_init / NULL");
    if (stride < 1 || first_island < 0 ||
        first_island + (h->de.islands - 1) * stride >= n_table)
        return fail(h, LASSIM_E_INVALID,
                    "de_migrate: table of %d islands, mine are %d from %d with stride %d",
                    n_table, h->de.islands, first_island, stride);
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    int rc;
    if ((rc = upload(h, h->de_table, table, (size_t)n_table * (h->de.dim + 1)))) return rc;
    de_migrate_ring_kernel<<<h->de.islands, 32, 0, h->stream>>>(h->de, h->de_X.p, h->de_F.p,
                                                                h->de_table.p, n_table, first_island, stride);
    CU(cudaGetLastError());
    h->launches += 1;
    CU(cudaStreamSynchronize(h->stream));      // `table` is the caller's
    return LASSIM_OK;
}

extern "C" int lassim_de_get(lassim_handle *h, double *X, double *F) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->de_ready) return fail(h, LASSIM_E_STATE, "de_get before de_init");
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    int rc;
    if (!h->de_evaluated && (rc = de_evaluate_population(h))) return rc;
    const size_t count = (size_t)h->de.islands * h->de.np;
    if (X) CU(cudaMemcpyAsync(X, h->de_X.p, sizeof(double) * count * h->de.dim,
                              cudaMemcpyDeviceToHost, h->stream));
    if (F) CU(cudaMemcpyAsync(F, h->de_F.p, sizeof(double) * count, cudaMemcpyDeviceToHost,
                              h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return LASSIM_OK;
}

extern "C" long long lassim_de_evaluations(const lassim_handle *h) {
    return h ? h->de_evaluations : -1;
}

// ---------------------------------------------------------------------------
// Evaluation without any allocation.
//
// Some callers of this library must not, or cannot, allocate device memory at
// the moment they want a population evaluated: a host process in which another
// framework's caching allocator already owns (almost) all of the GPU, so a raw
// cudaMalloc from a library fails or fragments its pool; a serving or control
// loop that cannot afford the device-wide synchronisation cudaMalloc/cudaFree
// imply while other streams are busy; a job confined to a hard memory budget
// (MIG slice, MPS limit) that was sized without this library's scratch.  For
// them the scratch of the evaluation kernel lives in a fixed array that is part
// of the loaded library image: every GPU that loads liblassim_b200.so holds
// LASSIM_NOALLOC_SCRATCH_BYTES of it until the process exits.
//
// In host code the array's name is only a stand-in for the symbol, not a
// device address; the address is resolved once per device with
// cudaGetSymbolAddress and cached here.
// ---------------------------------------------------------------------------
#define LASSIM_NOALLOC_SCRATCH_BYTES (8u << 20)
#define LASSIM_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_noalloc_scratch[LASSIM_NOALLOC_SCRATCH_BYTES];
static unsigned char *g_noalloc_addr[LASSIM_NOALLOC_MAX_DEVICES];   // per device, set once

namespace {

struct NoallocLayout {
    size_t off_done, off_status, off_counters, zero_bytes;
    size_t off_cost_main, off_pert_cost, off_steps, off_long, off_Y, off_E, bytes;
};

size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

NoallocLayout noalloc_layout(int N, int T, int K, int n) {
    NoallocLayout L;
    L.off_done = 32;                                       // [0,32): work counters, hand-over count, attempts
    L.off_status = L.off_done + sizeof(int) * (size_t)N;
    L.off_counters = (L.off_status + sizeof(int) * (size_t)N + 15) & ~(size_t)15;
    L.zero_bytes = L.off_counters + sizeof(unsigned long long) * 3 * (size_t)N;
    L.off_cost_main = align256(L.zero_bytes);
    L.off_pert_cost = align256(L.off_cost_main + sizeof(double) * (size_t)N);
    L.off_steps = align256(L.off_pert_cost + sizeof(double) * (size_t)N);
    L.off_long = align256(L.off_steps + sizeof(unsigned) * (size_t)N * K);
    L.off_Y = align256(L.off_long + sizeof(unsigned long long) * (size_t)N);
    L.off_E = align256(L.off_Y + sizeof(double) * (size_t)N * T * n);
    L.bytes = L.off_E + sizeof(double) * (size_t)N * K * n;
    return L;
}

}  // namespace

extern "C" int lassim_eval_device_noalloc(lassim_handle *h, int N, const double *dX,
                                          double *dcost, void *stream) {
    if (!h) return LASSIM_E_INVALID;
    if (!h->has_problem) return fail(h, LASSIM_E_STATE, "eval before set_problem");
    if (N < 0 || (N > 0 && (!dX || !dcost)))
        return fail(h, LASSIM_E_INVALID, "eval_device_noalloc: N=%d, X or cost is NULL", N);
    if (N == 0) return LASSIM_OK;
    if (h->device < 0 || h->device >= LASSIM_NOALLOC_MAX_DEVICES)
        return fail(h, LASSIM_E_UNSUPPORTED, "device index %d is beyond the %d devices the "
                    "allocation-free path keeps an address for", h->device,
                    LASSIM_NOALLOC_MAX_DEVICES);
    const int K = h->K, n = h->n, T = h->T;
    const NoallocLayout L = noalloc_layout(N, T, K, n);
    if (L.bytes > (size_t)LASSIM_NOALLOC_SCRATCH_BYTES) {
        const size_t per = 8 * ((size_t)T * n + (size_t)K * n) + 4 * (size_t)K + 64;
        return fail(h, LASSIM_E_INVALID,
                    "eval_device_noalloc: %d candidates of this problem need %zu bytes of "
                    "scratch but the built-in scratch holds %u bytes (about %zu candidates); "
                    "split the batch or use lassim_eval_device",
                    N, L.bytes, (unsigned)LASSIM_NOALLOC_SCRATCH_BYTES,
                    ((size_t)LASSIM_NOALLOC_SCRATCH_BYTES - 2048) / per);
    }
    DeviceGuard guard_(h->device);
    CU(guard_.err);
    // one-time set-up on this device: turn the symbol into a device address
    if (!g_noalloc_addr[h->device]) {
        unsigned char *addr = nullptr;
        CU(cudaGetSymbolAddress((void **)&addr, g_noalloc_scratch));
        if (!addr)
            return fail(h, LASSIM_E_CUDA, "cudaGetSymbolAddress returned a null address");
        g_noalloc_addr[h->device] = addr;
    }
    unsigned char *base = g_noalloc_addr[h->device];
    KernelChoice kc;
    if (!choose_kernel(h, &kc))
        return fail(h, LASSIM_E_UNSUPPORTED,
                    "state size n=%d is outside the kernels built (n <= 1024)", n);
    cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
    // the array keeps whatever the previous call left: clear what must be zero
    CU(cudaMemsetAsync(base, 0, L.zero_bytes, s));

    DevProblem P;
    P.n = n; P.T = T; P.K = K; P.U = h->U; P.dim = h->dim; P.nnz = h->nnz;
    P.times = h->d_times.p; P.y0 = h->d_y0.p; P.data = h->d_data.p;
    P.sigma2 = h->d_sigma2.p; P.kidx = h->d_kidx.p;
    P.row_ptr = h->d_row_ptr.p; P.col_idx = h->d_col_idx.p;
    P.pert = (K > 1) ? h->d_pert.p : nullptr;
    P.traj_prefix = h->d_traj_prefix.p; P.traj_tt = h->d_traj_tt.p;
    P.knock_ctrl = h->d_knock_ctrl.p;
    P.pert_factor = h->pert_factor; P.rtol = h->rtol; P.atol = h->atol;
    P.max_steps = (unsigned int)(h->max_steps > 4000000000LL ? 4000000000LL : h->max_steps);
    P.swing = h->jump_guard;

    P.core_x = nullptr; P.pmask = nullptr; P.pdata = nullptr; P.psigma2 = nullptr; P.py0 = nullptr;

    DevBatch B;
    B.N = N; B.mode = 0; B.X = dX; B.pgene = nullptr;
    B.work_begin = 0; B.work_count = (unsigned)N * (unsigned)K;
    B.cost = dcost;                                           // straight into the caller's buffer
    B.work_counter = reinterpret_cast<unsigned int *>(base);
    B.done = reinterpret_cast<int *>(base + L.off_done);
    B.status = reinterpret_cast<int *>(base + L.off_status);
    B.counters = reinterpret_cast<unsigned long long *>(base + L.off_counters);
    B.cost_main = reinterpret_cast<double *>(base + L.off_cost_main);
    B.pert_cost = reinterpret_cast<double *>(base + L.off_pert_cost);
    B.traj_steps = reinterpret_cast<unsigned int *>(base + L.off_steps);
    B.Y = reinterpret_cast<double *>(base + L.off_Y);
    B.E = reinterpret_cast<double *>(base + L.off_E);
    B.order = nullptr;
    B.long_list = reinterpret_cast<unsigned long long *>(base + L.off_long);

    // same kernels, same speculation policy as lassim_eval_device: same costs
    int rc = launch_kernels(h, kc, P, B, N, K, false, s, reinterpret_cast<unsigned int *>(base),
                            h->main, false);
    if (rc) return rc;
    return LASSIM_OK;
}

