"""ctypes binding of liblassim_b200.so (include/lassim_b200.h).

There is no CPU fallback: if the shared library is missing or a CUDA call
fails, these functions raise."""
import ctypes
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_PKG, "liblassim_b200.so")
_lib = None

c_double_p = ctypes.POINTER(ctypes.c_double)
c_int_p = ctypes.POINTER(ctypes.c_int)
c_uint_p = ctypes.POINTER(ctypes.c_uint)
c_ubyte_p = ctypes.POINTER(ctypes.c_ubyte)
c_ull_p = ctypes.POINTER(ctypes.c_ulonglong)

ST_STEP_BUDGET = 1
ST_NONFINITE = 2


class LassimError(RuntimeError):
    pass


def library_path():
    return _LIB_PATH


def load_library():
    """Load liblassim_b200.so (built in-tree by lassim_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise LassimError(
            "liblassim_b200.so is not built (%s); run `python -m "
            "lassim_b200.build` - there is no CPU fallback" % _LIB_PATH)
    # islands evolve concurrently on up to 16 x 2 streams: with the default of 8 hardware
    # queues streams would share queues and wait for each other.  Only read when the CUDA
    # context is created, so a host that initialised CUDA earlier keeps its own setting.
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
    lib = ctypes.CDLL(_LIB_PATH)
    H = ctypes.c_void_p
    lib.lassim_version.restype = ctypes.c_char_p
    lib.lassim_last_error.restype = ctypes.c_char_p
    lib.lassim_last_error.argtypes = [H]
    lib.lassim_create.argtypes = [ctypes.c_int, ctypes.POINTER(H)]
    lib.lassim_destroy.argtypes = [H]
    lib.lassim_set_problem.argtypes = [H, ctypes.c_int, ctypes.c_int, c_double_p,
                                       c_double_p, c_double_p, c_double_p,
                                       c_ubyte_p, c_double_p, ctypes.c_double]
    lib.lassim_set_mask.argtypes = [H, c_ubyte_p]
    lib.lassim_set_tolerances.argtypes = [H, ctypes.c_double, ctypes.c_double,
                                          ctypes.c_longlong]
    lib.lassim_set_speculation.argtypes = [H, ctypes.c_int]
    lib.lassim_set_speculation.restype = ctypes.c_int
    lib.lassim_set_series_cap.argtypes = [H, ctypes.c_uint]
    lib.lassim_set_series_cap.restype = ctypes.c_int
    lib.lassim_set_jump_guard.argtypes = [H, ctypes.c_double]
    lib.lassim_set_jump_guard.restype = ctypes.c_int
    lib.lassim_set_method.argtypes = [H, ctypes.c_int]
    lib.lassim_set_method.restype = ctypes.c_int
    lib.lassim_dim.argtypes = [H]
    lib.lassim_trajectories_per_candidate.argtypes = [H]
    lib.lassim_eval.argtypes = [H, ctypes.c_int, c_double_p, c_double_p,
                                c_double_p, c_double_p, c_int_p, c_ull_p]
    lib.lassim_eval_device.argtypes = [H, ctypes.c_int, ctypes.c_void_p,
                                       ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_void_p, ctypes.c_void_p]
    lib.lassim_eval_device_noalloc.argtypes = [H, ctypes.c_int, ctypes.c_void_p,
                                               ctypes.c_void_p, ctypes.c_void_p]
    lib.lassim_eval_device_noalloc.restype = ctypes.c_int
    lib.lassim_trajectories.argtypes = [H, ctypes.c_int, c_double_p, c_double_p,
                                        c_int_p]
    lib.lassim_set_peripheral.argtypes = [H, c_double_p, ctypes.c_int, c_double_p, c_double_p,
                                          c_double_p, c_ubyte_p]
    lib.lassim_eval_peripheral.argtypes = [H, ctypes.c_int, c_int_p, c_double_p, c_double_p,
                                           c_int_p]
    lib.lassim_peripheral_trajectories.argtypes = [H, ctypes.c_int, c_int_p, c_double_p,
                                                   c_double_p]
    lib.lassim_set_peripheral_joint.argtypes = [H, ctypes.c_int]
    lib.lassim_set_peripheral_joint.restype = ctypes.c_int
    for name in ("lassim_set_peripheral", "lassim_eval_peripheral",
                 "lassim_peripheral_trajectories"):
        getattr(lib, name).restype = ctypes.c_int
    lib.lassim_de_init.argtypes = [H, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                   ctypes.c_double, ctypes.c_ulonglong, c_double_p, c_double_p,
                                   c_double_p]
    lib.lassim_de_set_individual.argtypes = [H, ctypes.c_int, ctypes.c_int, c_double_p]
    lib.lassim_de_evolve.argtypes = [H, ctypes.c_int]
    lib.lassim_de_set_island_ids.argtypes = [H, ctypes.c_int, ctypes.c_int]
    lib.lassim_de_set_lanes.argtypes = [H, ctypes.c_int]
    lib.lassim_de_best.argtypes = [H, c_double_p]
    lib.lassim_de_migrate_ring.argtypes = [H, c_double_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    lib.lassim_de_get.argtypes = [H, c_double_p, c_double_p]
    lib.lassim_de_evaluations.argtypes = [H]
    lib.lassim_de_evaluations.restype = ctypes.c_longlong
    for name in ("lassim_de_init", "lassim_de_set_individual", "lassim_de_evolve",
                 "lassim_de_set_island_ids", "lassim_de_set_lanes",
                 "lassim_de_best", "lassim_de_migrate_ring", "lassim_de_get"):
        getattr(lib, name).restype = ctypes.c_int
    lib.lassim_last_end_states.argtypes = [H, ctypes.c_int, c_double_p]
    lib.lassim_last_traj_steps.argtypes = [H, ctypes.c_int, c_uint_p]
    lib.lassim_flops_per_attempt.argtypes = [H]
    lib.lassim_flops_per_attempt.restype = ctypes.c_longlong
    lib.lassim_kernel_name.argtypes = [H]
    lib.lassim_kernel_name.restype = ctypes.c_char_p
    lib.lassim_series_flops_per_attempt.argtypes = [H, ctypes.c_int]
    lib.lassim_series_flops_per_attempt.restype = ctypes.c_longlong
    lib.lassim_series_kernel_name.argtypes = [H, ctypes.c_int]
    lib.lassim_series_kernel_name.restype = ctypes.c_char_p
    lib.lassim_last_stage_parallel_attempts.argtypes = [H, c_ull_p]
    lib.lassim_last_stage_parallel_attempts.restype = ctypes.c_int
    lib.lassim_last_stage_parallel_ms.argtypes = [H, ctypes.POINTER(ctypes.c_float)]
    lib.lassim_last_stage_parallel_ms.restype = ctypes.c_int
    lib.lassim_last_capped_phase_ms.argtypes = [H, ctypes.POINTER(ctypes.c_float)]
    lib.lassim_last_capped_phase_ms.restype = ctypes.c_int
    lib.lassim_stream.argtypes = [H]
    lib.lassim_stream.restype = ctypes.c_void_p
    lib.lassim_launch_count.argtypes = [H]
    lib.lassim_launch_count.restype = ctypes.c_longlong
    lib.lassim_last_kernel_ms.argtypes = [H, ctypes.POINTER(ctypes.c_float)]
    lib.lassim_fp64_peak_probe.argtypes = [H, ctypes.c_int, c_double_p]
    for name in ("lassim_create", "lassim_destroy", "lassim_set_problem",
                 "lassim_set_mask", "lassim_set_tolerances", "lassim_dim",
                 "lassim_trajectories_per_candidate", "lassim_eval",
                 "lassim_eval_device", "lassim_trajectories",
                 "lassim_last_end_states", "lassim_last_traj_steps",
                 "lassim_last_kernel_ms", "lassim_fp64_peak_probe"):
        getattr(lib, name).restype = ctypes.c_int
    _lib = lib
    return lib


def _dp(a):
    return a.ctypes.data_as(c_double_p)


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None and a.shape != shape:
        raise ValueError("expected shape %s, got %s" % (shape, a.shape))
    return a


class Engine:
    """One evaluation engine on one CUDA device (a lassim_handle)."""

    def __init__(self, device=0):
        self._lib = load_library()
        self._h = ctypes.c_void_p()
        rc = self._lib.lassim_create(int(device), ctypes.byref(self._h))
        if rc != 0:
            msg = self._lib.lassim_last_error(None).decode()
            self._h = None
            raise LassimError("lassim_create(device=%d) failed: %s" % (device, msg))
        self.device = int(device)
        self.n = self.T = 0
        self.method = "dop853"

    # -- plumbing ----------------------------------------------------------
    def _check(self, rc, what):
        if rc != 0:
            raise LassimError("%s failed (%d): %s" % (
                what, rc, self._lib.lassim_last_error(self._h).decode()))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.lassim_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- problem -----------------------------------------------------------
    def set_problem(self, data, sigma, times, y0, mask, pert=None, pert_factor=0.0):
        y0 = _f64(y0).reshape(-1)
        n = y0.size
        times = _f64(times).reshape(-1)
        T = times.size
        data = _f64(data)
        if data.shape != (T, n):
            raise ValueError("data shape %s is not (T=%d, n=%d)" % (data.shape, T, n))
        sigma = _f64(np.broadcast_to(_f64(sigma), (n,)))
        mask = np.ascontiguousarray(np.asarray(mask).reshape(-1) != 0, dtype=np.uint8)
        if mask.size != n * n:
            raise ValueError("mask has %d entries, expected n*n=%d" % (mask.size, n * n))
        pp = None
        if pert is not None:
            pert = _f64(pert)
            if pert.shape != (n, n + 2):
                raise ValueError("perturbation data shape %s is not (n, n+2)=(%d, %d)"
                                 % (pert.shape, n, n + 2))
            pp = _dp(pert)
        self._check(self._lib.lassim_set_problem(
            self._h, n, T, _dp(times), _dp(y0), _dp(data), _dp(sigma),
            mask.ctypes.data_as(c_ubyte_p), pp, float(pert_factor)), "lassim_set_problem")
        self.n, self.T = n, T

    def set_mask(self, mask):
        mask = np.ascontiguousarray(np.asarray(mask).reshape(-1) != 0, dtype=np.uint8)
        if mask.size != self.n * self.n:
            raise ValueError("mask has %d entries, expected %d" % (mask.size, self.n ** 2))
        self._check(self._lib.lassim_set_mask(self._h, mask.ctypes.data_as(c_ubyte_p)),
                    "lassim_set_mask")

    def set_tolerances(self, rtol=1.49012e-8, atol=1.49012e-8, max_steps=20000000):
        self._check(self._lib.lassim_set_tolerances(self._h, rtol, atol, int(max_steps)),
                    "lassim_set_tolerances")

    def set_speculation(self, mode="auto"):
        """Kernel for the time-series trajectories: 'off' (the standard one), 'on'
        (speculative half-warp attempts), 'stages' (stage-parallel Gauss pair, one warp
        per trajectory), 'stages2' / 'stages3' (the same with two / three warps attempting
        different step sizes), 'handover' (standard kernel first, long time series continue
        in 'stages3') or 'auto' ('handover' up to 16,384 candidates, 'off' beyond)."""
        code = {"off": 0, "on": 1, "auto": 2, "stages": 3, "stages2": 4, "handover": 5, "stages3": 6,
                0: 0, 1: 1, 2: 2, 3: 3, 4: 4, 5: 5, 6: 6, False: 0, True: 1}.get(mode)
        if code is None:
            raise ValueError("time-series kernel mode %r (use 'off', 'on', 'stages' or 'auto')"
                             % (mode,))
        self._check(self._lib.lassim_set_speculation(self._h, code), "lassim_set_speculation")

    def set_series_cap(self, attempts=0):
        """Attempts a time series makes in the standard kernel before 'handover' moves it
        to the stage-parallel kernel; 0 = automatic (768 up to 4096 candidates)."""
        self._check(self._lib.lassim_set_series_cap(self._h, int(attempts)), "lassim_set_series_cap")

    def set_jump_guard(self, swing=0.5):
        """A step of the explicit pairs in which a logistic term swings by more than `swing`
        is charged hs * vmax * swing as its error (2 or more = off); see lassim_b200.h."""
        self._check(self._lib.lassim_set_jump_guard(self._h, float(swing)), "lassim_set_jump_guard")

    def set_method(self, method="dop853"):
        """'dop853' (Dormand-Prince 8(5,3), default) or 'dp45' (Dormand-Prince 5(4))."""
        code = {"dp45": 0, "dop853": 1}.get(str(method).lower())
        if code is None:
            raise ValueError("unknown integrator %r (use 'dp45' or 'dop853')" % (method,))
        self._check(self._lib.lassim_set_method(self._h, code), "lassim_set_method")
        self.method = str(method).lower()

    @property
    def dim(self):
        return self._lib.lassim_dim(self._h)

    @property
    def trajectories_per_candidate(self):
        return self._lib.lassim_trajectories_per_candidate(self._h)

    # -- evaluation ----------------------------------------------------------
    def eval(self, X, details=False):
        """cost[N] for host X[N, dim].  details=True returns a dict with
        cost, cost_main, pert_cost, status, counters."""
        X = _f64(X)
        if X.ndim == 1:
            X = X.reshape(1, -1)
        N = X.shape[0]
        if X.shape[1] != self.dim:
            raise ValueError("decision vectors have %d entries, problem dimension is %d"
                             % (X.shape[1], self.dim))
        cost = np.empty(N)
        if not details:
            self._check(self._lib.lassim_eval(self._h, N, _dp(X), _dp(cost), None, None,
                                              None, None), "lassim_eval")
            return cost
        cm = np.empty(N); pc = np.empty(N)
        st = np.empty(N, dtype=np.int32)
        cnt = np.empty((N, 3), dtype=np.uint64)
        self._check(self._lib.lassim_eval(
            self._h, N, _dp(X), _dp(cost), _dp(cm), _dp(pc),
            st.ctypes.data_as(c_int_p), cnt.ctypes.data_as(c_ull_p)), "lassim_eval")
        return dict(cost=cost, cost_main=cm, pert_cost=pc, status=st, counters=cnt)

    def eval_device(self, N, x_ptr, cost_ptr, cost_main_ptr=None, pert_cost_ptr=None,
                    status_ptr=None, counters_ptr=None, stream=None):
        """Device-resident evaluation: raw device addresses (ints), enqueued on
        `stream` (a cudaStream_t address; None = the engine's stream)."""
        self._check(self._lib.lassim_eval_device(
            self._h, int(N), x_ptr, cost_ptr, cost_main_ptr, pert_cost_ptr,
            status_ptr, counters_ptr, stream), "lassim_eval_device")

    def eval_device_noalloc(self, N, x_ptr, cost_ptr, stream=None):
        """Allocation-free variant of eval_device (cost only): for hosts that
        cannot call cudaMalloc at evaluation time.  Scratch is a fixed 8 MiB
        array inside the library; batches that do not fit raise LassimError.
        Calls that overlap on one GPU race on that array - order them."""
        self._check(self._lib.lassim_eval_device_noalloc(
            self._h, int(N), x_ptr, cost_ptr, stream), "lassim_eval_device_noalloc")

    def trajectories(self, X, return_status=False):
        X = _f64(X)
        if X.ndim == 1:
            X = X.reshape(1, -1)
        N = X.shape[0]
        if X.shape[1] != self.dim:
            raise ValueError("decision vectors have %d entries, problem dimension is %d"
                             % (X.shape[1], self.dim))
        Y = np.empty((N, self.T, self.n))
        st = np.empty(N, dtype=np.int32)
        self._check(self._lib.lassim_trajectories(self._h, N, _dp(X), _dp(Y),
                                                  st.ctypes.data_as(c_int_p)),
                    "lassim_trajectories")
        return (Y, st) if return_status else Y

    # -- peripheral genes against a fixed core ------------------------------
    def set_peripheral(self, core_x, data, sigma, y0, mask=None):
        """Peripheral genes of the problem currently set: core_x[dim], data[genes, T],
        sigma[genes], y0[genes], mask[genes, n] (None = every core gene may regulate)."""
        core_x = _f64(core_x).reshape(-1)
        if core_x.size != self.dim:
            raise ValueError("core vector has %d entries, the core problem has %d"
                             % (core_x.size, self.dim))
        data = _f64(data)
        if data.ndim != 2 or data.shape[1] != self.T:
            raise ValueError("peripheral data shape %s is not (genes, T=%d)" % (data.shape, self.T))
        genes = data.shape[0]
        sigma = _f64(np.broadcast_to(_f64(sigma), (genes,)))
        y0 = _f64(y0).reshape(-1)
        if y0.size != genes:
            raise ValueError("y0 has %d entries for %d genes" % (y0.size, genes))
        mp = None
        if mask is not None:
            mask = np.ascontiguousarray(np.asarray(mask) != 0, dtype=np.uint8)
            if mask.shape != (genes, self.n):
                raise ValueError("mask shape %s is not (genes, n)=(%d, %d)"
                                 % (mask.shape, genes, self.n))
            mp = mask.ctypes.data_as(c_ubyte_p)
        self._check(self._lib.lassim_set_peripheral(self._h, _dp(core_x), genes, _dp(data),
                                                    _dp(sigma), _dp(y0), mp),
                    "lassim_set_peripheral")
        self.p_genes = genes

    def set_peripheral_joint(self, joint=True):
        """True: integrate the joint (n+1)-state system per candidate (cross-check; call before
        set_peripheral); False (default): the core once, one thread per candidate."""
        self._check(self._lib.lassim_set_peripheral_joint(self._h, int(bool(joint))),
                    "lassim_set_peripheral_joint")

    def _peripheral_args(self, gene, Xp):
        Xp = _f64(Xp)
        if Xp.ndim == 1:
            Xp = Xp.reshape(1, -1)
        if Xp.shape[1] != self.n + 2:
            raise ValueError("peripheral candidates have %d entries, expected n + 2 = %d"
                             % (Xp.shape[1], self.n + 2))
        gene = np.ascontiguousarray(np.broadcast_to(np.asarray(gene, dtype=np.int32),
                                                    (Xp.shape[0],)))
        return gene, Xp

    def eval_peripheral(self, gene, Xp, return_status=False):
        """cost[N] of peripheral candidates Xp[N, n + 2] = [lambda, vmax, w(n)];
        gene[N] (or one int) says which gene each row belongs to."""
        gene, Xp = self._peripheral_args(gene, Xp)
        N = Xp.shape[0]
        cost = np.empty(N)
        st = np.empty(N, dtype=np.int32)
        self._check(self._lib.lassim_eval_peripheral(
            self._h, N, gene.ctypes.data_as(c_int_p), _dp(Xp), _dp(cost),
            st.ctypes.data_as(c_int_p)), "lassim_eval_peripheral")
        return (cost, st) if return_status else cost

    def peripheral_trajectories(self, gene, Xp):
        gene, Xp = self._peripheral_args(gene, Xp)
        N = Xp.shape[0]
        Y = np.empty((N, self.T, self.n + 1))
        self._check(self._lib.lassim_peripheral_trajectories(
            self._h, N, gene.ctypes.data_as(c_int_p), _dp(Xp), _dp(Y)),
            "lassim_peripheral_trajectories")
        return Y

    # -- differential evolution on the device ------------------------------
    def de_init(self, X0, lb, ub, variant=2, f=0.8, cr=0.9, seed=0):
        """X0[islands, individuals, dim]: initial populations of this GPU's islands."""
        X0 = _f64(X0)
        if X0.ndim != 3 or X0.shape[2] != self.dim:
            raise ValueError("populations of shape %s, expected (islands, individuals, %d)"
                             % (X0.shape, self.dim))
        lb = _f64(lb).reshape(-1); ub = _f64(ub).reshape(-1)
        if lb.size != self.dim or ub.size != self.dim:
            raise ValueError("bounds of length %d/%d for dimension %d" % (lb.size, ub.size, self.dim))
        self._check(self._lib.lassim_de_init(self._h, X0.shape[0], X0.shape[1], int(variant),
                                             float(f), float(cr), int(seed) & (2 ** 64 - 1),
                                             _dp(lb), _dp(ub), _dp(X0)), "lassim_de_init")
        self._de_shape = X0.shape

    def de_set_individual(self, island, index, x):
        x = _f64(x).reshape(-1)
        if x.size != self.dim:
            raise ValueError("individual of length %d for dimension %d" % (x.size, self.dim))
        self._check(self._lib.lassim_de_set_individual(self._h, int(island), int(index), _dp(x)),
                    "lassim_de_set_individual")

    def de_set_island_ids(self, first_island=0, stride=1):
        """Position of this GPU's islands in the whole archipelago (keys the random streams)."""
        self._check(self._lib.lassim_de_set_island_ids(self._h, int(first_island), int(stride)),
                    "lassim_de_set_island_ids")

    def de_set_lanes(self, lanes=0):
        """Groups of islands that evolve concurrently on streams of their own (0 = automatic,
        1 = all islands in lockstep); does not change results."""
        self._check(self._lib.lassim_de_set_lanes(self._h, int(lanes)), "lassim_de_set_lanes")

    def de_evolve(self, generations):
        self._check(self._lib.lassim_de_evolve(self._h, int(generations)), "lassim_de_evolve")

    def de_best(self):
        I, _, D = self._de_shape
        table = np.empty((I, D + 1))
        self._check(self._lib.lassim_de_best(self._h, _dp(table)), "lassim_de_best")
        return table

    def de_migrate_ring(self, table, first_island=0, stride=1):
        table = _f64(table)
        self._check(self._lib.lassim_de_migrate_ring(self._h, _dp(table), table.shape[0],
                                                     int(first_island), int(stride)),
                    "lassim_de_migrate_ring")

    def de_get(self):
        I, NP, D = self._de_shape
        X = np.empty((I, NP, D)); F = np.empty((I, NP))
        self._check(self._lib.lassim_de_get(self._h, _dp(X), _dp(F)), "lassim_de_get")
        return X, F

    @property
    def de_evaluations(self):
        return int(self._lib.lassim_de_evaluations(self._h))

    def last_end_states(self, N):
        K = self.trajectories_per_candidate
        E = np.empty((N, K, self.n))
        self._check(self._lib.lassim_last_end_states(self._h, N, _dp(E)),
                    "lassim_last_end_states")
        return E

    def last_traj_steps(self, N):
        K = self.trajectories_per_candidate
        s = np.empty((N, K), dtype=np.uint32)
        self._check(self._lib.lassim_last_traj_steps(self._h, N, s.ctypes.data_as(c_uint_p)),
                    "lassim_last_traj_steps")
        return s

    @property
    def flops_per_attempt(self):
        return int(self._lib.lassim_flops_per_attempt(self._h))

    @property
    def kernel_name(self):
        return self._lib.lassim_kernel_name(self._h).decode()

    def series_flops_per_attempt(self, N):
        """Flops per attempt of the kernel that integrates the time series of a batch of N."""
        return int(self._lib.lassim_series_flops_per_attempt(self._h, int(N)))

    def last_stage_parallel_attempts(self):
        """Attempts the stage-parallel kernel made in the last evaluation (sum)."""
        v = ctypes.c_ulonglong()
        self._check(self._lib.lassim_last_stage_parallel_attempts(self._h, ctypes.byref(v)),
                    "lassim_last_stage_parallel_attempts")
        return int(v.value)

    def last_stage_parallel_ms(self):
        """Device time of the stage-parallel kernel of the last evaluation (ms)."""
        ms = ctypes.c_float()
        self._check(self._lib.lassim_last_stage_parallel_ms(self._h, ctypes.byref(ms)),
                    "lassim_last_stage_parallel_ms")
        return float(ms.value)

    def last_capped_phase_ms(self):
        """Device time from the start of the last evaluation to the launch of its
        stage-parallel kernel (ms): the first attempts of every time series in the standard
        kernel."""
        ms = ctypes.c_float()
        self._check(self._lib.lassim_last_capped_phase_ms(self._h, ctypes.byref(ms)),
                    "lassim_last_capped_phase_ms")
        return float(ms.value)

    def series_kernel_name(self, N):
        return self._lib.lassim_series_kernel_name(self._h, int(N)).decode()

    @property
    def stream(self):
        """Address of the engine's own cudaStream_t (e.g. for torch.cuda.ExternalStream)."""
        return int(self._lib.lassim_stream(self._h))

    @property
    def launch_count(self):
        return int(self._lib.lassim_launch_count(self._h))

    def last_kernel_ms(self):
        ms = ctypes.c_float()
        self._check(self._lib.lassim_last_kernel_ms(self._h, ctypes.byref(ms)),
                    "lassim_last_kernel_ms")
        return float(ms.value)

    def fp64_peak_tflops(self, iters=200000):
        v = ctypes.c_double()
        self._check(self._lib.lassim_fp64_peak_probe(self._h, int(iters), ctypes.byref(v)),
                    "lassim_fp64_peak_probe")
        return float(v.value)
