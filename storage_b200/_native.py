"""ctypes binding of liblsmc_b200.so (include/lsmc_b200.h) — the same entry points a C# host binds with P/Invoke.

There is NO CPU fallback: if the CUDA library is missing or no B200 is visible, every call raises.
"""
import ctypes as C
import os

import numpy as np

from . import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LSMC_B200_LIB") or os.path.join(_HERE, "liblsmc_b200.so")   # override: development builds only

EXPORTS = (
    "lsmc_version", "lsmc_create", "lsmc_destroy", "lsmc_last_error", "lsmc_comm_unique_id", "lsmc_comm_init", "lsmc_comm_init_local",
    "lsmc_calculate_multi", "lsmc_intrinsic", "lsmc_host_alloc", "lsmc_host_free",
    "lsmc_calculate", "lsmc_set_problem", "lsmc_get_layout", "lsmc_request_outputs", "lsmc_simulate", "lsmc_backward",
    "lsmc_forward", "lsmc_get_results", "lsmc_get_timings", "lsmc_set_kernel_timing", "lsmc_mt19937_normals", "lsmc_get_grid", "lsmc_get_inventory_space",
    "lsmc_get_regression_coeffs", "lsmc_get_sim_panel", "lsmc_debug_set_regression_coeffs", "lsmc_debug_keep_values",
    "lsmc_debug_get_values", "lsmc_npv_stats_noalloc",
)


class LsmcError(RuntimeError):
    """status -> the exception the C# shim re-throws (include/lsmc_b200.h lsmc_status)."""
    NAMES = {1: "ArgumentException", 2: "InventoryConstraintsCannotBeFulfilledException", 3: "OperationCanceledException",
             4: "ApplicationException", 5: "NotSupportedException"}

    def __init__(self, status, message):
        super().__init__("%s (lsmc status %d): %s" % (self.NAMES.get(status, "error"), status, message))
        self.status = status
        self.message = message


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("liblsmc_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                              "the LSMC path has no CPU fallback")
        _lib = C.CDLL(LIB_PATH)
        _lib.lsmc_last_error.restype = C.c_char_p
        _lib.lsmc_last_error.argtypes = [C.c_void_p]
        _lib.lsmc_version.restype = C.c_char_p
        _lib.lsmc_destroy.argtypes = [C.c_void_p]
        _lib.lsmc_destroy.restype = None
    return _lib


_dp = C.POINTER(C.c_double)


class Context:
    """One lsmc_ctx: one GPU, one stream, all device memory of a valuation."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        rc = lib().lsmc_create(C.c_int(device), C.byref(self._h))
        if rc != 0:
            raise LsmcError(rc, lib().lsmc_last_error(None).decode())
        self._keep = []

    def close(self):
        if self._h:
            lib().lsmc_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc != 0:
            raise LsmcError(rc, lib().lsmc_last_error(self._h).decode())

    # -- multi GPU ---------------------------------------------------------------------------------------------
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        rc = lib().lsmc_comm_unique_id(buf)
        if rc != 0:
            raise LsmcError(rc, lib().lsmc_last_error(None).decode())
        return buf.raw

    def comm_init(self, rank, world, unique_id: bytes):
        buf = C.create_string_buffer(unique_id, 128)
        self._check(lib().lsmc_comm_init(self._h, C.c_int(rank), C.c_int(world), buf))

    # -- one call ----------------------------------------------------------------------------------------------
    def calculate(self, problem, regression_rng, valuation_rng, results=None, panels=(), decision_index=False,
                  on_progress=None):
        if results is None:
            results = _abi.Results(problem.n_periods, problem.n_sim_periods, int(problem.n_sims), panels=panels,
                                   decision_index=decision_index)
        cp, cr, cv, cres = problem.c_struct(), regression_rng.c_struct(), valuation_rng.c_struct(), results.c_struct()
        cb = _make_cb(on_progress)
        self._problem = problem
        self._check(lib().lsmc_calculate(self._h, C.byref(cp), C.byref(cr), C.byref(cv), cb, None, C.byref(cres)))
        return results.sync_scalars()

    # -- staged ------------------------------------------------------------------------------------------------
    def request_outputs(self, panels=False, decision_index=False):
        self._check(lib().lsmc_request_outputs(self._h, C.c_uint32((1 if panels else 0) | (2 if decision_index else 0))))

    def set_problem(self, problem):
        cp = problem.c_struct()
        self._check(lib().lsmc_set_problem(self._h, C.byref(cp)))
        self._problem = problem

    def layout(self):
        lay = _abi.LsmcLayout()
        self._check(lib().lsmc_get_layout(self._h, C.byref(lay)))
        return lay

    def simulate(self, which, rng):
        cr = rng.c_struct()
        self._check(lib().lsmc_simulate(self._h, C.c_int(which), C.byref(cr)))

    def backward(self, on_progress=None):
        self._check(lib().lsmc_backward(self._h, _make_cb(on_progress), None))

    def forward(self, on_progress=None):
        self._check(lib().lsmc_forward(self._h, _make_cb(on_progress), None))

    def get_results(self, results=None, panels=(), decision_index=False):
        pb = self._problem
        if results is None:
            results = _abi.Results(pb.n_periods, pb.n_sim_periods, int(pb.n_sims), panels=panels, decision_index=decision_index)
        cres = results.c_struct()
        self._check(lib().lsmc_get_results(self._h, C.byref(cres)))
        return results.sync_scalars()

    def set_kernel_timing(self, on):
        """CUDA events around every sweep launch of the next backward passes (timings()['backward_sweep_kernel_ms'])."""
        self._check(lib().lsmc_set_kernel_timing(self._h, C.c_int(1 if on else 0)))

    def timings(self):
        t = _abi.LsmcTimings()
        self._check(lib().lsmc_get_timings(self._h, C.byref(t)))
        return {name: getattr(t, name) for name, _ in _abi.LsmcTimings._fields_}

    # -- introspection -----------------------------------------------------------------------------------------
    def grid(self, period):
        lay = self.layout()
        out = np.zeros(lay.max_grid)
        count = C.c_int32()
        self._check(lib().lsmc_get_grid(self._h, C.c_int(period), out.ctypes.data_as(_dp), C.byref(count)))
        return out[:count.value].copy()

    def inventory_space(self):
        P = self.layout().n_periods
        lo, hi = np.zeros(P), np.zeros(P)
        self._check(lib().lsmc_get_inventory_space(self._h, lo.ctypes.data_as(_dp), hi.ctypes.data_as(_dp)))
        return lo, hi

    def regression_coeffs(self, period):
        g = len(self.grid(period + 1))
        out = np.zeros((g, self._problem.n_basis))
        self._check(lib().lsmc_get_regression_coeffs(self._h, C.c_int(period), out.ctypes.data_as(_dp)))
        return out

    def sim_panel(self, which, factor=-1):
        lay = self.layout()
        out = np.zeros((lay.n_sim_periods, lay.n_sims))
        self._check(lib().lsmc_get_sim_panel(self._h, C.c_int(which), C.c_int(factor), out.ctypes.data_as(_dp)))
        return out

    # -- test hooks --------------------------------------------------------------------------------------------
    def debug_set_regression_coeffs(self, period, coeffs):
        a = np.ascontiguousarray(coeffs, dtype=np.float64)
        self._check(lib().lsmc_debug_set_regression_coeffs(self._h, C.c_int(period), a.ctypes.data_as(_dp)))

    def debug_keep_values(self, keep=True):
        self._check(lib().lsmc_debug_keep_values(self._h, C.c_int(int(keep))))

    def debug_values(self, period):
        g = len(self.grid(period))
        out = np.zeros((g, self.layout().n_sims))
        self._check(lib().lsmc_debug_get_values(self._h, C.c_int(period), out.ctypes.data_as(_dp)))
        return out


class _PinnedBlock:
    """Owner of one lsmc_host_alloc block; numpy arrays created over it keep it alive through `.base`."""

    def __init__(self, nbytes):
        self.ptr = C.c_void_p()
        rc = lib().lsmc_host_alloc(C.c_size_t(nbytes), C.byref(self.ptr))
        if rc != 0:
            raise LsmcError(rc, lib().lsmc_last_error(None).decode())
        self.nbytes = nbytes

    def __del__(self):
        try:
            if self.ptr:
                lib().lsmc_host_free(self.ptr)
                self.ptr = C.c_void_p()
        except Exception:
            pass


def pinned_zeros(shape, dtype=np.float64):
    """A zero-filled numpy array in page-locked host memory (lsmc_host_alloc): result panels allocated this way are written
    by the copy engine directly (include/lsmc_b200.h)."""
    dtype = np.dtype(dtype)
    count = int(np.prod(shape))
    block = _PinnedBlock(max(count * dtype.itemsize, 1))
    lib().lsmc_host_free.argtypes = [C.c_void_p]
    buf = (C.c_char * block.nbytes).from_address(block.ptr.value)
    arr = np.frombuffer(buf, dtype=dtype, count=count).reshape(shape)
    arr[...] = 0
    _PINNED_OWNERS[arr.ctypes.data] = block   # keeps the block alive for the process lifetime of the view
    return arr


_PINNED_OWNERS = {}


def release_pinned(arr):
    """Free the page-locked block behind an array made by pinned_zeros (the array must not be used afterwards)."""
    _PINNED_OWNERS.pop(arr.ctypes.data, None)


def intrinsic(problem):
    """lsmc_intrinsic: (npv, profile[P][6]) of the deterministic valuation on the forward curve (host only)."""
    cp = problem.c_struct()
    npv = C.c_double()
    profile = np.zeros((int(problem.n_periods), 6))
    rc = lib().lsmc_intrinsic(C.byref(cp), C.byref(npv), profile.ctypes.data_as(_dp))
    if rc != 0:
        raise LsmcError(rc, lib().lsmc_last_error(None).decode())
    return npv.value, profile


def mt19937_normals(seed, antithetic, n_sims, n_steps, n_factors):
    """lsmc_mt19937_normals: the standard normals of the reference's seeded stream, [n_steps][n_factors][n_sims] (host only)."""
    out = np.zeros((int(n_steps), int(n_factors), int(n_sims)))
    fn = lib().lsmc_mt19937_normals
    fn.argtypes = [C.c_uint64, C.c_int, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]
    fn.restype = C.c_int
    rc = fn(C.c_uint64(int(seed) & 0xFFFFFFFF), C.c_int(1 if antithetic else 0), C.c_int64(int(n_sims)), C.c_int32(int(n_steps)),
            C.c_int32(int(n_factors)), out.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise LsmcError(rc, "lsmc_mt19937_normals: invalid argument")
    return out


class Group:
    """n contexts of ONE process joined by lsmc_comm_init_local: one call values the whole problem on all of them
    (lsmc_calculate_multi).  `devices` may name the same GPU twice (tests on a one-GPU box)."""

    def __init__(self, devices):
        self.contexts = [Context(d) for d in devices]
        self._arr = (C.c_void_p * len(self.contexts))(*[c._h for c in self.contexts])
        rc = lib().lsmc_comm_init_local(self._arr, C.c_int(len(self.contexts)))
        if rc != 0:
            raise LsmcError(rc, lib().lsmc_last_error(self.contexts[0]._h).decode())

    def calculate(self, problem, regression_rng, valuation_rng, results=None, panels=(), decision_index=False, on_progress=None):
        if results is None:
            results = _abi.Results(problem.n_periods, problem.n_sim_periods, int(problem.n_sims), panels=panels,
                                   decision_index=decision_index)
        cp, cr, cv, cres = problem.c_struct(), regression_rng.c_struct(), valuation_rng.c_struct(), results.c_struct()
        cb = _make_cb(on_progress)
        rc = lib().lsmc_calculate_multi(self._arr, C.c_int(len(self.contexts)), C.byref(cp), C.byref(cr), C.byref(cv), cb, None, C.byref(cres))
        if rc != 0:
            raise LsmcError(rc, lib().lsmc_last_error(self.contexts[0]._h).decode())
        return results.sync_scalars()

    def close(self):
        for c in self.contexts:
            c.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def npv_stats_noalloc(d_pv_by_sim, n_sims, antithetic_pairs, d_out4, cuda_stream=0):
    """lsmc_npv_stats_noalloc (include/lsmc_b200.h): device pointers as integers; enqueues on `cuda_stream`, allocates nothing."""
    err = C.create_string_buffer(256)
    fn = lib().lsmc_npv_stats_noalloc
    fn.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
    fn.restype = C.c_int
    rc = fn(C.c_void_p(d_pv_by_sim), C.c_int64(n_sims), C.c_int(int(bool(antithetic_pairs))), C.c_void_p(d_out4),
            C.c_void_p(cuda_stream), err, C.c_size_t(len(err)))
    if rc != 0:
        raise LsmcError(rc, err.value.decode())


def comm_setup(ctx, rank, world, dist):
    """Join `ctx` to the communicator of a one-process-per-GPU job: rank 0 creates the NCCL unique id and broadcasts it
    through the caller's torch.distributed process group (`dist`; any backend, only Python objects travel)."""
    if world <= 1:
        return
    uid = [Context.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.comm_init(rank, world, uid[0])


def _make_cb(on_progress):
    if on_progress is None:
        return _abi.PROGRESS_CB(0)

    def trampoline(_user, value):
        ret = on_progress(value)
        return 1 if ret is False else 0
    return _abi.PROGRESS_CB(trampoline)
