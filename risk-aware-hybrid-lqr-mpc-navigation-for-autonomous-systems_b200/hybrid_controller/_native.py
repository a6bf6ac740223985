"""ctypes binding of libb2c.so (include/b2c.h) -- the only route from the host classes to the GPU.

There is no CPU fallback: if the library or a CUDA device is missing, the first use raises
``B2CUnavailable`` (importing this module stays cheap and never touches CUDA, so the drop-in
package can be imported on a GPU-less box, e.g. for `pytest -m "not gpu"`).
"""
import ctypes as C
import os
import threading

import numpy as np

# B2C_LIB: developer override (instrumented builds under tools/); always a CUDA build of the same ABI
LIB_PATH = os.environ.get("B2C_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb2c.so")

LIN_JACOBIAN, LIN_EULER, LIN_EXACT = 0, 1, 2
GUARD_NONE, GUARD_LQR, GUARD_MPC = 0, 1, 2
ST_OK, ST_MAXITER, ST_NONFINITE, ST_INFEASIBLE = 0, 1, 2, 3
MPC_LTV, MPC_LTI = 0, 1


class B2CUnavailable(RuntimeError):
    """libb2c.so cannot be used (not built, or no CUDA device)."""


class B2CError(RuntimeError):
    """An entry point returned an error code."""


class MpcParams(C.Structure):
    _fields_ = [("kind", C.c_int32), ("N", C.c_int32), ("M", C.c_int32), ("block_size", C.c_int32),
                ("soft", C.c_int32), ("max_iter", C.c_int32), ("ramp_up_steps", C.c_int32),
                ("engine", C.c_int32),
                ("Qd", C.c_double * 3), ("Rd", C.c_double * 2), ("Pd", C.c_double * 3),
                ("d_safe", C.c_double), ("rho", C.c_double), ("v_max", C.c_double),
                ("w_max", C.c_double), ("dt", C.c_double), ("tol", C.c_double)]


class RolloutParams(C.Structure):
    _fields_ = [("mode", C.c_int32), ("ticks", C.c_int32), ("mpc_rate", C.c_int32), ("min_dwell", C.c_int32),
                ("table_rows", C.c_int32), ("reserved", C.c_int32), ("dt", C.c_double),
                ("plant_v_max", C.c_double), ("plant_w_max", C.c_double),
                ("lqr_Qd", C.c_double * 3), ("lqr_Rd", C.c_double * 2), ("lqr_v_max", C.c_double),
                ("lqr_w_max", C.c_double),
                ("risk_d_safe", C.c_double), ("risk_d_trigger", C.c_double), ("risk_alpha", C.c_double),
                ("risk_beta", C.c_double), ("risk_threshold_low", C.c_double),
                ("mpc", MpcParams)]


ROLL_LQR, ROLL_MPC, ROLL_HYBRID = 0, 1, 2

_P = C.c_void_p
_D3, _D2 = C.c_double * 3, C.c_double * 2
# name -> (restype, argtypes); mirrors include/b2c.h one to one
SIGNATURES = {
    "b2c_create": (C.c_int, [C.POINTER(_P), C.c_int]),
    "b2c_create_on_stream": (C.c_int, [C.POINTER(_P), C.c_int, _P]),
    "b2c_destroy": (None, [_P]),
    "b2c_last_error": (C.c_char_p, []),
    "b2c_version": (C.c_int, []),
    "b2c_current_device": (C.c_int, [C.POINTER(C.c_int)]),
    "b2c_sync": (C.c_int, [_P]),
    "b2c_stream": (_P, [_P]),
    "b2c_launch_count": (C.c_int64, [_P]),
    "b2c_host_alloc": (C.c_int, [C.POINTER(_P), C.c_int64]),
    "b2c_host_free": (C.c_int, [_P]),
    "b2c_linearize_dev": (C.c_int, [_P, C.c_int64, C.c_int, C.c_int, C.c_double, _P, _P, _P, _P]),
    "b2c_linearize_packed_dev": (C.c_int, [_P, C.c_int64, C.c_int, C.c_int, C.c_double, _P, _P, _P, _P]),
    "b2c_linearize_host": (C.c_int, [_P, C.c_int64, C.c_int, C.c_int, C.c_double, _P, _P, _P, _P]),
    "b2c_dare_dev": (C.c_int, [_P, C.c_int64, C.c_int, C.c_int, C.c_double, _D3, _D2, _P, _P, _P, _P, _P, _P]),
    "b2c_dare_host": (C.c_int, [_P, C.c_int64, C.c_int, C.c_int, C.c_double, _D3, _D2, _P, _P, _P, _P, _P, _P]),
    "b2c_lqr_control_dev": (C.c_int, [_P, C.c_int64, C.c_double, C.c_double, _P, _P, _P, _P, _P, _P]),
    "b2c_lqr_control_host": (C.c_int, [_P, C.c_int64, C.c_double, C.c_double, _P, _P, _P, _P, _P, _P]),
    "b2c_mpc_scratch_bytes": (C.c_int64, [C.POINTER(MpcParams), C.c_int64]),
    "b2c_mpc_dev": (C.c_int, [_P, C.c_int64, C.POINTER(MpcParams)] + [_P] * 12),
    "b2c_mpc_host": (C.c_int, [_P, C.c_int64, C.POINTER(MpcParams)] + [_P] * 12),
    "b2c_mpc_host_async": (C.c_int, [_P, C.c_int64, C.POINTER(MpcParams)] + [_P] * 12),
    "b2c_wait": (C.c_int, [_P]),
    # table form: (ctx, B, params, table, rows, k0, x0, obs, step_count, u0, cost, status, iters, slack_used, useq, xpred)
    "b2c_mpc_table_dev": (C.c_int, [_P, C.c_int64, C.POINTER(MpcParams), _P, C.c_int32] + [_P] * 11),
    "b2c_mpc_table_host": (C.c_int, [_P, C.c_int64, C.POINTER(MpcParams), _P, C.c_int32] + [_P] * 11),
    "b2c_mpc_table_host_async": (C.c_int, [_P, C.c_int64, C.POINTER(MpcParams), _P, C.c_int32] + [_P] * 11),
    "b2c_mpc_dev_static": (C.c_int, [_P, C.c_int64, C.POINTER(MpcParams)] + [_P] * 12),
    "b2c_mpc_static_max_instances": (C.c_int64, [C.POINTER(MpcParams)]),
    "b2c_rollout_dev": (C.c_int, [_P, C.c_int64, C.POINTER(RolloutParams)] + [_P] * 11),
    "b2c_rollout_host": (C.c_int, [_P, C.c_int64, C.POINTER(RolloutParams)] + [_P] * 11),
    "b2c_rollout_generate_dev": (C.c_int, [_P, C.c_int64, C.c_int64, C.c_uint64, _P, C.c_int32, C.c_int32, _P, C.c_int32,
                                           C.c_double, C.c_double, C.c_double, _P, _P, _P]),
    "b2c_measure_fma_peak": (C.c_int, [_P, C.c_int, C.POINTER(C.c_double)]),
}

_lib = None
_lock = threading.Lock()


def load_library():
    """dlopen libb2c.so and attach prototypes (no CUDA call is made)."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise B2CUnavailable(
                    f"{LIB_PATH} is not built; run `python -c 'import __graft_entry__ as g; g.build()'` "
                    "(there is no CPU fallback)")
            lib = C.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)
                fn.restype, fn.argtypes = res, args
            _lib = lib
    return _lib


def _ptr(a):
    """host numpy array / torch tensor / int / None -> void*"""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    return int(a)


class Context:
    """Owns one b2c_ctx (stream + scratch) on one device."""

    def __init__(self, device=-1, stream=None):
        self.lib = load_library()
        h = _P()
        # stream=None: the context creates its own non-blocking stream; an int (0 = default stream,
        # or e.g. torch.cuda.current_stream().cuda_stream) is used as given
        rc = (self.lib.b2c_create_on_stream(C.byref(h), int(device), _P(int(stream))) if stream is not None
              else self.lib.b2c_create(C.byref(h), int(device)))
        if rc != 0:
            raise B2CUnavailable(self.lib.b2c_last_error().decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.b2c_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def call(self, name, *args):
        rc = getattr(self.lib, name)(self.h, *args)
        if rc != 0:
            raise B2CError(f"{name} failed ({rc}): {self.lib.b2c_last_error().decode()}")

    def sync(self):
        self.call("b2c_sync")

    @property
    def stream_handle(self):
        """the cudaStream_t the context launches on, as an integer"""
        self.lib.b2c_stream.restype = _P
        return int(self.lib.b2c_stream(self.h) or 0)

    @property
    def launches(self):
        return int(self.lib.b2c_launch_count(self.h))

    def fma_peak(self, bits=64):
        out = C.c_double()
        self.call("b2c_measure_fma_peak", int(bits), C.byref(out))
        return out.value


_default_ctx = {}


def default_context(device=-1):
    """Process-wide context per device for the scalar (batch = 1) drop-in calls.  device < 0 means the calling
    thread's current CUDA device, resolved at every call: the cache is keyed by the real device id."""
    if device < 0:
        lib = load_library()
        cur = C.c_int(-1)
        device = cur.value if lib.b2c_current_device(C.byref(cur)) == 0 else -1
    ctx = _default_ctx.get(device)
    if ctx is None:
        ctx = _default_ctx[device] = Context(device)
    return ctx


def mpc_params(kind, N, M, block_size, Qd, Rd, Pd, d_safe, rho, v_max, w_max, dt, soft=True,
               max_iter=0, ramp_up_steps=10, tol=0.0, engine=0):
    p = MpcParams()
    p.engine = engine
    p.kind, p.N, p.M, p.block_size = kind, N, M, block_size
    p.soft, p.max_iter, p.ramp_up_steps = int(bool(soft)), max_iter, ramp_up_steps
    p.Qd[:] = [float(x) for x in Qd]
    p.Rd[:] = [float(x) for x in Rd]
    p.Pd[:] = [float(x) for x in Pd]
    p.d_safe, p.rho, p.v_max, p.w_max, p.dt, p.tol = d_safe, rho, v_max, w_max, dt, tol
    return p


def _f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape is not None:
        a = a.reshape(shape)
    return a


# ---- host-buffer wrappers (numpy in / numpy out) used by the drop-in classes -------------------
def linearize_host(ctx, v, th, dt, method=LIN_EULER, guard=GUARD_NONE):
    v, th = _f64(v).ravel(), _f64(th).ravel()
    B = v.size
    Ad, Bd = np.empty((B, 3, 3)), np.empty((B, 3, 2))
    ctx.call("b2c_linearize_host", B, method, guard, float(dt), _ptr(v), _ptr(th), _ptr(Ad), _ptr(Bd))
    return Ad, Bd


def dare_host(ctx, v, th, dt, Qd, Rd, method=LIN_EULER, guard=GUARD_LQR, want_P=True):
    v, th = _f64(v).ravel(), _f64(th).ravel()
    B = v.size
    K = np.empty((B, 2, 3))
    P = np.empty((B, 3, 3)) if want_P else None
    it, st = np.empty(B, np.int32), np.empty(B, np.int32)
    ctx.call("b2c_dare_host", B, method, guard, float(dt), _D3(*map(float, Qd)), _D2(*map(float, Rd)),
             _ptr(v), _ptr(th), _ptr(K), _ptr(P), _ptr(it), _ptr(st))
    return K, P, it, st


def mpc_host(ctx, params, x0, xref, uref, obs, step_count=None, want_seq=True):
    """x0 (B,3), xref (B,N+1,3), uref (B,N,2), obs (B,M,3) -> dict of numpy outputs."""
    N, M = params.N, params.M
    x0 = _f64(x0).reshape(-1, 3)
    B = x0.shape[0]
    xref, uref = _f64(xref, (B, N + 1, 3)), _f64(uref, (B, N, 2))
    obs = _f64(obs, (B, M, 3)) if M else np.zeros((B, 0, 3))
    if step_count is not None:
        step_count = np.ascontiguousarray(step_count, dtype=np.int32).reshape(B).copy()
    out = dict(u0=np.empty((B, 2)), cost=np.empty(B), status=np.empty(B, np.int32),
               iters=np.empty(B, np.int32), slack_used=np.empty(B, np.int32),
               useq=np.empty((B, N, 2)) if want_seq else None,
               xpred=np.empty((B, N + 1, 3)) if want_seq else None, step_count=step_count)
    ctx.call("b2c_mpc_host", B, C.byref(params), _ptr(x0), _ptr(xref), _ptr(uref), _ptr(obs),
             _ptr(step_count), _ptr(out["u0"]), _ptr(out["cost"]), _ptr(out["status"]),
             _ptr(out["iters"]), _ptr(out["slack_used"]), _ptr(out["useq"]), _ptr(out["xpred"]))
    return out


def mpc_table_host(ctx, params, table, k0, x0, obs, step_count=None, want_seq=True):
    """Table form of solve_with_ltv (b2c_mpc_table_host): table (rows,6) = ReferenceTrajectoryGenerator.generate(),
    k0 (B,) start rows, x0 (B,3), obs (B,M,3).  Instance i solves on get_trajectory_segment(k0[i], N+1) of the table;
    results are bit-identical to mpc_host on the materialised segments."""
    N, M = params.N, params.M
    table = np.ascontiguousarray(table, dtype=np.float64).reshape(-1, 6)
    x0 = _f64(x0).reshape(-1, 3)
    B = x0.shape[0]
    k0 = np.ascontiguousarray(np.broadcast_to(np.asarray(k0, np.int32), (B,)))
    obs = _f64(obs, (B, M, 3)) if M else np.zeros((B, 0, 3))
    if step_count is not None:
        step_count = np.ascontiguousarray(step_count, dtype=np.int32).reshape(B).copy()
    out = dict(u0=np.empty((B, 2)), cost=np.empty(B), status=np.empty(B, np.int32),
               iters=np.empty(B, np.int32), slack_used=np.empty(B, np.int32),
               useq=np.empty((B, N, 2)) if want_seq else None,
               xpred=np.empty((B, N + 1, 3)) if want_seq else None, step_count=step_count)
    ctx.call("b2c_mpc_table_host", B, C.byref(params), _ptr(table), int(table.shape[0]), _ptr(k0), _ptr(x0), _ptr(obs),
             _ptr(step_count), _ptr(out["u0"]), _ptr(out["cost"]), _ptr(out["status"]),
             _ptr(out["iters"]), _ptr(out["slack_used"]), _ptr(out["useq"]), _ptr(out["xpred"]))
    return out
