"""b2c_mpc_dev_static: the allocation-free variant of b2c_mpc_dev (include/b2c.h; no counterpart in the reference).
Called through the C ABI and compared with the existing entry point on two
calls in a row (the second with other inputs and another shape, so stale contents of the shared static scratch would
show), plus the size limit.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _soa(a, B, dev):
    import torch
    return torch.from_numpy(np.ascontiguousarray(np.asarray(a, float).reshape(B, -1).T)).to(dev)


def _run(ctx, name, prm, B, N, ins, dev, sc0):
    import torch
    u0 = torch.full((2, B), np.nan, dtype=torch.float64, device=dev)
    cost = torch.full((B,), np.nan, dtype=torch.float64, device=dev)
    st, it, sl = (torch.full((B,), -7, dtype=torch.int32, device=dev) for _ in range(3))
    useq = torch.full((2 * N, B), np.nan, dtype=torch.float64, device=dev)
    xpred = torch.full((3 * (N + 1), B), np.nan, dtype=torch.float64, device=dev)
    sc = torch.from_numpy(sc0.copy()).to(dev)
    torch.cuda.synchronize()
    ctx.call(name, B, C.byref(prm), *(t.data_ptr() for t in ins), sc.data_ptr(), u0.data_ptr(), cost.data_ptr(),
             st.data_ptr(), it.data_ptr(), sl.data_ptr(), useq.data_ptr(), xpred.data_ptr())
    ctx.sync()
    return {k: v.cpu().numpy() for k, v in dict(u0=u0, cost=cost, status=st, iters=it, slack=sl, useq=useq,
                                                xpred=xpred, sc=sc).items()}


def test_mpc_dev_static_matches_mpc_dev_on_two_calls_in_a_row():
    import torch
    from hybrid_controller import _native as nat, workloads as wl
    dev = torch.device("cuda", 0)
    ctx = nat.Context(0)
    lib = nat.load_library()
    # (N, M, block, soft, B): config-4 shape with soft rows, then the run_simulation shape with move blocking and hard rows
    for call, (N, M, bs, soft, B) in enumerate(((20, 5, 1, True, 1500), (6, 3, 2, False, 900))):
        x0, xr, ur, ob = wl.config4_instances(B, N=N, seed=31 + call, obstacles=wl.DENSE_OBSTACLES[:M])
        x0[B // 3, 0] = np.nan                                            # one fallback instance
        ins = [_soa(a, B, dev) for a in (x0, xr, ur, ob)]
        sc0 = (np.arange(B) % 13).astype(np.int32)
        prm = nat.mpc_params(nat.MPC_LTV, N, M, bs, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02,
                             soft=soft, engine=1)
        assert lib.b2c_mpc_static_max_instances(C.byref(prm)) >= B
        ref = _run(ctx, "b2c_mpc_dev", prm, B, N, ins, dev, sc0)          # existing entry point, same (general) kernel
        # warm call (loads the kernel's module, fills torch's caching allocator for the output tensors); then the free
        # device memory the DRIVER reports must not drop across a call: the scratch b2c_mpc_dev would have had to
        # cudaMalloc for this batch is 5.8 MB / 1.0 MB
        _run(ctx, "b2c_mpc_dev_static", prm, B, N, ins, dev, sc0)
        free0 = torch.cuda.mem_get_info(dev)[0]
        out = _run(ctx, "b2c_mpc_dev_static", prm, B, N, ins, dev, sc0)
        assert free0 - torch.cuda.mem_get_info(dev)[0] < (1 << 19)
        assert ref["status"][B // 3] != 0 and (ref["status"] == 0).sum() > B // 2
        for k in ref:                                                    # same kernel, same arithmetic: every bit
            assert np.array_equal(out[k], ref[k], equal_nan=True), (call, k)
        if soft and bs == 1:                                             # and the default engine within the tests' tolerance
            prm0 = nat.mpc_params(nat.MPC_LTV, N, M, bs, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02)
            fast = _run(ctx, "b2c_mpc_dev", prm0, B, N, ins, dev, sc0)
            ok = ref["status"] == 0
            assert np.array_equal(fast["status"], out["status"])
            assert np.abs(fast["u0"][:, ok] - out["u0"][:, ok]).max() <= 1e-9
            assert np.array_equal(fast["u0"][:, ~ok], out["u0"][:, ~ok], equal_nan=True)     # the fallback law's outputs
            assert (np.abs(fast["cost"][ok] - out["cost"][ok]) <= 1e-10 * np.abs(out["cost"][ok])).all()
    # the limit: one instance too many is refused with a clear message, nothing is launched
    prm = nat.mpc_params(nat.MPC_LTV, 20, 5, 1, [15, 15, 50], [.1, .1], [30, 30, 40], .3, 5000., 2., 3., .02)
    too_many = lib.b2c_mpc_static_max_instances(C.byref(prm)) + 1
    n0 = ctx.launches
    with pytest.raises(nat.B2CError, match="exceed the static array"):
        ctx.call("b2c_mpc_dev_static", too_many, C.byref(prm), *([None] * 12))
    assert ctx.launches == n0
    ctx.close()
