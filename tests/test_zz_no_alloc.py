"""mg_residual_norm_noalloc (include/mgengine.h): the allocation-free variant of the residual norm, whose
partial sums live in a static device array, against the existing entry point mg_residual_norm, through the
C ABI, two calls in a row.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from evostencils_b200 import engine, lowering, problems

pytestmark = pytest.mark.gpu

TOL = 1e-12            # relative: the norm tolerance of tests/test_gpu_parity.py


def _setup(problem, level, seed):
    ws = engine.Workspace(problem.dimension, level - 1, level, 1, "f64", 0)
    table = lowering.operator_table(problem.system_operator(level, 0))
    for lv in (level - 1, level):
        ws.set_operator(lv, lowering.operator_table(problem.system_operator(lv, 0)))
    rng = np.random.default_rng(seed)
    shape = (2 ** level + 1,) * problem.dimension
    ws.upload("sol", level, 0, rng.standard_normal(shape))
    ws.upload("rhs", level, 0, rng.standard_normal(shape))
    return ws, table[0][0]


def _pointer(ws, kind, level):
    ptr, n, pitch, nz, z0 = C.c_void_p(), C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
    ws._check(ws.lib.mg_field_pointer(ws._h, engine._FIELD[kind], level, 0, C.byref(ptr), C.byref(n), C.byref(pitch),
                                      C.byref(nz), C.byref(z0)))
    return ptr.value, n.value, pitch.value


@pytest.mark.parametrize("name,level", [("poisson2d", 6), ("poisson2d", 9), ("poisson3d", 5), ("poisson2d", 4)])
def test_no_alloc_norm_matches_the_workspace_norm_twice_in_a_row(name, level):
    """Levels in this order on purpose: the big 2D grid leaves 511 row sums in the shared static array, the
    smaller grids after it must not see them."""
    import torch
    problem = problems.POISSON_2D if name == "poisson2d" else problems.POISSON_3D
    ws, entries = _setup(problem, level, seed=level)
    out = torch.full((2,), -1.0, dtype=torch.float64, device="cuda:0")
    torch.cuda.set_device(0)
    for call in range(2):
        want = ws.residual_norm(level)
        u_ptr, n, pitch = _pointer(ws, "sol", level)
        f_ptr, _, _ = _pointer(ws, "rhs", level)
        assert n == 2 ** level + 1
        engine.residual_norm_noalloc(problem.dimension, n, pitch, entries, u_ptr, f_ptr, out.data_ptr() + 8 * call)
        torch.cuda.synchronize()
        got = float(out[call].item())
        assert want > 0 and abs(got - want) <= TOL * want, (call, got, want)
        # new data for the second call: same entry point, same scratch
        rng = np.random.default_rng(100 + level + call)
        ws.upload("sol", level, 0, rng.standard_normal((n,) * problem.dimension))
    assert float(out[0].item()) != float(out[1].item())
    ws.close()


def test_no_alloc_norm_rejects_grids_beyond_the_static_scratch():
    """3D n = 1025: 1023^2 interior rows > MG_NOALLOC_MAX_ROWS = 266240 -- refused before anything is launched."""
    with pytest.raises(engine.EngineError, match="too large for the static scratch"):
        engine.residual_norm_noalloc(3, 1025, 1040, [((0, 0, 0), 1.0)], 4096, 4096, 4096)
    with pytest.raises(engine.EngineError, match="null device pointer"):
        engine.residual_norm_noalloc(2, 65, 80, [((0, 0), 1.0)], 0, 4096, 4096)
