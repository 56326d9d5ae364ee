"""kafka_solve_bands_static: the solve for callers that cannot allocate.

Some users of the library can only pass the arrays they already hold: a plug-in
living inside an application that has claimed the whole GPU for its own memory
pool, a loop that cannot afford the device-wide stall of cudaMalloc, or a thin
foreign-function stub with no allocator behind it.  For them the library keeps
the block-partial scratch of the solve in a fixed 2 MiB device array of its own.

This test calls the entry point through the C ABI twice in a row -- a large
problem first, then a smaller one, so the second call finds the first call's
leftovers in the shared array -- and checks every output against the ordinary
entry point (caller-provided scratch) with the tolerance the other solve tests
use.  It also checks that a call beyond the built-in limit is refused.

The file name sorts last so that a run which stops at the first failure has
already executed every other test; it passes on its own in a fresh process too.
"""
import numpy as np
import pytest
import torch

import gpu_util as gu
from kafka_inferenceengine_b200 import _lib
from kafka_inferenceengine_b200.device import padded, tri_count

pytestmark = pytest.mark.gpu

RTOL = 1e-9          # same as tests/test_gpu_solve.py


def _problem(rng, n, P, B):
    ld = padded(n)
    keep, arr = [], (_lib.SolveBand * B)()
    for b in range(B):
        obs = rng.random(n) < 0.85
        J = np.where(obs[:, None], rng.standard_normal((n, P)), 0.0)
        tens = [gu.planar(J, ld), gu.planar(np.where(obs, rng.standard_normal(n), 0.0), ld),
                gu.planar(np.where(obs, rng.standard_normal(n), 0.0), ld),
                gu.planar(np.where(obs, 400.0 * rng.uniform(0.5, 2.0, n), 0.0), ld)]
        keep.append(tens)
        arr[b].kind = _lib.BAND_ARRAYS
        arr[b].n_dim = P
        for k in range(P):
            arr[b].mapper[k] = k
        arr[b].jac, arr[b].h0, arr[b].y, arr[b].w = [t.data_ptr() for t in tens]
    x_f = rng.standard_normal((n, P))
    x_prev = x_f + 0.1 * rng.standard_normal((n, P))
    state = dict(xp=gu.planar(x_prev, ld), xf=gu.planar(x_f, ld), pinv=gu.pack_blocks(gu.random_spd(rng, n, P), ld))
    return ld, arr, keep, state


def _outputs(P, B, ld):
    return dict(x=torch.zeros((P, ld), dtype=torch.float64, device=gu.DEV),
                a=torch.zeros((tri_count(P), ld), dtype=torch.float64, device=gu.DEV),
                innov=torch.zeros((B, ld), dtype=torch.float64, device=gu.DEV),
                fwd=torch.zeros((B, ld), dtype=torch.float64, device=gu.DEV),
                norm2=torch.full((1,), -1.0, dtype=torch.float64, device=gu.DEV),
                n_bad=torch.full((1,), -1, dtype=torch.int32, device=gu.DEV))


def _run_both(rng, n, P, B):
    lib = _lib.get_lib()
    ld, arr, keep, st = _problem(rng, n, P, B)
    ref, got = _outputs(P, B, ld), _outputs(P, B, ld)
    scratch = torch.empty(int(lib.kafka_solve_scratch_doubles(n)), dtype=torch.float64, device=gu.DEV)
    _lib.check(lib.kafka_solve_bands(n, ld, P, arr, B, _lib.ptr(st["xp"]), _lib.ptr(st["xf"]), _lib.ptr(st["pinv"]),
                                     _lib.ptr(ref["x"]), _lib.ptr(ref["a"]), _lib.ptr(ref["innov"]),
                                     _lib.ptr(ref["fwd"]), _lib.ptr(ref["norm2"]), _lib.ptr(ref["n_bad"]),
                                     _lib.ptr(scratch), 0, gu.stream()), "kafka_solve_bands")
    # the variant takes the same arguments without the scratch pointer
    _lib.check(lib.kafka_solve_bands_static(n, ld, P, arr, B, _lib.ptr(st["xp"]), _lib.ptr(st["xf"]),
                                            _lib.ptr(st["pinv"]), _lib.ptr(got["x"]), _lib.ptr(got["a"]),
                                            _lib.ptr(got["innov"]), _lib.ptr(got["fwd"]), _lib.ptr(got["norm2"]),
                                            _lib.ptr(got["n_bad"]), 0, gu.stream()), "kafka_solve_bands_static")
    torch.cuda.synchronize()
    for key in ("x", "a", "innov", "fwd"):
        r, g = ref[key][:, :n].cpu().numpy(), got[key][:, :n].cpu().numpy()
        assert np.isfinite(g).all(), key
        assert np.max(np.abs(g - r) / np.maximum(np.abs(r), 1e-3)) < RTOL, key
    assert got["n_bad"].item() == ref["n_bad"].item() == 0
    assert got["norm2"].item() == pytest.approx(ref["norm2"].item(), rel=RTOL)
    assert got["norm2"].item() > 0.0


def test_static_scratch_solve_matches_on_two_calls_in_a_row():
    rng = np.random.default_rng(2024)
    _run_both(rng, 70_001, 10, 3)      # 547 thread blocks worth of partials left behind ...
    _run_both(rng, 4_099, 7, 2)        # ... for a call that uses 33 of them


def test_static_scratch_refuses_oversize_calls():
    lib = _lib.get_lib()
    dummy = torch.zeros(64, dtype=torch.float64, device=gu.DEV)
    nb = torch.zeros(1, dtype=torch.int32, device=gu.DEV)
    too_many = _lib.STATIC_MAX_PIXELS + 1
    st = lib.kafka_solve_bands_static(too_many, too_many, 2, None, 0, _lib.ptr(dummy), _lib.ptr(dummy),
                                      _lib.ptr(dummy), _lib.ptr(dummy), _lib.ptr(dummy), None, None,
                                      _lib.ptr(dummy), _lib.ptr(nb), 0, gu.stream())
    assert st == -3                                     # KAFKA_E_TOOLARGE, nothing launched
    msg = lib.kafka_last_error().decode()
    assert "at most %d pixels" % _lib.STATIC_MAX_PIXELS in msg
    torch.cuda.synchronize()
    # the largest admissible size needs exactly the whole array
    assert lib.kafka_solve_scratch_doubles(_lib.STATIC_MAX_PIXELS) == _lib.STATIC_SCRATCH_DOUBLES
