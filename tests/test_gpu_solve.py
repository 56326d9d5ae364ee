"""GPU parity: the fused normal-equation build + Cholesky solve
(SURVEY section 8 rows a8/a9, config 5 "supplied Jacobians") against oracle B
(fp64 dense) and the reference's own output, through the C ABI."""
import numpy as np
import pytest
import torch

import golden_io as gio
import gpu_util as gu
import kafka_inferenceengine_b200 as kb
from kafka_inferenceengine_b200 import _lib
from kafka_inferenceengine_b200 import synthetic as syn
from kafka_inferenceengine_b200.device import padded, tri_count
from oracle import kafka_cpu as kc

pytestmark = pytest.mark.gpu

RTOL = 1e-9          # posterior mean / A against the fp64 oracle (north_star asks 1e-6)


def _problem(rng, n, P, B, frac_obs=0.85):
    Js = [rng.standard_normal((n, P)) for _ in range(B)]
    H0s = [rng.standard_normal(n) for _ in range(B)]
    ms = [rng.random(n) < frac_obs for _ in range(B)]
    ws = [np.where(m, 400.0 * rng.uniform(0.5, 2.0, n), 0.0) for m in ms]
    ys = [np.where(m, rng.standard_normal(n), 0.0) for m in ms]
    for b in range(B):                      # unobserved rows are empty (utils.py:199-215)
        Js[b][~ms[b]] = 0.0
        H0s[b][~ms[b]] = 0.0
    Pinv = gu.random_spd(rng, n, P)
    x_f = rng.standard_normal(n * P)
    x_prev = x_f + 0.1 * rng.standard_normal(n * P)
    return Js, H0s, ws, ys, Pinv, x_f, x_prev


def _call(n, P, bands, x_prev, x_f, pinv_t, flags=0, want_fwd=True, count_lu=False):
    """Straight ctypes call of kafka_solve_bands.  bands: list of dicts.  ``count_lu``: also return
    the number of pixels solved by pivoted elimination (KAFKA_SOLVE_COUNT_LU)."""
    lib = _lib.get_lib()
    ld = padded(n)
    B = len(bands)
    arr = (_lib.SolveBand * max(B, 1))()
    keep = []
    for b, bd in enumerate(bands):
        y = gu.planar(bd["y"], ld)
        w = gu.planar(bd["w"], ld)
        keep += [y, w]
        arr[b].y, arr[b].w = y.data_ptr(), w.data_ptr()
        if "wcm" in bd:
            arr[b].kind = _lib.BAND_WCM
            for k in range(6):
                arr[b].row[k] = bd["wcm"][k]
            arr[b].mapper[0], arr[b].mapper[1] = bd["mapper"]
        elif "uni_J" in bd:
            arr[b].kind = _lib.BAND_UNIFORM
            J = torch.from_numpy(np.ascontiguousarray(bd["uni_J"], dtype=np.float64)).to(gu.DEV)
            h0 = torch.tensor([bd["uni_H0"]], dtype=torch.float64, device=gu.DEV)
            keep += [J, h0]
            arr[b].jac, arr[b].h0 = J.data_ptr(), h0.data_ptr()
            arr[b].n_dim = len(bd["mapper"])
            for k, sl in enumerate(bd["mapper"]):
                arr[b].mapper[k] = int(sl)
        elif "row" in bd:
            arr[b].kind = _lib.BAND_LINEAR
            for p in range(P):
                arr[b].row[p] = bd["row"][p]
            arr[b].offset = bd["offset"]
        else:
            arr[b].kind = _lib.BAND_ARRAYS
            J = gu.planar(bd["J"], ld)
            h0 = gu.planar(bd["H0"], ld)
            keep += [J, h0]
            arr[b].jac, arr[b].h0 = J.data_ptr(), h0.data_ptr()
            mp = bd.get("mapper", np.arange(P))
            arr[b].n_dim = len(mp)
            for k, s in enumerate(mp):
                arr[b].mapper[k] = int(s)
    xp = gu.planar(x_prev.reshape(n, P), ld)
    xf = gu.planar(x_f.reshape(n, P), ld)
    x_out = torch.zeros((P, ld), dtype=torch.float64, device=gu.DEV)
    a_out = torch.zeros((tri_count(P), ld), dtype=torch.float64, device=gu.DEV)
    innov = torch.zeros((max(B, 1), ld), dtype=torch.float64, device=gu.DEV)
    fwd = torch.zeros((max(B, 1), ld), dtype=torch.float64, device=gu.DEV)
    norm2 = torch.zeros(1, dtype=torch.float64, device=gu.DEV)
    n_bad = torch.full((2,), -7, dtype=torch.int32, device=gu.DEV)
    scratch = torch.full((int(lib.kafka_solve_scratch_doubles(n)),), float("nan"), dtype=torch.float64, device=gu.DEV)
    if count_lu:
        flags |= _lib.SOLVE_COUNT_LU
    _lib.check(lib.kafka_solve_bands(n, ld, P, arr, B, _lib.ptr(xp), _lib.ptr(xf), _lib.ptr(pinv_t), _lib.ptr(x_out),
                                     _lib.ptr(a_out), _lib.ptr(innov), _lib.ptr(fwd) if want_fwd else None,
                                     _lib.ptr(norm2), _lib.ptr(n_bad), _lib.ptr(scratch), flags, gu.stream()),
               "kafka_solve_bands")
    torch.cuda.synchronize()
    counts = n_bad.cpu().numpy()
    assert count_lu or counts[1] == -7                   # without the flag the second counter is not the library's
    res = (gu.unplanar(x_out, n).ravel(), gu.unpack_blocks(a_out, n, P), innov[:B, :n].cpu().numpy().ravel(),
           fwd[:B, :n].cpu().numpy().ravel(), float(norm2.item()), int(counts[0]))
    return res + (int(counts[1]),) if count_lu else res


@pytest.mark.parametrize("pair", [0, _lib.SOLVE_PAIR])
@pytest.mark.parametrize("P", list(range(1, 17)))
def test_supplied_jacobian_all_state_sizes(P, pair):
    """Config 5 sweep axis n_params 1..16 (x bands 1..13); ``pair`` forces the two-lanes-per-pixel
    kernel (the default from 14 parameters) at every size."""
    rng = np.random.default_rng(P)
    B = 1 + (5 * P) % 13
    n = 1500 + 37 * P
    Js, H0s, ws, ys, Pinv, x_f, x_prev = _problem(rng, n, P, B)
    bands = [dict(J=Js[b], H0=H0s[b], y=ys[b], w=ws[b]) for b in range(B)]
    x, A, innov, fwd, norm2, n_bad = _call(n, P, bands, x_prev, x_f, gu.pack_blocks(Pinv),
                                           flags=pair)
    xo, Ao, innov_o, fwd_o = kc.solve_dense64(ys, ws, H0s, Js, x_prev, x_f, Pinv)
    assert n_bad == 0
    assert np.max(gu.rel(A, Ao, floor=1e-6)) < RTOL
    assert np.max(np.abs(x - xo) / np.maximum(np.abs(xo), 1e-3)) < RTOL
    assert np.array_equal(innov, innov_o)
    assert np.allclose(fwd, fwd_o, rtol=1e-9, atol=1e-10)
    assert norm2 == pytest.approx(float(((xo - x_prev) ** 2).sum()), rel=1e-9)


@pytest.mark.parametrize("P,flags", [(7, 0), (7, _lib.SOLVE_PAIR), (15, 0), (16, 0),
                                     (15, _lib.SOLVE_PAIR)])
def test_norm_is_deterministic_and_edge_sizes(P, flags):
    rng = np.random.default_rng(99)
    for n in (1, 31, 128, 129, 5000):
        B = 2
        Js, H0s, ws, ys, Pinv, x_f, x_prev = _problem(rng, n, P, B)
        bands = [dict(J=Js[b], H0=H0s[b], y=ys[b], w=ws[b]) for b in range(B)]
        r1 = _call(n, P, bands, x_prev, x_f, gu.pack_blocks(Pinv), flags=flags)
        r2 = _call(n, P, bands, x_prev, x_f, gu.pack_blocks(Pinv), flags=flags)
        assert r1[4] == r2[4] and np.array_equal(r1[0], r2[0]) and np.array_equal(r1[1], r2[1])
        xo, Ao, _, _ = kc.solve_dense64(ys, ws, H0s, Js, x_prev, x_f, Pinv)
        assert np.max(np.abs(r1[0] - xo) / np.maximum(np.abs(xo), 1e-3)) < RTOL


@pytest.mark.parametrize("P", [2, 7, 14])
def test_pair_kernel_on_uniform_linear_and_wcm_bands(P):
    """Every band kind through both kernels: a shared linearisation (KAFKA_BAND_UNIFORM), a constant
    row and (P=2) the Water-Cloud model; the two kernels agree to rounding and with the oracle."""
    rng = np.random.default_rng(40 + P)
    n = 1111
    m = [rng.random(n) < 0.8 for _ in range(3)]
    ws = [np.where(mm, 200.0, 0.0) for mm in m]
    ys = [np.where(mm, rng.standard_normal(n), 0.0) for mm in m]
    D = min(P, 4)
    mapper = rng.permutation(P)[:D]
    uJ, uH0 = rng.standard_normal(D), 0.37
    row, off = rng.standard_normal(P), -0.2
    Pinv = gu.random_spd(rng, n, P)
    x_f = 1.0 + 0.1 * rng.standard_normal(n * P)
    x_prev = x_f + 0.02 * rng.standard_normal(n * P)
    bands = [dict(uni_J=uJ, uni_H0=uH0, mapper=mapper, y=ys[0], w=ws[0]), dict(row=row, offset=off, y=ys[1], w=ws[1])]
    Jf = [np.zeros((n, P)), np.where(m[1][:, None], row[None, :], 0.0)]
    Jf[0][:, mapper] = uJ
    Jf[0][~m[0]] = 0.0
    H0f = [np.where(m[0], uH0, 0.0), np.where(m[1], x_prev.reshape(n, P) @ row + off, 0.0)]
    if P == 2:
        prm = list(kc.WCM_PARAMETERS["VV"]) + [np.cos(np.deg2rad(23.0))]
        bands.append(dict(wcm=prm, mapper=(0, 1), y=ys[2], w=ws[2]))
        H0w, Jw = kc.sar_operator(x_prev, 2, m[2], 0, round_f32=False)
        # sar_operator uses the VV constants of sar_forward_model.py:60; same numbers as prm
        Jf.append(Jw)
        H0f.append(H0w)
    nb = len(bands)
    res = [_call(n, P, bands, x_prev, x_f, gu.pack_blocks(Pinv), flags=f) for f in (0, _lib.SOLVE_PAIR)]
    assert np.allclose(res[0][0], res[1][0], rtol=1e-11, atol=1e-13)
    assert np.array_equal(res[0][1], res[1][1]) and np.array_equal(res[0][2], res[1][2])
    assert np.allclose(res[0][3], res[1][3], rtol=1e-11, atol=1e-12)
    xo, Ao, innov_o, fwd_o = kc.solve_dense64(ys[:nb], ws[:nb], H0f, Jf, x_prev, x_f, Pinv)
    for x, A, innov, fwd, norm2, n_bad in res:
        assert n_bad == 0
        assert np.max(gu.rel(A, Ao, floor=1e-6)) < 1e-8
        assert np.max(np.abs(x - xo) / np.maximum(np.abs(xo), 1e-3)) < 1e-8
        assert np.allclose(innov, innov_o, rtol=1e-9, atol=1e-10)


@pytest.mark.parametrize("P", [3, 7, 14, 16])
def test_pair_kernel_on_shared_linearisation_mapped_arrays_and_broadcast_prior(P):
    """The two-lanes-per-pixel kernel on a shared linearisation (KAFKA_BAND_UNIFORM) on some state slots,
    per-pixel arrays through a state mapper and per-pixel arrays on every slot, with per-pixel prior
    blocks and with one block shared by all pixels.  Same results as the one-thread kernel (same
    operations in the same order) and as the oracle."""
    rng = np.random.default_rng(70 + P)
    n = 1237
    m = [rng.random(n) < 0.8 for _ in range(3)]
    ws = [np.where(mm, 200.0, 0.0) for mm in m]
    ys = [np.where(mm, rng.standard_normal(n), 0.0) for mm in m]
    D = min(P, 4)
    map_u, map_a = rng.permutation(P)[:D], np.sort(rng.permutation(P)[:D])
    uJ, uH0 = rng.standard_normal(D), 0.37
    Ja = np.where(m[1][:, None], rng.standard_normal((n, D)), 0.0)
    H0a = np.where(m[1], rng.standard_normal(n), 0.0)
    Jb = np.where(m[2][:, None], rng.standard_normal((n, P)), 0.0)
    H0b = np.where(m[2], rng.standard_normal(n), 0.0)
    x_f = 1.0 + 0.1 * rng.standard_normal(n * P)
    x_prev = x_f + 0.02 * rng.standard_normal(n * P)
    bands = [dict(uni_J=uJ, uni_H0=uH0, mapper=map_u, y=ys[0], w=ws[0]),
             dict(J=Ja, H0=H0a, y=ys[1], w=ws[1], mapper=map_a), dict(J=Jb, H0=H0b, y=ys[2], w=ws[2])]
    Jf = [np.zeros((n, P)), np.zeros((n, P)), Jb]
    Jf[0][:, map_u] = uJ
    Jf[0][~m[0]] = 0.0
    Jf[1][:, map_a] = Ja
    H0f = [np.where(m[0], uH0, 0.0), H0a, H0b]
    Pinv = gu.random_spd(rng, n, P)
    one = gu.random_spd(rng, 1, P)[0]
    r, c = np.tril_indices(P)
    cases = [(gu.pack_blocks(Pinv), Pinv, 0),
             (torch.from_numpy(np.ascontiguousarray(one[r, c])).to(gu.DEV), np.repeat(one[None], n, axis=0), _lib.PRIOR_BROADCAST)]
    for packed, blocks, extra in cases:
        res = [_call(n, P, bands, x_prev, x_f, packed, flags=f | extra) for f in (0, _lib.SOLVE_PAIR)]
        assert np.allclose(res[0][0], res[1][0], rtol=1e-11, atol=1e-13)
        assert np.array_equal(res[0][1], res[1][1]) and np.array_equal(res[0][2], res[1][2])
        assert np.allclose(res[0][3], res[1][3], rtol=1e-11, atol=1e-12)
        xo, Ao, innov_o, fwd_o = kc.solve_dense64(ys, ws, H0f, Jf, x_prev, x_f, blocks)
        for x, A, innov, fwd, norm2, n_bad in res:
            assert n_bad == 0
            assert np.max(gu.rel(A, Ao, floor=1e-6)) < 1e-8
            assert np.max(np.abs(x - xo) / np.maximum(np.abs(xo), 1e-3)) < 1e-8
            assert np.allclose(innov, innov_o, rtol=1e-9, atol=1e-10)
            assert np.allclose(fwd, fwd_o, rtol=1e-9, atol=1e-10)
            assert norm2 == pytest.approx(float(((xo - x_prev) ** 2).sum()), rel=1e-9)


def test_no_bands_and_empty_image():
    rng = np.random.default_rng(1)
    n, P = 300, 5
    _, _, _, _, Pinv, x_f, x_prev = _problem(rng, n, P, 1)
    x, A, _, _, norm2, n_bad = _call(n, P, [], x_prev, x_f, gu.pack_blocks(Pinv))
    assert np.allclose(x, x_f, rtol=1e-12, atol=1e-13) and np.array_equal(A, Pinv) and n_bad == 0
    lib = _lib.get_lib()
    norm2 = torch.ones(1, dtype=torch.float64, device=gu.DEV)
    nb = torch.ones(1, dtype=torch.int32, device=gu.DEV)
    st = lib.kafka_solve_bands(0, 32, P, None, 0, None, None, None, None, None, None, None, _lib.ptr(norm2),
                               _lib.ptr(nb), None, 0, gu.stream())
    assert st == 0 and norm2.item() == 0.0 and nb.item() == 0


@pytest.mark.parametrize("pair", [0, _lib.SOLVE_PAIR])
def test_mapper_linear_and_broadcast_prior(pair):
    """TIP layout: band Jacobians live on 4 of 7 slots (utils.py:148-153); a linear
    band in tuple form; one prior block shared by all pixels (kafka_test_S2.py:106-118)."""
    rng = np.random.default_rng(5)
    n, P = 2111, 7
    mappers = [np.array([0, 1, 6, 2]), np.array([3, 4, 6, 5])]
    m = [rng.random(n) < 0.8 for _ in range(3)]
    J4 = [np.where(m[b][:, None], rng.standard_normal((n, 4)), 0.0) for b in range(2)]
    H0 = [np.where(m[b], rng.standard_normal(n), 0.0) for b in range(2)]
    ws = [np.where(mm, 300.0, 0.0) for mm in m]
    ys = [np.where(mm, rng.standard_normal(n), 0.0) for mm in m]
    row = rng.standard_normal(P)
    off = 0.3
    _, _, inv = kc.tip_prior()
    x_f = np.tile(kc.TIP_MEAN, n) + 0.01 * rng.standard_normal(n * P)
    x_prev = x_f + 0.05 * rng.standard_normal(n * P)
    bands = [dict(J=J4[b], H0=H0[b], y=ys[b], w=ws[b], mapper=mappers[b]) for b in range(2)]
    bands.append(dict(row=row, offset=off, y=ys[2], w=ws[2]))
    r, c = np.tril_indices(P)
    pinv_b = torch.from_numpy(np.ascontiguousarray(inv.astype(np.float64)[r, c])).to(gu.DEV)
    x, A, innov, fwd, norm2, n_bad = _call(n, P, bands, x_prev, x_f, pinv_b, flags=_lib.PRIOR_BROADCAST | pair)
    Jf = []
    for b in range(2):
        J = np.zeros((n, P))
        J[:, mappers[b]] = J4[b]
        Jf.append(J)
    Jf.append(np.where(m[2][:, None], row[None, :], 0.0))
    H0f = H0 + [np.where(m[2], x_prev.reshape(n, P) @ row + off, 0.0)]
    Pinv = np.repeat(inv[None].astype(np.float64), n, axis=0)
    xo, Ao, innov_o, fwd_o = kc.solve_dense64(ys, ws, H0f, Jf, x_prev, x_f, Pinv)
    assert n_bad == 0
    assert np.max(gu.rel(A, Ao, floor=1e-6)) < RTOL
    assert np.max(np.abs(x - xo) / np.maximum(np.abs(xo), 1e-3)) < RTOL
    assert np.allclose(innov, innov_o, rtol=1e-12, atol=1e-13)
    assert np.allclose(fwd, fwd_o, rtol=1e-9, atol=1e-10)


@pytest.mark.parametrize("pair", [0, _lib.SOLVE_PAIR])
@pytest.mark.parametrize("P", list(range(1, 17)))
def test_indefinite_blocks_get_the_lu_solution(P, pair):
    """A block that is not positive definite is solved as the reference's sparse LU solves it
    (solvers.py:133-134): x = A^-1 b by elimination with partial pivoting.  Every state size,
    both solve kernels; failing pixels at the first / last lane of a warp, at tile borders, in the
    last (ragged) tile; a singular block and a block holding a NaN stay at the forecast and are
    the only ones counted as such.  Forward-modelled output and the norm follow the new state."""
    rng = np.random.default_rng(100 + P)
    B = 1 + P % 3
    n = 1000 + 13 * P
    Js, H0s, ws, ys, Pinv, x_f, x_prev = _problem(rng, n, P, B)
    indef = np.array(sorted(set([0, 31, 32, 63, 64, 127, 128, 129, 500, n - 1]) | set(rng.integers(0, n, 25).tolist())))
    singular, with_nan = 300, 700
    indef = indef[(indef != singular) & (indef != with_nan)]
    for i in indef:
        # SPD prior block minus a large multiple of one direction: one negative eigenvalue that the
        # observations cannot lift, so the Cholesky factorisation meets a negative pivot
        v = rng.standard_normal(P)
        v /= np.linalg.norm(v)
        lift = sum(ws[b][i] * float(Js[b][i] @ Js[b][i]) for b in range(B))
        Pinv[i] = Pinv[i] - (3.0 * np.trace(Pinv[i]) + 3.0 * lift + 1.0) * np.outer(v, v)
    Pinv[singular] = 0.0
    for b in range(B):
        ws[b][singular] = 0.0
    Pinv[with_nan][0, 0] = np.nan
    bands = [dict(J=Js[b], H0=H0s[b], y=ys[b], w=ws[b]) for b in range(B)]
    x, A, innov, fwd, norm2, n_bad, n_lu = _call(n, P, bands, x_prev, x_f, gu.pack_blocks(Pinv), flags=pair,
                                                 count_lu=True)
    assert n_bad == 2 and n_lu == len(indef)
    Pinv_o = np.where(np.isnan(Pinv), 1.0, Pinv)
    Pinv_o[singular] = np.eye(P)                                   # numpy refuses a singular block outright
    xo, Ao, innov_o, fwd_o = kc.solve_dense64(ys, ws, H0s, Js, x_prev, x_f, Pinv_o)
    Ao[singular] = 0.0
    ev = np.linalg.eigvalsh(Ao[indef])
    assert (ev.min(axis=1) < 0).all() and (np.abs(ev).min(axis=1) / np.abs(ev).max(axis=1)).min() > 1e-7
    xr, xor_, xfr = x.reshape(n, P), xo.reshape(n, P).copy(), x_f.reshape(n, P)
    xor_[[singular, with_nan]] = xfr[[singular, with_nan]]
    assert np.array_equal(xr[singular], xfr[singular]) and np.array_equal(xr[with_nan], xfr[with_nan])
    good = np.ones(n, dtype=bool)
    good[with_nan] = False
    assert np.max(gu.rel(A[good], Ao[good], floor=1e-6)) < RTOL
    # elimination on an indefinite block: errors grow with its condition number (<= 1e7 here)
    assert np.max(np.abs(xr - xor_) / np.maximum(np.abs(xor_), 1e-3)) < 1e-7
    keep = np.ones(n, dtype=bool)
    keep[indef] = False
    assert np.max(np.abs(xr[keep] - xor_[keep]) / np.maximum(np.abs(xor_[keep]), 1e-3)) < RTOL
    assert norm2 == pytest.approx(float(((xor_.ravel() - x_prev) ** 2).sum()), rel=1e-9)
    fr = fwd.reshape(B, n)
    fo = np.stack([H0s[b] + np.einsum("np,np->n", Js[b], xor_ - xfr) for b in range(B)])   # solvers.py:137
    assert np.allclose(fr, fo, rtol=1e-7, atol=1e-8)
    # the second run gives the same bits (the corrections of the partial sums are applied in a fixed order)
    again = _call(n, P, bands, x_prev, x_f, gu.pack_blocks(Pinv), flags=pair, count_lu=True)
    assert again[4] == norm2 and np.array_equal(again[0], x, equal_nan=True)


def test_indefinite_blocks_without_optional_outputs_and_broadcast_prior():
    """The pivoted elimination rebuilds A and b from the inputs: it must not depend on a_out / fwd
    being asked for, and has to honour a broadcast prior block."""
    rng = np.random.default_rng(17)
    n, P, B = 777, 6, 2
    Js, H0s, ws, ys, Pinv, x_f, x_prev = _problem(rng, n, P, B)
    block = -np.eye(P) * 50.0                      # one prior block for all pixels, negative definite
    block[0, 0] = 0.5
    Pinv[:] = block
    bands = [dict(J=Js[b], H0=H0s[b], y=ys[b], w=ws[b]) for b in range(B)]
    pinv_b = torch.from_numpy(np.ascontiguousarray(block[np.tril_indices(P)])).to(gu.DEV)
    x, _, _, _, norm2, n_bad, n_lu = _call(n, P, bands, x_prev, x_f, pinv_b, flags=_lib.PRIOR_BROADCAST, want_fwd=False,
                                           count_lu=True)
    xo, Ao, _, _ = kc.solve_dense64(ys, ws, H0s, Js, x_prev, x_f, Pinv)
    n_indef = int((np.linalg.eigvalsh(Ao).min(axis=1) <= 0).sum())
    assert n_indef > n // 2 and n_lu == n_indef and n_bad == 0
    assert np.max(np.abs(x - xo) / np.maximum(np.abs(xo), 1e-3)) < 1e-7
    assert norm2 == pytest.approx(float(((xo - x_prev) ** 2).sum()), rel=1e-9)


def test_reference_signature_solver_vs_unmodified_reference():
    """variational_kalman_multiband with the reference's own argument list on the
    inputs of the solve_s2 fixture: operators come from this package's builder
    (called like the reference's), result compared with the unmodified
    reference's float32 output and with oracle B."""
    g = gio.load("solve_s2")
    mask = g["mask"]
    P = 10
    n = int(mask.sum())
    gps = gio.emulators(g, syn.EmulatorData)
    obs = gio.FixtureObservations(g, gps)
    d0 = obs.dates[0]
    bd = [obs.get_band_data(d0, b) for b in range(10)]
    Hs = [kb.create_prosail_observation_operator(P, gps[b], None, bd[b].mask, mask, g["x_lin"], b) for b in range(10)]
    Pinv = kc.blocks_to_sparse(np.repeat(g["prior_inv_block"][None].astype(np.float64), n, axis=0))
    x, _, A, innov, fwd = kb.variational_kalman_multiband(
        [b.observations for b in bd], [b.mask for b in bd], mask, [b.uncertainty for b in bd], Hs, P,
        g["x_lin"], g["x0"], None, Pinv, None)
    x = np.asarray(x)
    # the reference's own numbers (float32 solve): diag(A) to float32 rounding, x to its LU error
    assert np.allclose(A.diagonal(), g["ref_A_diag"], rtol=3e-7)
    assert np.median(gu.rel(x, g["ref_x"])) < 5e-6 and np.allclose(x, g["ref_x"], rtol=1e-3, atol=1e-5)
    assert np.allclose(innov, g["ref_innovations"], rtol=1e-6, atol=1e-7)
    assert np.allclose(fwd, g["ref_fwd"], rtol=1e-3, atol=1e-5)
    # oracle B on the same float32 operators: 1e-6 target, met with margin
    bands = gio.compacted_bands(obs, mask)[d0]
    Js, H0s = [], []
    for H0, H in Hs:
        Hc = H.tocoo()
        J = np.zeros((n, P))
        J[Hc.row, Hc.col - P * Hc.row] = Hc.data
        Js.append(J)
        H0s.append(H0.astype(np.float64))
    blocks = np.repeat(g["prior_inv_block"][None].astype(np.float64), n, axis=0)
    xo, Ao, _, _ = kc.solve_dense64([b[0] for b in bands], [b[1] for b in bands], H0s, Js, g["x_lin"], g["x0"], blocks)
    assert np.max(gu.rel(x, xo)) < 1e-9
    assert np.max(gu.rel(A.blocks(), Ao, floor=1e-3)) < 1e-9


def test_large_image_properties():
    """BASELINE-sized pixel count (config 5 upper range, 4e6 px here): results do
    not depend on where a pixel sits in the image, and a zero-weight band changes
    nothing (linearity in the weights)."""
    n, P, B = 4_000_000, 10, 3
    gen = torch.Generator(device=gu.DEV).manual_seed(3)
    ld = padded(n)
    lib = _lib.get_lib()

    def rnd(rows):
        return torch.randn((rows, ld), generator=gen, dtype=torch.float64, device=gu.DEV)
    J = [rnd(P) for _ in range(B)]
    h0 = [rnd(1) for _ in range(B)]
    y = [rnd(1) for _ in range(B)]
    w = [torch.full((1, ld), 400.0, dtype=torch.float64, device=gu.DEV) for _ in range(B)]
    w[2].zero_()
    xp, xf = rnd(P), rnd(P)
    L = tri_count(P)
    r, c = np.tril_indices(P)
    pinv = torch.zeros((L, ld), dtype=torch.float64, device=gu.DEV)
    for k in range(L):
        pinv[k] = 4.0 if r[k] == c[k] else 0.0

    def run(nb, sel=None):
        arr = (_lib.SolveBand * nb)()
        views = []
        for b in range(nb):
            tens = [t if sel is None else t[:, sel].contiguous() for t in (J[b], h0[b], y[b], w[b])]
            views.append(tens)
            arr[b].kind = _lib.BAND_ARRAYS
            arr[b].n_dim = P
            for k in range(P):
                arr[b].mapper[k] = k
            arr[b].jac, arr[b].h0, arr[b].y, arr[b].w = [t.data_ptr() for t in tens]
        a, b_, pv = [t if sel is None else t[:, sel].contiguous() for t in (xp, xf, pinv)]
        m = a.shape[1]
        xo = torch.empty_like(a)
        ao = torch.empty_like(pv)
        n2 = torch.zeros(1, dtype=torch.float64, device=gu.DEV)
        nbad = torch.zeros(1, dtype=torch.int32, device=gu.DEV)
        scratch = torch.empty(int(lib.kafka_solve_scratch_doubles(m)), dtype=torch.float64, device=gu.DEV)
        _lib.check(lib.kafka_solve_bands(m, m, P, arr, nb, _lib.ptr(a), _lib.ptr(b_), _lib.ptr(pv), _lib.ptr(xo),
                                         _lib.ptr(ao), None, None, _lib.ptr(n2), _lib.ptr(nbad), _lib.ptr(scratch), 0,
                                         gu.stream()), "solve")
        assert nbad.item() == 0
        return xo, ao, n2.item()
    x3, a3, n3 = run(3)
    x2, a2, n2 = run(2)
    assert torch.equal(x3, x2) and torch.equal(a3, a2) and n3 == n2      # zero-weight band is a no-op
    sel = torch.arange(1_000_000, 1_000_000 + 4096, device=gu.DEV)
    xs, as_, _ = run(2, sel)
    assert torch.equal(xs, x2[:, sel]) and torch.equal(as_, a2[:, sel])
    # residual of the normal equations on a sample: A x = rhs to rounding
    s = slice(0, 2048)
    Ab = gu.unpack_blocks(a2[:, s], 2048, P)
    xs_ = x2[:, s].t().cpu().numpy()
    rhs = 4.0 * xf[:, s].t().cpu().numpy()
    for b in range(2):
        Jb = J[b][:, s].t().cpu().numpy()
        yp = y[b][0, s].cpu().numpy() + (Jb * xp[:, s].t().cpu().numpy()).sum(1) - h0[b][0, s].cpu().numpy()
        rhs += 400.0 * yp[:, None] * Jb
    res = np.einsum("npq,nq->np", Ab, xs_) - rhs
    assert np.max(np.abs(res)) < 1e-9 * np.max(np.abs(rhs))


@pytest.mark.parametrize("pair", [0, _lib.SOLVE_PAIR])
def test_indefinite_blocks_vs_unmodified_reference(pair):
    """tests/golden/solve_indefinite.npz: the UNMODIFIED reference's variational_kalman_multiband
    (solvers.py:100-145) and blend_prior (kf_tools.py:75-96) on forecast blocks that are not positive
    definite for every third pixel.  Its sparse LU returns the solution of those systems; the
    library reaches the same numbers through the pivoted elimination of its second stage (to the
    float32 accuracy of the reference's LU) and agrees with oracle B to 1e-9."""
    from kafka_inferenceengine_b200.device import BlockDiag
    g = gio.load("solve_indefinite")
    P = 7
    blocks = g["blocks"]
    n = blocks.shape[0]
    mask = g["mask"]
    bands, ys, ws, H0s, Js = [], [], [], [], []
    for b in range(g["obs"].shape[0]):
        y, r, m = kc.gather_band(g["obs"][b], g["inv_var"][b], g["obs_mask"][b], mask)
        w = kc.gate_inverse_variance(r, m)
        J, H0 = g["J"][b].astype(np.float64), g["H0"][b].astype(np.float64)
        bands.append(dict(J=J, H0=H0, y=y, w=w))
        ys.append(y)
        ws.append(w)
        H0s.append(H0)
        Js.append(J)
    x, A, innov, fwd, norm2, n_bad, n_lu = _call(n, P, bands, g["x_lin"], g["x_f"], gu.pack_blocks(blocks), flags=pair,
                                                 count_lu=True)
    xo, Ao, innov_o, fwd_o = kc.solve_dense64(ys, ws, H0s, Js, g["x_lin"], g["x_f"], blocks)
    ev = np.linalg.eigvalsh(Ao)
    assert n_bad == 0 and n_lu == int((ev.min(axis=1) < 0).sum()) and n_lu >= 12
    assert np.max(np.abs(x - xo) / np.maximum(np.abs(xo), 1e-3)) < RTOL
    assert norm2 == pytest.approx(float(((xo - g["x_lin"]) ** 2).sum()), rel=1e-9)
    assert np.allclose(fwd, fwd_o, rtol=1e-8, atol=1e-9)
    # against the reference's float32 LU: errors of the order of cond(A) * 6e-8
    cond = np.abs(ev).max(axis=1) / np.abs(ev).min(axis=1)
    err = np.abs(x - g["ref_x"]).reshape(n, P).max(axis=1) / np.abs(g["ref_x"]).reshape(n, P).max(axis=1)
    assert (err < 2e-6 * np.maximum(cond, 10.0)).all(), (err / cond).max()
    idx = np.arange(P)
    assert np.allclose(A[:, idx, idx].ravel(), g["ref_A_diag"], rtol=3e-6, atol=1e-5)
    if not pair:
        xb, Cb = kb.blend_prior(g["prior_mean"], BlockDiag.from_blocks(g["prior_inv_blocks"], gu.DEV), g["x_f"],
                                BlockDiag.from_blocks(blocks, gu.DEV))
        xbo, Cbo = kc.blend_prior(g["prior_mean"], g["prior_inv_blocks"], g["x_f"], blocks)
        assert np.max(np.abs(np.asarray(xb) - xbo) / np.maximum(np.abs(xbo), 1e-3)) < 1e-9
        evc = np.linalg.eigvalsh(Cbo)
        condc = np.abs(evc).max(axis=1) / np.abs(evc).min(axis=1)
        errb = np.abs(np.asarray(xb) - g["ref_blend_x"]).reshape(n, P).max(axis=1) / \
            np.abs(g["ref_blend_x"]).reshape(n, P).max(axis=1)
        assert (errb < 2e-6 * np.maximum(condc, 10.0)).all(), (errb / condc).max()
        assert np.allclose(Cb.blocks()[:, idx, idx].ravel(), g["ref_blend_C_diag"], rtol=1e-6)
