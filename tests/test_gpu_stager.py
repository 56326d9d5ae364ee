"""Host -> device path of one date's rasters (engine.ObservationStager): rasters handed
over in page-locked memory (kb.pinned_raster) are fetched by the DMA engine in place,
pageable ones go through the package's staging buffers; both must give bit-identical
compacted rows, equal to the reference's gathers (solvers.py:90-93)."""
import numpy as np
import pytest
import scipy.sparse as sp

import gpu_util as gu

pytestmark = pytest.mark.gpu

import kafka_inferenceengine_b200 as kb                          # noqa: E402
from kafka_inferenceengine_b200 import synthetic as syn          # noqa: E402
from kafka_inferenceengine_b200.device import PixelIndex         # noqa: E402
from kafka_inferenceengine_b200.engine import ObservationStager  # noqa: E402
from oracle import kafka_cpu as kc                               # noqa: E402


def _pinned(a):
    p = kb.pinned_raster(a.shape, a.dtype)
    p[...] = a
    return p


def _rasters(rng, shape, n_bands):
    out = []
    for _ in range(n_bands):
        obs = rng.random(shape)
        m = rng.random(shape) < 0.8
        inv = np.where(m, 1.0 / (0.05 * obs) ** 2, np.inf)
        out.append((obs, inv, m))
    return out


@pytest.mark.parametrize("shape", [(600, 700), (90, 100)])     # above / below the small-raster threshold
def test_pinned_and_pageable_sources_give_the_reference_gathers(shape):
    rng = np.random.default_rng(5)
    state_mask = rng.random(shape) < 0.6
    N = int(state_mask.sum())
    pix = PixelIndex(state_mask, gu.DEV)
    ras = _rasters(rng, shape, 4)
    for pinned in (False, True, "mixed", "dia"):
        data = []
        for b, (obs, inv, m) in enumerate(ras):
            pin = pinned is True or (pinned == "mixed" and b % 2 == 0) or pinned == "dia"
            f = _pinned if pin else (lambda a: a)
            unc = f(inv)
            if pinned == "dia":
                # the reference's data contract: a (ny*nx)^2 sparse matrix with the inverse variance on its diagonal
                unc = sp.dia_matrix((f(inv.reshape(1, -1)), [0]), shape=(inv.size, inv.size))
            data.append(syn.BandData(f(obs), unc, f(m), {"band": b}, None))
        st = ObservationStager(pix)
        bufs = st.upload(data)
        for b, (obs, inv, m) in enumerate(ras):
            y, r, mm = kc.gather_band(obs, inv.ravel(), m, state_mask)
            assert np.array_equal(bufs.y[b, :N].cpu().numpy(), y)
            assert np.array_equal(bufs.w[b, :N].cpu().numpy(), kc.gate_inverse_variance(r, mm))
            assert np.array_equal(bufs.m[b, :N].cpu().numpy().astype(bool), mm)
        assert st.h2d_bytes == 4 * state_mask.size * 17
        slot = st._slots[0]
        if pinned in (True, "dia"):
            assert slot.h_obs is None and slot.h_msk is None          # nothing was staged
        else:
            assert slot.h_obs is not None
        # a second date through the same stager (the other buffer set), sources changed in place
        if pinned is True:
            for d, (obs, inv, m) in zip(data, ras):
                d.observations[...] = obs[::-1]
            bufs = st.upload(data)
            y, r, mm = kc.gather_band(ras[1][0][::-1], ras[1][1].ravel(), ras[1][2], state_mask)
            assert np.array_equal(bufs.y[1, :N].cpu().numpy(), y)


@pytest.mark.parametrize("dtype", [np.int16, np.float32])
def test_raw_reflectance_from_pinned_rasters(dtype):
    rng = np.random.default_rng(6)
    shape = (640, 512)
    state_mask = rng.random(shape) < 0.7
    N = int(state_mask.sum())
    pix = PixelIndex(state_mask, gu.DEV)
    raws = [rng.integers(-20, 6000, size=shape).astype(dtype) for _ in range(3)]
    extra = rng.random(shape) < 0.7
    res = []
    for pin in (False, True):
        f = _pinned if pin else (lambda a: a)
        data = [syn.BandData(f(r), kb.ScaledReflectance(), f(extra) if b == 2 else None, {"band": b}, None)
                for b, r in enumerate(raws)]
        st = ObservationStager(pix)
        bufs = st.upload(data)
        res.append([t[:, :N].cpu().numpy() for t in (bufs.y, bufs.w, bufs.m)])
        assert st.h2d_bytes == 3 * state_mask.size * np.dtype(dtype).itemsize + state_mask.size
        if pin:
            assert st._slots[0].h_raw is None and st._slots[0].h_msk is None
    for a, b in zip(*res):
        assert np.array_equal(a, b)
    for b, raw in enumerate(raws):
        rho, inv, m = kc.s2_ingest(raw, ext_mask=extra if b == 2 else None)
        y, r, mm = kc.gather_band(rho.astype(np.float64), inv.ravel(), m, state_mask)
        assert np.array_equal(res[1][0][b], y)
        assert np.array_equal(res[1][1][b], kc.gate_inverse_variance(r, mm))
        assert np.array_equal(res[1][2][b].astype(bool), mm)
