"""GPU parity: state-mask compaction, pixel indexing and layout conversion
 -- bit-exact against the oracle."""
import ctypes as C
import hashlib

import numpy as np
import pytest
import torch

import golden_io as gio
import gpu_util as gu
from kafka_inferenceengine_b200 import _lib
from kafka_inferenceengine_b200.device import BlockDiag, DeviceVector, PixelIndex, padded, tri_count
from oracle import kafka_cpu as kc

pytestmark = pytest.mark.gpu


def _masks():
    rng = np.random.default_rng(3)
    yield "barrax", gio.barrax_mask()
    yield "empty", np.zeros((7, 9), dtype=bool)
    yield "full", np.ones((33, 65), dtype=bool)
    yield "single", np.eye(1, dtype=bool)
    yield "ragged", rng.random((61, 67)) < 0.3            # 4087 px: not a multiple of 16
    yield "one_tile_edge", rng.random((64, 64)) < 0.5      # exactly 4096
    yield "big", rng.random((1500, 1400)) < 0.27           # many CTAs, scan over 500+ blocks


@pytest.mark.parametrize("name,mask", list(_masks()), ids=[m[0] for m in _masks()])
def test_compaction_bit_exact(name, mask):
    pix = PixelIndex(mask, gu.DEV)
    want = kc.compact_mask(mask)
    assert pix.n_pix == want.size == int(mask.sum())
    got = pix.idx[:pix.n_pix].cpu().numpy()
    assert got.dtype == np.int32
    assert np.array_equal(got, want)
    if name == "barrax":
        g = gio.load("barrax_mask")
        assert hashlib.sha256(got.astype(np.int64).tobytes()).hexdigest() == str(g["idx_sha256"])


def test_compaction_unaligned_mask_pointer():
    rng = np.random.default_rng(4)
    base = torch.from_numpy((rng.random(10007) < 0.4).astype(np.uint8)).to(gu.DEV)
    lib = _lib.get_lib()
    for off in (1, 3, 8):
        view = base[off:]
        n = view.numel()
        scratch = torch.empty(int(lib.kafka_compact_scratch_elems(n)), dtype=torch.int32, device=gu.DEV)
        idx = torch.empty(n, dtype=torch.int32, device=gu.DEV)
        cnt = C.c_int64(0)
        _lib.check(lib.kafka_compact_mask(_lib.ptr(view), n, _lib.ptr(idx), n, _lib.ptr(scratch), C.byref(cnt),
                                          gu.stream()), "compact")
        want = np.flatnonzero(view.cpu().numpy())
        assert cnt.value == want.size
        assert np.array_equal(idx[:cnt.value].cpu().numpy(), want)


def test_compaction_capacity_error():
    lib = _lib.get_lib()
    m = torch.ones(100, dtype=torch.uint8, device=gu.DEV)
    scratch = torch.empty(int(lib.kafka_compact_scratch_elems(100)), dtype=torch.int32, device=gu.DEV)
    idx = torch.empty(10, dtype=torch.int32, device=gu.DEV)
    cnt = C.c_int64(0)
    st = lib.kafka_compact_mask(_lib.ptr(m), 100, _lib.ptr(idx), 10, _lib.ptr(scratch), C.byref(cnt), gu.stream())
    assert st == -3 and cnt.value == 100


def test_gather_band_matches_oracle_and_gates_inf():
    rng = np.random.default_rng(5)
    mask = gio.barrax_mask()
    pix = PixelIndex(mask, gu.DEV)
    obs = rng.random(mask.shape)
    omask = rng.random(mask.shape) < 0.8
    with np.errstate(divide="ignore"):
        inv = np.where(omask, 1.0 / (0.05 * obs) ** 2, np.inf)       # Sentinel2_Observations.py:175-178
    y = torch.empty(pix.ld, dtype=torch.float64, device=gu.DEV)
    w = torch.empty_like(y)
    m = torch.empty(pix.ld, dtype=torch.uint8, device=gu.DEV)
    pix.gather_band(torch.from_numpy(obs.ravel()).to(gu.DEV), torch.from_numpy(inv.ravel()).to(gu.DEV),
                    torch.from_numpy(omask.ravel().astype(np.uint8)).to(gu.DEV), y, w, m)
    yo, ro, mo = kc.gather_band(obs, inv.ravel(), omask, mask)
    n = pix.n_pix
    assert np.array_equal(y[:n].cpu().numpy(), yo)
    assert np.array_equal(w[:n].cpu().numpy(), kc.gate_inverse_variance(ro, mo))
    assert np.array_equal(m[:n].cpu().numpy().astype(bool), mo)
    assert np.isfinite(w[:n].cpu().numpy()).all()


def test_gather_bands_one_launch_equals_per_band_gathers():
    rng = np.random.default_rng(15)
    mask = gio.barrax_mask()
    pix = PixelIndex(mask, gu.DEV)
    nb, n_r, n = 4, mask.size, pix.n_pix
    obs = rng.random((nb, n_r))
    omask = rng.random((nb, n_r)) < 0.7
    with np.errstate(divide="ignore"):
        inv = np.where(omask, 1.0 / (0.05 * obs) ** 2, np.inf)
    d_obs, d_inv = torch.from_numpy(obs).to(gu.DEV), torch.from_numpy(inv).to(gu.DEV)
    d_msk = torch.from_numpy(omask.astype(np.uint8)).to(gu.DEV)
    y = torch.full((nb, pix.ld), -1.0, dtype=torch.float64, device=gu.DEV)
    w = torch.full_like(y, -1.0)
    m = torch.full((nb, pix.ld), 9, dtype=torch.uint8, device=gu.DEV)
    pix.gather_bands(d_obs, d_inv, d_msk, nb, y, w, m)
    for b in range(nb):
        yo, ro, mo = kc.gather_band(obs[b].reshape(mask.shape), inv[b], omask[b].reshape(mask.shape), mask)
        assert np.array_equal(y[b, :n].cpu().numpy(), yo)
        assert np.array_equal(w[b, :n].cpu().numpy(), kc.gate_inverse_variance(ro, mo))
        assert np.array_equal(m[b, :n].cpu().numpy().astype(bool), mo)
    assert (y[:, n:] == -1.0).all() and (m[:, n:] == 9).all()          # padding columns untouched


def test_scatter_output_matches_oracle():
    rng = np.random.default_rng(6)
    mask = gio.barrax_mask()
    pix = PixelIndex(mask, gu.DEV)
    P = 10
    x = rng.standard_normal(pix.n_pix * P)
    blocks = gu.random_spd(rng, pix.n_pix, P)
    xv = DeviceVector.from_interleaved(x, P, gu.DEV)
    A = BlockDiag.from_blocks(blocks, gu.DEV)
    mean_r, unc_r = pix.scatter_output(xv, A, P)
    diag = blocks[:, np.arange(P), np.arange(P)].ravel()
    for ii in range(P):
        assert np.array_equal(mean_r[ii].cpu().numpy(), kc.scatter_param(x, ii, P, mask))
        assert np.allclose(unc_r[ii].cpu().numpy(), kc.scatter_unc(diag, ii, P, mask), rtol=2e-7, atol=0)


@pytest.mark.parametrize("P", [1, 2, 7, 10, 16])
@pytest.mark.parametrize("n", [1, 31, 128, 1000])
def test_interleave_roundtrip_bit_exact(P, n):
    rng = np.random.default_rng(P * 1000 + n)
    x = rng.standard_normal(n * P)
    v = DeviceVector.from_interleaved(x, P, gu.DEV)
    assert v.ld == padded(n) and v.ld >= n
    assert np.array_equal(v.data[:, :n].cpu().numpy(), x.reshape(n, P).T)     # planar[p][i] = x[P*i+p]
    assert np.array_equal(np.asarray(v), x)
    assert np.array_equal(v[3 % P::P], x[3 % P::P])


@pytest.mark.parametrize("P", [1, 3, 7, 10, 16])
def test_blocks_pack_unpack_diagonal(P):
    rng = np.random.default_rng(P)
    n = 777
    blocks = gu.random_spd(rng, n, P)
    A = BlockDiag.from_blocks(blocks, gu.DEV)
    assert A.data.shape == (tri_count(P), padded(n))
    r, c = np.tril_indices(P)
    assert np.array_equal(A.data[:, :n].cpu().numpy(), blocks[:, r, c].T)
    assert np.array_equal(A.blocks(), blocks)
    assert np.array_equal(A.diagonal(), blocks[:, np.arange(P), np.arange(P)].ravel())
    S = kc.blocks_to_sparse(blocks)
    B = BlockDiag.from_scipy(S, P, gu.DEV)
    assert np.array_equal(B.blocks(), blocks)
    assert (abs(A.to_scipy() - S)).max() == 0
    v = rng.standard_normal(n * P)
    assert np.allclose(A.dot(v), S.dot(v), rtol=1e-12)
    C2 = A + B * 1.0
    assert np.array_equal(C2.blocks(), 2 * blocks)
    bb = BlockDiag.from_block(blocks[0], n, gu.DEV)
    assert bb.is_broadcast and np.array_equal(bb.blocks()[5], blocks[0])
    assert np.array_equal((bb + A).blocks(), blocks + blocks[0][None])
