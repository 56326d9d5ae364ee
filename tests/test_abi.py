"""CPU-only checks of the C-ABI boundary: the library loads, exports every symbol
that include/kafka_b200.h declares, and validates arguments before touching the
GPU (no compute calls here)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from kafka_inferenceengine_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "kafka_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(kafka_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.get_lib()
    names = declared_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), "missing export: " + name
    # and the python binding declares a prototype for each of them
    assert sorted(_lib.SIGNATURES) == names


def test_version_and_error_string():
    lib = _lib.get_lib()
    assert lib.kafka_version() >= 100
    assert isinstance(lib.kafka_last_error(), bytes)


def test_struct_layouts_match_header():
    # kafka_gp_band: pointer + 2 int32 + 16 int32; kafka_solve_band: see header
    assert C.sizeof(_lib.GpBand) == 8 + 4 + 4 + 4 * _lib.MAX_DIM
    assert C.sizeof(_lib.SolveBand) == 4 + 4 + 4 * _lib.MAX_DIM + 8 * _lib.MAX_PARAMS + 8 + 4 * 8


def test_bad_arguments_are_rejected_without_a_gpu():
    lib = _lib.get_lib()
    assert lib.kafka_gp_table_doubles(250, 17) == -1
    assert lib.kafka_gp_table_doubles(250, 10) == 2 * 10 + 4 + 250 * 14
    assert lib.kafka_gp_table_doubles(250, 7) == 2 * 8 + 4 + 250 * 12
    assert lib.kafka_solve_scratch_doubles(-1) == -1
    assert lib.kafka_compact_scratch_elems(4097) == 4
    st = lib.kafka_interleaved_to_planar(None, 10, 0, None, 32, None)
    assert st == -1 and b"n_params" in lib.kafka_last_error()
    st = lib.kafka_interleaved_to_planar(None, 10, 4, None, 5, None)
    assert st == -1 and b"ld" in lib.kafka_last_error()
    st = lib.kafka_solve_bands(10, 32, 17, None, 0, None, None, None, None, None, None, None, None, None, None, 0, None)
    assert st == -1
    st = lib.kafka_solve_bands(10, 32, 4, None, 17, None, None, None, None, None, None, None, None, None, None, 0, None)
    assert st == -1
    st = lib.kafka_advance_carry_slot(None, None, None, 32, 10, 7, 9, None, None, None, None, None)
    assert st == -1 and b"slot" in lib.kafka_last_error()
    with pytest.raises(_lib.KafkaCudaError):
        _lib.check(st, "kafka_advance_carry_slot")


def test_gp_pack_host_matches_numpy():
    rng = np.random.default_rng(0)
    for T, D in ((5, 1), (17, 4), (30, 7), (250, 10)):
        X = rng.standard_normal((T, D))
        theta = rng.standard_normal(D + 2) * 0.3
        alpha = rng.standard_normal(T)
        tab = _lib.gp_pack(X, theta, alpha)
        DP = (D + 1) & ~1
        assert tab.size == 2 * DP + 4 + T * (DP + 4)
        e = np.exp(theta)
        c = X.mean(axis=0)
        assert np.allclose(tab[:D], e[:D], rtol=1e-15)
        assert np.allclose(tab[DP:DP + D], c, rtol=1e-14, atol=1e-16)
        assert tab[2 * DP] == pytest.approx(e[D], rel=1e-15)
        cbar = np.log(np.abs(alpha)).mean()
        L = 1024.0 / np.log(2.0)                     # rows are in units of ln2/1024
        assert tab[2 * DP + 2] == 0.0                                # not a "wide" training set
        assert tab[2 * DP + 3] == pytest.approx(L * cbar, rel=1e-14)
        rows = tab[2 * DP + 4:].reshape(T, DP + 4)
        Xc = X - tab[DP:DP + D]
        assert np.allclose(rows[:, :D], L * e[:D] * Xc, rtol=4e-16, atol=0)
        assert np.allclose(rows[:, DP], L * (theta[D] - 0.5 * (e[:D] * Xc * Xc).sum(axis=1)), rtol=1e-14)
        assert np.array_equal(rows[:, DP + 1], alpha)
        # the pair the linearise loop reads: exponent with ln|alpha| - cbar folded in, and the sign bit of
        # the weight as an integer in the slot's low word (XOR-ed into the result's high word)
        e_plain = theta[D] - 0.5 * (e[:D] * Xc * Xc).sum(axis=1)
        assert np.allclose(rows[:, DP + 2], L * (e_plain + np.log(np.abs(alpha)) - cbar), rtol=1e-13, atol=1e-9)
        assert np.array_equal(rows[:, DP + 3].view(np.uint64), np.where(alpha < 0, 0x80000000, 0).astype(np.uint64))
        if DP != D:
            assert np.all(rows[:, D] == 0.0)
    with pytest.raises(ValueError):
        _lib.gp_pack(np.zeros((3, 17)), np.zeros(19), np.zeros(3))
    # a training point ~2000 length scales from the centre: flagged, the kernels then keep the
    # range-checked exponential
    X = np.zeros((4, 2))
    X[0] = 6000.0
    tab = _lib.gp_pack(X, np.zeros(4), np.ones(4))
    assert tab[2 * 2 + 2] == 1.0
    # a zero weight gets an exponent far below underflow and no sign offset; a non-finite one is refused
    tab = _lib.gp_pack(np.zeros((3, 2)), np.zeros(4), np.array([0.0, 2.0, -2.0]))
    rows = tab[2 * 2 + 4:].reshape(3, 2 + 4)
    assert rows[0, 4] == -2.0 ** 29 and rows[0, 5].view(np.uint64) == 0 and rows[2, 5].view(np.uint64) == 0x80000000
    with pytest.raises(_lib.KafkaCudaError):
        _lib.gp_pack(np.zeros((3, 2)), np.zeros(4), np.array([1.0, np.nan, 1.0]))


def test_header_constants_match_the_python_binding():
    """The numbers a binder copies out of include/kafka_b200.h are the ones _lib.py uses, and the
    built-in scratch of kafka_solve_bands_static is exactly as large as its pixel limit needs."""
    text = open(HEADER).read()

    def const(name):
        m = re.search(r"#define\s+%s\s+\(?(-?\d+)u?\)?" % name, text)
        assert m, name
        return int(m.group(1))
    assert const("KAFKA_MAX_PARAMS") == _lib.MAX_PARAMS and const("KAFKA_MAX_DIM") == _lib.MAX_DIM
    assert const("KAFKA_ROUND_F32") == _lib.ROUND_F32 and const("KAFKA_PRIOR_BROADCAST") == _lib.PRIOR_BROADCAST
    assert const("KAFKA_SOLVE_PAIR") == _lib.SOLVE_PAIR and const("KAFKA_SOLVE_COUNT_LU") == _lib.SOLVE_COUNT_LU
    assert const("KAFKA_STATIC_SCRATCH_DOUBLES") == _lib.STATIC_SCRATCH_DOUBLES
    assert const("KAFKA_STATIC_MAX_PIXELS") == _lib.STATIC_MAX_PIXELS
    assert const("KAFKA_COMM_ID_BYTES") == _lib.COMM_ID_BYTES
    lib = _lib.get_lib()
    assert lib.kafka_solve_scratch_doubles(_lib.STATIC_MAX_PIXELS) == _lib.STATIC_SCRATCH_DOUBLES
    assert lib.kafka_solve_scratch_doubles(_lib.STATIC_MAX_PIXELS + 1) > _lib.STATIC_SCRATCH_DOUBLES
    assert lib.kafka_solve_scratch_doubles(0) > 0            # the second stage's own slots


def test_new_entry_points_reject_bad_arguments_without_a_gpu():
    lib = _lib.get_lib()
    st = lib.kafka_output_rasters(None, None, 32, 10, 4, None, 100, None, None, 0, None, None)
    assert st == -1 and b"raster" in lib.kafka_last_error()
    st = lib.kafka_output_rasters(None, None, 32, 10, 17, None, 100, C.c_void_p(256), None, 0, None, None)
    assert st == -1 and b"n_params" in lib.kafka_last_error()
    # the two rasters have to be one allocation, mean first
    st = lib.kafka_output_rasters(None, None, 32, 10, 4, None, 100, C.c_void_p(4096), C.c_void_p(4096 + 8), 0, None, None)
    assert st == -1 and b"one allocation" in lib.kafka_last_error()
    st = lib.kafka_upload_rasters(3, None, None, None, None)
    assert st == -1
