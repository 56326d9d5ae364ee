"""Run the UNMODIFIED reference on chunks of the benchmark workload and time it
(TEST INFRASTRUCTURE: used by bench.py's ``--impl reference`` arm and
``cpu_baseline`` leg only).

The reference package is imported through ``ref_shim`` from ``/root/reference``
or, on the GPU box, from the copy ``pip install --target baseline/_ref`` made at
build time (``__graft_entry__.build``).  Its emulator class (``gp_emulator``) is
not vendored, so the GP objects handed to it are ``oracle.gp.GaussianProcess``;
everything else on the timed path is the reference's own code:

(A) ``LinearKalman.do_all_bands`` (kafka/linear_kf.py:245-323) with
    ``create_prosail_observation_operator`` (kafka/inference/utils.py:181-219) and
    ``variational_kalman_multiband`` (kafka/inference/solvers.py:100-145) -- the whole
    assimilation of one date for one chunk, iterated to convergence;
(B) ``variational_kalman_multiband`` alone with pre-built operators -- the
    "scipy spsolve path" of BASELINE.json configs[4].

Chunks are square rasters of at most 256 x 256 pixels, one ``LinearKalman`` per
chunk, as the reference's drivers do (kafka_test_S2.py:202, kafka_test_Py36.py:241).
"""
import collections
import logging
import time

import numpy as np
import scipy.sparse as sp

from . import ref_shim
from . import synthetic as osyn
from .gp import GaussianProcess

BandData = collections.namedtuple("BandData", "observations uncertainty mask metadata emulator")
_kafka = None


def reference():
    global _kafka
    if _kafka is None:
        _kafka = ref_shim.load()
        for name in ("kafka.linear_kf.linear_kf", "kafka.inference.solvers", "kafka.inference.utils"):
            logging.getLogger(name).setLevel(logging.WARNING)
    return _kafka


def chunk_inputs(gp_arrays, y, inv, m, side):
    """Collaborator objects of one chunk: band data in the readers' contract (``(ny, nx)``
    rasters, inverse variance as a (ny*nx)^2 sparse diagonal holding inf where masked,
    Sentinel2_Observations.py:176-179), all-True state mask, prior."""
    kafka = reference()
    n = side * side
    B = y.shape[0]
    gps = [GaussianProcess(a[0], a[1], a[2]) for a in gp_arrays]
    data = [BandData(y[b].reshape(side, side), sp.diags(inv[b], format="csr"), m[b].reshape(side, side),
                     {"band": b}, gps[b]) for b in range(B)]
    state_mask = np.ones((side, side), dtype=bool)
    blk = osyn.sail_prior_block()
    x0 = np.tile(osyn.SAIL_MEAN, n)
    Pinv = kafka.inference.utils.block_diag([blk for _ in range(n)], dtype=np.float32)
    return data, state_mask, x0, Pinv


def time_do_all_bands(args):
    """(A) one chunk through the unmodified LinearKalman.do_all_bands.  Returns
    (seconds, pixels, solver iterations)."""
    gp_arrays, y, inv, m, side = args
    kafka = reference()
    data, state_mask, x0, Pinv = chunk_inputs(gp_arrays, y, inv, m, side)
    P = osyn.SAIL_MEAN.size
    kf = kafka.LinearKalman(None, None, state_mask, kafka.inference.utils.create_prosail_observation_operator,
                            osyn.S2_PARAMETERS, state_propagation=None, prior=None, linear=False)
    counter = _IterCounter()
    lg = logging.getLogger("kafka.linear_kf.linear_kf")
    old = lg.level
    lg.setLevel(logging.INFO)
    lg.addHandler(counter)
    try:
        t0 = time.perf_counter()
        x, _, A, _ = kf.do_all_bands(None, data, x0, None, Pinv)
        dt_ = time.perf_counter() - t0
    finally:
        lg.removeHandler(counter)
        lg.setLevel(old)
    assert np.all(np.isfinite(x)) and x.shape == (side * side * P,)
    return dt_, side * side, counter.n


def time_multiband_solve(args):
    """(B) one chunk through the unmodified variational_kalman_multiband with operators
    built beforehand (outside the timed region).  Returns (seconds, pixels)."""
    gp_arrays, y, inv, m, side = args
    kafka = reference()
    data, state_mask, x0, Pinv = chunk_inputs(gp_arrays, y, inv, m, side)
    P = osyn.SAIL_MEAN.size
    ku = kafka.inference.utils
    x_lin = x0 + 0.3 * np.tile(osyn.SAIL_SIGMA, side * side) * np.random.default_rng(5).standard_normal(x0.size)
    Hs = [ku.create_prosail_observation_operator(P, d.emulator, None, d.mask, state_mask, x_lin, b)
          for b, d in enumerate(data)]
    t0 = time.perf_counter()
    x, _, A, innov, fwd = kafka.inference.solvers.variational_kalman_multiband(
        [d.observations for d in data], [d.mask for d in data], state_mask, [d.uncertainty for d in data],
        Hs, P, x_lin, x0, None, Pinv.tocsr(), None)
    dt_ = time.perf_counter() - t0
    assert np.all(np.isfinite(x))
    return dt_, side * side


class _IterCounter(logging.Handler):
    def __init__(self):
        super().__init__()
        self.n = 0

    def emit(self, record):
        if "convergence norm" in record.getMessage():
            self.n += 1


def time_supplied_jacobian_solve(n_params, n_bands, side, seed):
    """BASELINE configs[4]'s CPU arm: the unmodified variational_kalman_multiband on one chunk of
    ``side x side`` pixels with SUPPLIED random Jacobians in the solver's (H0, H csr) tuple form
    (J ~ N(0,1), inverse variance 400, P_f^-1 = 4 I; the synthetic inputs of SURVEY section 8d).
    Returns (seconds, pixels)."""
    kafka = reference()
    rng = np.random.default_rng(seed)
    n = side * side
    P = n_params
    state_mask = np.ones((side, side), dtype=bool)
    indptr = np.arange(0, (n + 1) * P, P)
    indices = np.arange(n * P)
    Hs, obs, unc, masks = [], [], [], []
    for _ in range(n_bands):
        J = rng.standard_normal((n, P)).astype(np.float32)
        Hs.append((rng.standard_normal(n).astype(np.float32), sp.csr_matrix((J.ravel(), indices, indptr), shape=(n, n * P))))
        obs.append(rng.standard_normal((side, side)))
        unc.append(sp.diags(np.full(n, 400.0), format="csr"))
        masks.append(state_mask)
    x_lin = rng.standard_normal(n * P)
    x_f = rng.standard_normal(n * P)
    Pinv = sp.diags(np.full(n * P, 4.0), format="csr").astype(np.float32)
    t0 = time.perf_counter()
    x, _, A, innov, fwd = kafka.inference.solvers.variational_kalman_multiband(
        obs, masks, state_mask, unc, Hs, P, x_lin, x_f, None, Pinv, None)
    dt_ = time.perf_counter() - t0
    assert np.all(np.isfinite(x)) and x.shape == (n * P,)
    return dt_, n
