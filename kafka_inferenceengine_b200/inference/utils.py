"""Counterpart of kafka/inference/utils.py: time-grid iteration, observation
operator builders and ``block_diag``, with the emulator evaluated by the CUDA
kernels of libkafka_b200.so.

The reference's operator builders are plain Python callables returning
``(H0, H)`` with ``H`` a float32 CSR matrix (utils.py:130-219).  Arbitrary Python
callables cannot run per pixel on the device and there is no CPU fallback, so
the builders here are *descriptor objects*: ``LinearKalman`` recognises them and
runs the matching fused kernels; calling one directly (same signature as the
reference) evaluates that single band on the GPU and returns the reference's
``(H0, H)`` tuple for inspection / tests.
"""
import datetime as dt                      # noqa: F401  (parity with the reference's namespace)
import logging
import weakref

import numpy as np
import scipy.sparse as sp
import torch

from .. import _lib
from ..device import BlockDiag, DeviceVector, default_device, padded

LOG = logging.getLogger(__name__)

__all__ = ["iterate_time_grid", "run_emulator", "create_uncertainty",
           "create_linear_observation_operator", "create_nonlinear_observation_operator",
           "create_prosail_observation_operator", "LinearBandOperator", "block_diag",
           "EmulatorTable", "emulator_table", "ObservationOperator"]


def iterate_time_grid(time_grid, the_dates):
    """kafka/inference/utils.py:44-65 (host logic, same yields and log lines)."""
    is_first = True
    istart_date = time_grid[0]
    dates = np.array(the_dates)
    for timestep in time_grid[1:]:
        sel = np.logical_and(dates >= istart_date, dates < timestep)
        locate_times = dates[sel]
        LOG.info("Doing timestep from {} -> {}".format(istart_date.strftime("%Y-%m-%d"),
                                                      timestep.strftime("%Y-%m-%d")))
        LOG.info("# of Observations: %d" % len(locate_times))
        for iobs in locate_times:
            LOG.info("\t->{}".format(iobs.strftime("%Y-%m-%d")))
        istart_date = timestep
        yield timestep, locate_times, is_first
        is_first = False


# --------------------------------------------------------------------------
# emulator tables (device) --------------------------------------------------
# --------------------------------------------------------------------------
class EmulatorTable(object):
    """Device copy of one GP emulator in the layout the kernels stream
    (include/kafka_b200.h: kafka_gp_pack_host).  Built from the fields of a
    ``gp_emulator.GaussianProcess``-like object: ``inputs``, ``theta`` (log
    hyper-parameters) and ``invQt``."""

    def __init__(self, emulator, device=None):
        device = device or default_device()
        self.n_train, self.n_dim = np.asarray(emulator.inputs).shape
        host = _lib.gp_pack(emulator.inputs, emulator.theta, emulator.invQt)
        self.table = torch.from_numpy(host).to(device)
        self._inv_q = None
        self._emulator = weakref.ref(emulator) if _weakrefable(emulator) else (lambda: emulator)

    def inv_q(self):
        if self._inv_q is None:
            em = self._emulator()
            self._inv_q = torch.from_numpy(np.ascontiguousarray(em.invQ, dtype=np.float64)).to(self.table.device)
        return self._inv_q


def _weakrefable(obj):
    try:
        weakref.ref(obj)
        return True
    except TypeError:
        return False


_TABLES = {}


def emulator_table(emulator, device=None):
    """Cached ``EmulatorTable`` for an emulator object."""
    device = device or default_device()
    key = (id(emulator), str(device))
    hit = _TABLES.get(key)
    if hit is not None and hit[0]() is emulator:
        return hit[1]
    tab = EmulatorTable(emulator, device)
    ref = weakref.ref(emulator, lambda _r, k=key: _TABLES.pop(k, None)) if _weakrefable(emulator) \
        else (lambda e=emulator: e)
    _TABLES[key] = (ref, tab)
    return tab


def run_emulator(gp, x, tol=None):
    """kafka/inference/utils.py:68-106: emulator mean and Jacobian at the rows of
    ``x``.  The reference evaluates unique rows only and scatters the result
    back (value-neutral); above 1e6 unique rows it substitutes an UNSEEDED random
    5000-row look-up table (utils.py:75-84), which is an approximation this
    implementation does not reproduce: every row is evaluated exactly."""
    x = np.atleast_2d(np.asarray(x, dtype=np.float64))
    tab = emulator_table(gp)
    n, D = x.shape
    if D != tab.n_dim:
        raise ValueError("emulator expects %d inputs, got %d" % (tab.n_dim, D))
    dev = tab.table.device
    ld = padded(n)
    xin = torch.zeros((D, ld), dtype=torch.float64, device=dev)
    xin[:, :n] = torch.from_numpy(np.ascontiguousarray(x.T)).to(dev)
    mean = torch.empty(ld, dtype=torch.float64, device=dev)
    jac = torch.empty((D, ld), dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.get_lib().kafka_gp_predict(_lib.ptr(xin), ld, n, _lib.ptr(tab.table), tab.n_train, D,
                                                   None, _lib.ptr(mean), _lib.ptr(jac), None, _lib.stream_ptr()),
                   "kafka_gp_predict")
    return mean[:n].cpu().numpy(), jac[:, :n].t().contiguous().cpu().numpy()


def create_uncertainty(uncertainty, mask):
    """kafka/inference/utils.py:109-116."""
    good_obs = int(np.asarray(mask).sum())
    r = np.ones(good_obs) * uncertainty * uncertainty
    return sp.dia_matrix((r, 0), shape=(good_obs, good_obs))


# --------------------------------------------------------------------------
# observation operators ------------------------------------------------------
# --------------------------------------------------------------------------
class LinearBandOperator(object):
    """Payload for ``create_linear_observation_operator``: the band observes
    ``row . x + offset`` of every pixel's state (passed where the reference
    passes an emulator)."""

    def __init__(self, row, offset=0.0):
        self.row = np.asarray(row, dtype=np.float64)
        self.offset = float(offset)


class ObservationOperator(object):
    """Descriptor of one of the supported operator families.

    kind ``"gp"``     GP emulator fed by the state slots ``mapper(band)``
    kind ``"linear"`` constant row + offset (tuple-form linear operator)
    """

    def __init__(self, name, kind, state_mappers=None):
        self.__name__ = name
        self.kind = kind
        self.state_mappers = state_mappers

    def mapper(self, n_params, band, n_dim):
        if self.state_mappers is None:
            if n_dim != n_params:
                raise ValueError("%s feeds the full %d-parameter state to the emulator, which expects %d inputs"
                                 % (self.__name__, n_params, n_dim))
            return np.arange(n_params, dtype=np.int32)
        if band >= len(self.state_mappers):
            raise ValueError("%s has no state mapper for band %d" % (self.__name__, band))   # utils.py:148-153
        return np.asarray(self.state_mappers[band], dtype=np.int32)

    # the reference's call signature (utils.py:130-131,181-182)
    def __call__(self, n_params, emulator, metadata, mask, state_mask, x_forecast, band):
        LOG.info("Creating the ObsOp for band %d" % band)
        state_mask = np.asarray(state_mask, dtype=bool)
        obs = np.asarray(mask, dtype=bool)[state_mask]
        x = DeviceVector.from_interleaved(x_forecast, n_params)
        n = x.n_pix
        if self.kind == "linear":
            row = emulator.row
            xs = np.asarray(x).reshape(n, n_params)
            H0 = np.where(obs, xs @ row + emulator.offset, 0.0)
            J = np.where(obs[:, None], row[None, :], 0.0)
        else:
            LOG.info("Running emulators")
            tab = emulator_table(emulator, x.device)
            mp = self.mapper(n_params, band, tab.n_dim)
            h0, jac = linearize_bands(x, [(tab, mp)], torch.from_numpy(obs.astype(np.uint8)).to(x.device), True)
            H0 = h0[0, :n].cpu().numpy()
            J = np.zeros((n, n_params))
            J[:, mp] = jac[0, :, :n].t().cpu().numpy()
            LOG.info("Storing emulators in H matrix")
        indptr = np.concatenate([[0], np.cumsum(np.where(obs, n_params, 0))])
        cols = (n_params * np.flatnonzero(obs)[:, None] + np.arange(n_params)[None, :]).ravel()
        H = sp.csr_matrix((J[obs].astype(np.float32).ravel(), cols, indptr), shape=(n, n_params * n))
        LOG.info("\tDone!")
        return (H0.astype(np.float32), H)


def linearize_bands(x, bands, obs_mask, round_f32, h0=None, jac=None, want_jac=True):
    """Run kafka_gp_linearize for ``bands`` = [(EmulatorTable, mapper)] sharing
    one input dimension.  ``obs_mask``: uint8 (n_bands, ld) or (ld,) device tensor
    or None.  Returns device tensors h0 (n_bands, ld) and jac (n_bands, D, ld)."""
    lib = _lib.get_lib()
    nb = len(bands)
    D = bands[0][0].n_dim
    ld = x.ld
    arr = (_lib.GpBand * nb)()
    for b, (tab, mp) in enumerate(bands):
        if tab.n_dim != D or len(mp) != D:
            raise ValueError("bands linearised together must share the emulator input dimension")
        arr[b].table = tab.table.data_ptr()
        arr[b].n_train = tab.n_train
        arr[b].n_dim = D
        for d in range(D):
            arr[b].mapper[d] = int(mp[d])
    if obs_mask is not None:
        if obs_mask.dim() == 1:
            m = torch.zeros((nb, ld), dtype=torch.uint8, device=x.device)
            m[:, :obs_mask.numel()] = obs_mask[None, :]
            obs_mask = m
        assert obs_mask.shape == (nb, ld) and obs_mask.dtype == torch.uint8
    if h0 is None:
        h0 = torch.empty((nb, ld), dtype=torch.float64, device=x.device)
    if jac is None and want_jac:
        jac = torch.empty((nb, D, ld), dtype=torch.float64, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.kafka_gp_linearize(_lib.ptr(x.data), ld, x.n_pix, arr, nb, _lib.ptr(obs_mask),
                                          _lib.ptr(h0), _lib.ptr(jac), _lib.ROUND_F32 if round_f32 else 0,
                                          _lib.stream_ptr()), "kafka_gp_linearize")
    return h0, jac


# utils.py:181-219: full state -> emulator
create_prosail_observation_operator = ObservationOperator("create_prosail_observation_operator", "gp")
# utils.py:130-177: two-band TIP, spectral state mappers (utils.py:148-153)
create_nonlinear_observation_operator = ObservationOperator(
    "create_nonlinear_observation_operator", "gp",
    state_mappers=(np.array([0, 1, 6, 2]), np.array([3, 4, 6, 5])))
# utils.py:119-126 is unusable in the reference; this is the linear
# operator in the (H0, H) tuple form the solver understands (solvers.py:83-85).
create_linear_observation_operator = ObservationOperator("create_linear_observation_operator", "linear")


def block_diag(mats, format=None, dtype=None):
    """kafka/inference/utils.py:240-339 for equally shaped dense blocks: returns a
    device ``BlockDiag`` (use ``.to_scipy()`` for the sparse matrix).  A list of
    the SAME block object is stored once (broadcast)."""
    mats = list(mats) if not isinstance(mats, np.ndarray) else mats
    if len(mats) == 0:
        raise ValueError("block_diag needs at least one block")
    first = np.asarray(mats[0])
    if all(m is mats[0] for m in mats):
        bd = BlockDiag.from_block(first if dtype is None else first.astype(dtype), len(mats))
    else:
        blocks = np.asarray(mats, dtype=dtype)
        bd = BlockDiag.from_blocks(blocks)
    return bd if format is None else bd.to_scipy().asformat(format)
