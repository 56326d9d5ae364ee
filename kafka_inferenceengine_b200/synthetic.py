"""Seeded synthetic inputs for the BASELINE.json configurations.

The reference ships one (unfinished) synthetic observation source,
``BHRObservationsTest`` (kafka/input_output/observations.py:313-333); everything
else reads satellite granules through GDAL, which is out of scope here.  This
module provides the same duck-typed collaborators (``dates``,
``bands_per_observation``, ``get_band_data`` -> namedtuple with
``observations, uncertainty, mask, metadata, emulator``; data contract of
kafka/input_output/Sentinel2_Observations.py:148-185) filled with seeded
synthetic data.  No network, no datasets: random-init
emulators of the right shape, synthetic reflectances.

Pure numpy.  The forward model used to synthesise reflectances is passed in as
``predict(emulator, X) -> mean`` so that tests can use the CPU oracle and the
benchmark the CUDA emulator kernel.
"""
import datetime as dt
from collections import namedtuple

import numpy as np
import scipy.sparse as sp

BandData = namedtuple("BandData", "observations uncertainty mask metadata emulator")

# kafka_test_S2.py:136-137
S2_PARAMETERS = ['n', 'cab', 'car', 'cbrown', 'cw', 'cm', 'lai', 'ala', 'bsoil', 'psoil']
# kafka_test.py:166-167 (TIP)
TIP_PARAMETERS = ['w_vis', 'x_vis', 'a_vis', 'w_nir', 'x_nir', 'a_nir', 'TeLAI']

# kafka_test_S2.py:84-94
SAIL_MEAN = np.array([2.1, np.exp(-60. / 100.), np.exp(-7.0 / 100.), 0.1,
                      np.exp(-50 * 0.0176), np.exp(-100. * 0.002),
                      np.exp(-4. / 2.), 70. / 90., 0.5, 0.9])
SAIL_SIGMA = np.array([0.01, 0.2, 0.01, 0.05, 0.01, 0.01, 0.50, 0.1, 0.1, 0.1])
# kafka/inference/kf_tools.py:108-109
TIP_SIGMA = np.array([0.12, 0.7, 0.0959, 0.15, 1.5, 0.2, 0.5])
TIP_MEAN = np.array([0.17, 1.0, 0.1, 0.7, 2.0, 0.18, np.exp(-0.5 * 1.5)])


class EmulatorData(object):
    """Minimal GP container with ``gp_emulator.GaussianProcess``'s field names
    (``inputs`` (T,D), ``theta`` = LOG hyper-parameters (D+2), ``invQt`` (T))."""

    def __init__(self, inputs, targets, theta):
        self.inputs = np.ascontiguousarray(inputs, dtype=np.float64)
        self.targets = np.ascontiguousarray(targets, dtype=np.float64)
        self.theta = np.ascontiguousarray(theta, dtype=np.float64)
        self.n, self.D = self.inputs.shape
        e = np.exp(self.theta)
        Xs = self.inputs * np.sqrt(e[:self.D])
        d2 = ((Xs[:, None, :] - Xs[None, :, :]) ** 2).sum(axis=2)
        self.Q = e[self.D] * np.exp(-0.5 * d2) + e[self.D + 1] * np.eye(self.n)
        self.invQt = np.linalg.solve(self.Q, self.targets)

    @property
    def invQ(self):
        if not hasattr(self, "_invQ"):
            self._invQ = np.linalg.inv(self.Q)
        return self._invQ


def make_emulator(rng, mean, sigma, n_train=250, w_scale=0.03, sigma_f2=0.05,
                  nugget=1e-6, offset=0.2, amp=0.1, cls=EmulatorData):
    """Smooth synthetic emulator around a prior: inputs ~ N(mean,(2 sigma)^2),
    targets offset+amp*tanh(z.w) with z the sigma-normalised offset, inverse
    squared length scales 1/(4 sigma)^2."""
    mean = np.asarray(mean, dtype=np.float64)
    sigma = np.asarray(sigma, dtype=np.float64)
    D = mean.shape[0]
    X = mean + 2.0 * sigma * rng.standard_normal((n_train, D))
    w = rng.standard_normal(D) * w_scale
    t = offset + amp * np.tanh(((X - mean) / sigma) @ w)
    theta = np.concatenate([np.log(1.0 / (4.0 * sigma) ** 2),
                            [np.log(sigma_f2), np.log(nugget)]])
    return cls(X, t, theta)


def inverse_variance_matrix(inv_var_raster):
    """The (ny*nx)^2 sparse diagonal the reference's readers hand over
    (Sentinel2_Observations.py:175-179): 1/sigma^2, ``inf`` where masked."""
    return sp.diags(inv_var_raster.ravel(), format="csr")


class SyntheticObservations(object):
    """S2-style observation source (Sentinel2_Observations.py:82-185 contract)."""

    def __init__(self, state_mask, dates, emulators, x_true, predict, rng,
                 state_mappers=None, rel_sigma=0.05, p_observed=0.9,
                 min_sigma=None, sparse_uncertainty=True):
        self.state_mask = np.asarray(state_mask, dtype=bool)
        self.dates = list(dates)
        self.emulators = list(emulators)
        self.bands_per_observation = {d: len(self.emulators) for d in self.dates}
        self._data = {}
        n_params = x_true.shape[1]
        ny, nx = self.state_mask.shape
        for d in self.dates:
            bands = []
            for b, gp in enumerate(self.emulators):
                cols = np.arange(n_params) if state_mappers is None else state_mappers[b]
                rho_c = predict(gp, x_true[:, cols])
                sig = rel_sigma * np.abs(rho_c)
                if min_sigma is not None:
                    sig = np.maximum(sig, min_sigma)
                rho_c = rho_c + sig * rng.standard_normal(rho_c.shape[0])
                rho = np.zeros((ny, nx))
                rho[self.state_mask] = rho_c
                mask = (rho > 0) & (rng.random((ny, nx)) < p_observed)      # :168
                rho = np.where(mask, rho, 0.0)
                r = rel_sigma * rho                                        # :174
                if min_sigma is not None:
                    r = np.where(mask, np.maximum(r, min_sigma), 0.0)
                with np.errstate(divide="ignore"):
                    inv_var = 1.0 / r ** 2                                 # :178 (inf where masked)
                unc = inverse_variance_matrix(inv_var) if sparse_uncertainty else inv_var
                bands.append(BandData(rho, unc, mask, {"band": b}, gp))
            self._data[d] = bands

    def get_band_data(self, date, band):
        return self._data[date][band]


class LinearBandOperator(object):
    """A linear observation operator in the (H0, H) tuple form the solver
    understands (solvers.py:83-85): row ``h`` (length n_params) per pixel plus a
    constant offset, i.e. H0 = h.x_prev + c.  Stands in for the reference's dead
    ``create_linear_observation_operator`` (utils.py:119-126)."""

    def __init__(self, row, offset=0.0):
        self.row = np.asarray(row, dtype=np.float64)
        self.offset = float(offset)


class LinearObservations(object):
    """Config-1 source: one band observing a linear combination of the state."""

    def __init__(self, state_mask, dates, operator, x_true, rng, sigma_obs=0.05,
                 p_observed=1.0):
        self.state_mask = np.asarray(state_mask, dtype=bool)
        self.dates = list(dates)
        self.bands_per_observation = {d: 1 for d in self.dates}
        self._data = {}
        ny, nx = self.state_mask.shape
        for d in self.dates:
            yc = x_true @ operator.row + operator.offset
            yc = yc + sigma_obs * rng.standard_normal(yc.shape[0])
            y = np.zeros((ny, nx))
            y[self.state_mask] = yc
            mask = rng.random((ny, nx)) < p_observed
            with np.errstate(divide="ignore"):
                inv_var = np.where(mask, 1.0 / sigma_obs ** 2, np.inf)
            self._data[d] = [BandData(np.where(mask, y, 0.0),
                                      inverse_variance_matrix(inv_var), mask,
                                      {"band": 0}, operator)]

    def get_band_data(self, date, band):
        return self._data[date][band]


class MemoryOutput(object):
    """Keeps what ``KafkaOutput.dump_data`` would write
    (input_output/observations.py:354-394): per parameter the mean and
    1/sqrt(diag(P^-1)), here as compacted vectors instead of GeoTIFFs."""

    def __init__(self, parameter_list):
        self.parameter_list = parameter_list
        self.output = {}

    def dump_data(self, timestep, x_analysis, P_analysis, P_analysis_inv,
                  state_mask, n_params):
        x = np.asarray(x_analysis, dtype=np.float64)
        d = np.asarray(P_analysis_inv.diagonal(), dtype=np.float64)
        self.output[timestep] = {"x": x.copy(), "diag": d.copy()}


class BroadcastPrior(object):
    """``SAILPrior``-style prior (kafka_test_S2.py:77-118): one mean vector and
    one inverse covariance block repeated for every state pixel.
    ``make_block_diag(block, n)`` builds the matrix; ``make_vector(mean, n)``
    (optional) the mean vector -- by default a host ``np.tile`` like the
    reference; both results are cached, the prior does not depend on the date."""

    def __init__(self, mean, inv_cov_block, n_pixels, make_block_diag, make_vector=None):
        self.mean = np.asarray(mean, dtype=np.float64)
        self.inv_cov = np.asarray(inv_cov_block, dtype=np.float64)
        self.n_pixels = int(n_pixels)
        self._make = make_block_diag
        self._make_vector = make_vector or (lambda m, n: np.tile(m, n))
        self._cache = None

    def process_prior(self, time, inv_cov=True):
        if self._cache is None:
            self._cache = (self._make_vector(self.mean, self.n_pixels), self._make(self.inv_cov, self.n_pixels))
        return self._cache


def day_grid(start, n, step=1):
    return [start + dt.timedelta(days=i * step) for i in range(n)]
