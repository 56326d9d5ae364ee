"""Seeded synthetic workload of the benchmark for the CPU legs (TEST INFRASTRUCTURE).

The ``--impl reference`` arm and the ``cpu_baseline`` leg of ``bench.py`` must run
without the product package (no ``kafka_inferenceengine_b200`` import, no CUDA
library mapped), so the few constants and generators they need are restated
here, from the reference's own driver: the PROSAIL prior of ``kafka_test_S2.py``
lines 84-94 and the data contract of ``kafka/input_output/Sentinel2_Observations.py``
lines 166-179.  ``tests/test_host_logic.py`` checks them against the product
package's ``synthetic`` module.
"""
import numpy as np

from .gp import make_synthetic_gp

# kafka_test_S2.py:84-94 (SAILPrior): mean and standard deviation of the transformed PROSAIL state
SAIL_MEAN = np.array([2.1, np.exp(-60. / 100.), np.exp(-7.0 / 100.), 0.1,
                      np.exp(-50 * 0.0176), np.exp(-100. * 0.002),
                      np.exp(-4. / 2.), 70. / 90., 0.5, 0.9])
SAIL_SIGMA = np.array([0.01, 0.2, 0.01, 0.05, 0.01, 0.01, 0.50, 0.1, 0.1, 0.1])
# kafka_test_S2.py:136-137
S2_PARAMETERS = ['n', 'cab', 'car', 'cbrown', 'cw', 'cm', 'lai', 'ala', 'bsoil', 'psoil']


def make_emulators(n_bands=10, n_train=250, seed=1234):
    """The benchmark's emulators (same seed and draw order as bench.make_emulators)."""
    rng = np.random.default_rng(seed)
    return [make_synthetic_gp(rng, SAIL_MEAN, SAIL_SIGMA, n_train=n_train) for _ in range(n_bands)]


def sail_prior_block():
    """Inverse covariance block of the prior, float32 like the driver builds it
    (kafka_test_S2.py:93-94)."""
    return np.diag(1.0 / SAIL_SIGMA ** 2).astype(np.float32)


def synth_observations(gps, n_pix, seed):
    """(rho, inverse variance, mask), each (n_bands, n_pix): x_true ~ N(prior),
    reflectance = emulator(x_true) with 5 % noise, observed with probability 0.9
    where rho > 0, inverse variance 1/(0.05 rho)^2 and inf where masked
    (Sentinel2_Observations.py:168-178)."""
    rng = np.random.default_rng(seed)
    xt = SAIL_MEAN + SAIL_SIGMA * rng.standard_normal((n_pix, SAIL_MEAN.size))
    r = np.array([g.predict(xt, do_unc=False)[0] for g in gps])
    r = r * (1 + 0.05 * rng.standard_normal(r.shape))
    mk = (r > 0) & (rng.random(r.shape) < 0.9)
    r = np.where(mk, r, 0.0)
    with np.errstate(divide="ignore"):
        iv = np.where(mk, 1.0 / (0.05 * r) ** 2, np.inf)
    return r, iv, mk
