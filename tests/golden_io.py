"""Rebuild inputs from the committed fixtures in tests/golden (made by
oracle/make_golden.py from the unmodified reference)."""
import datetime as dt
import os

import numpy as np

from kafka_inferenceengine_b200 import synthetic as syn

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)


def barrax_mask():
    d = load("barrax_mask")
    ny, nx = d["shape"]
    return np.unpackbits(d["bits"])[:ny * nx].astype(bool).reshape(ny, nx)


def emulators(d, cls, prefix="gp"):
    out = []
    b = 0
    while "%s_inputs_%d" % (prefix, b) in d:
        out.append(cls(d["%s_inputs_%d" % (prefix, b)], d["%s_targets_%d" % (prefix, b)],
                       d["%s_theta_%d" % (prefix, b)]))
        b += 1
    return out


class FixtureObservations(object):
    """Observation collaborator (Sentinel2_Observations.py:148-185 contract)
    replaying the rasters stored in a fixture."""

    def __init__(self, d, emulator_objs):
        self.dates = [dt.datetime.fromordinal(int(o)) for o in d["obs_dates"]]
        self.bands_per_observation = {}
        self._data = {}
        for di, date in enumerate(self.dates):
            bands = []
            b = 0
            while "obs_y_%d_%d" % (di, b) in d:
                inv_var = d["obs_invvar_%d_%d" % (di, b)]
                bands.append(syn.BandData(d["obs_y_%d_%d" % (di, b)],
                                          syn.inverse_variance_matrix(inv_var),
                                          d["obs_mask_%d_%d" % (di, b)], {"band": b},
                                          emulator_objs[b] if len(emulator_objs) > 1
                                          else emulator_objs[0]))
                b += 1
            self._data[date] = bands
            self.bands_per_observation[date] = b

    def get_band_data(self, date, band):
        return self._data[date][band]


def grid(d):
    return [dt.datetime.fromordinal(int(o)) for o in d["grid"]]


def compacted_bands(obs, state_mask):
    """{date: [(y, w, m) per band]} with gated weights, via the oracle gather."""
    from oracle import kafka_cpu as kc
    out = {}
    for date in obs.dates:
        rows = []
        for b in range(obs.bands_per_observation[date]):
            bd = obs.get_band_data(date, b)
            y, r, m = kc.gather_band(bd.observations, bd.uncertainty.diagonal(), bd.mask, state_mask)
            rows.append((y, kc.gate_inverse_variance(r, m), m))
        out[date] = rows
    return out


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), 1e-300)
