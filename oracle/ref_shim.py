"""Import the UNMODIFIED reference behind stub modules (TEST INFRASTRUCTURE).

The package is taken from ``/root/reference`` (the build container) or, where
that path does not exist (the GPU box), from ``baseline/_ref``: the same files,
installed there by ``oracle/install_ref.py`` (``pip install --target``) when
``__graft_entry__.build()`` runs in the build container; that directory is
git-ignored and travels with the working tree.  Used by ``oracle/make_golden.py``,
by the validation tests that compare the oracle with the reference (they skip
when neither location exists) and by ``oracle/ref_runner.py``, which times the
reference for ``bench.py``'s CPU legs.  Nothing in the product package, the
``-m gpu`` parity tests or ``smoke()`` needs it.

The reference does not import as-is here (no gdal / matplotlib /
BRDF_descriptors / SIAC; NumPy 2 and SciPy 1.18 dropped names it uses).  Recipe
(SURVEY.md section 8c), no reference file is edited or copied:

1. empty stub modules for ``gdal, osr, ogr, osgeo.*, matplotlib(.pyplot),
   BRDF_descriptors, SIAC(.kernels)``           (imports at inference/utils.py:35-38,
   solvers.py:25, input_output/observations.py:47,54)
2. ``np.int = int``                                        (inference/utils.py:310)
3. ``scipy.sparse.sputils`` and friends aliased to their private successors
                                                           (inference/utils.py:286-293)
4. ``np.vstack`` accepting a ``set``                        (inference/utils.py:72)
"""
import os
import sys
import types

_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# the source tree in the build container; elsewhere (the GPU box) the unmodified package as
# ``oracle/install_ref.py`` installed it with pip into the git-ignored baseline/_ref
_CANDIDATES = ("/root/reference", os.path.join(_REPO, "baseline", "_ref"))
REFERENCE_ROOT = next((c for c in _CANDIDATES if os.path.isdir(os.path.join(c, "kafka"))), _CANDIDATES[0])


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "kafka"))


def _stub(name, **attrs):
    mod = sys.modules.get(name)
    if mod is None:
        mod = types.ModuleType(name)
        sys.modules[name] = mod
    for k, v in attrs.items():
        setattr(mod, k, v)
    return mod


def load():
    """Return the reference's ``kafka`` package (imported from REFERENCE_ROOT)."""
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    import numpy as np
    import scipy.sparse as sp

    for name in ("gdal", "osr", "ogr", "osgeo", "osgeo.gdal", "osgeo.osr",
                 "osgeo.ogr", "matplotlib", "matplotlib.pyplot", "SIAC",
                 "SIAC.kernels", "netCDF4", "dask", "dask.distributed"):
        _stub(name)
    sys.modules["osgeo"].gdal = sys.modules["osgeo.gdal"]
    sys.modules["osgeo"].osr = sys.modules["osgeo.osr"]
    sys.modules["osgeo"].ogr = sys.modules["osgeo.ogr"]
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    _stub("BRDF_descriptors", RetrieveBRDFDescriptors=object)
    _stub("SIAC.kernels", Kernels=object)
    sys.modules["SIAC"].kernels = sys.modules["SIAC.kernels"]

    if not hasattr(np, "int"):
        np.int = int
    if not hasattr(np, "bool"):
        np.bool = bool
    import scipy.sparse._sputils as _spu
    _stub("scipy.sparse.sputils", upcast=_spu.upcast,
          get_index_dtype=_spu.get_index_dtype)
    sp.sputils = sys.modules["scipy.sparse.sputils"]
    for short, priv, cls in (("csr", "_csr", "csr_matrix"), ("csc", "_csc", "csc_matrix"),
                             ("bsr", "_bsr", "bsr_matrix"), ("coo", "_coo", "coo_matrix"),
                             ("dia", "_dia", "dia_matrix")):
        m = __import__("scipy.sparse." + priv, fromlist=[cls])
        _stub("scipy.sparse." + short, **{cls: getattr(m, cls)})

    if not getattr(np.vstack, "_accepts_set", False):
        _orig = np.vstack

        def vstack(tup, *a, **k):
            if isinstance(tup, (set, frozenset)):
                tup = list(tup)
            return _orig(tup, *a, **k)
        vstack._accepts_set = True
        np.vstack = vstack

    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import kafka  # noqa: the reference package
    assert os.path.realpath(kafka.__file__).startswith(os.path.realpath(REFERENCE_ROOT))
    return kafka
