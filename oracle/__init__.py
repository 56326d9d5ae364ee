"""CPU oracle for the KaFKA per-pixel assimilation path.

THIS PACKAGE IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import it.  Nothing under
``kafka_inferenceengine_b200/`` imports it, and the product path raises when
the CUDA library is missing instead of falling back to this code.

Contents
--------
``gp.py``          stand-in for the un-vendored ``gp_emulator.GaussianProcess``
                   (ARD squared-exponential GP: predict / variance / Jacobian /
                   Hessian).  PARITY UNPINNED for this piece: the package is not
                   in ``/root/reference``, has no pinned version and no golden
                   vectors; see the module header.
``kafka_cpu.py``   numpy/scipy restatement of the reference's filter equations:
                   oracle A' (scipy sparse + float32 SuperLU, as the reference
                   does it) and oracle B (fp64, one dense block per pixel).
``ref_shim.py``    imports the UNMODIFIED reference from ``/root/reference``
                   behind stub modules (works only where that path exists, i.e.
                   in the build container, never on the GPU box).
``make_golden.py`` runs the unmodified reference under the shim and writes the
                   fixtures committed under ``tests/golden/``.
``synthetic.py``   seeded synthetic inputs for the BASELINE.json configs.
"""
