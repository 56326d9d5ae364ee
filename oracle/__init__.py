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
``ref_shim.py``    imports the UNMODIFIED reference behind stub modules, from
                   ``/root/reference`` (build container) or from ``baseline/_ref``
                   (the pip-installed copy that travels to the GPU box).
``install_ref.py`` the recipe that makes ``baseline/_ref`` (run by
                   ``__graft_entry__.build()`` where the source tree exists).
``ref_runner.py``  times the unmodified reference on chunks of the benchmark
                   workload (``bench.py --impl reference`` / ``cpu_baseline``).
``make_golden.py`` runs the unmodified reference under the shim and writes the
                   fixtures committed under ``tests/golden/``.
``synthetic.py``   the benchmark's seeded workload restated for the CPU legs,
                   which must not import the product package.
"""
