"""Recipe that puts the UNMODIFIED reference package where the GPU box can import it
(TEST INFRASTRUCTURE; run by ``__graft_entry__.build()`` in the build container).

    python -m oracle.install_ref

``/root/reference`` exists only in the build container.  The CPU legs of ``bench.py``
time the reference's own code on the GPU box's host cores, so the package is
installed with pip, from a scratch copy (the source tree is read-only and setuptools
writes ``build/`` and ``*.egg-info`` next to ``setup.py``), into ``baseline/_ref``:
git-ignored, so no reference source enters the history, but part of the working tree
that ``gpurun`` ships.  Nothing is edited; ``oracle/ref_shim.py`` supplies the stub
modules the package needs at import.
"""
import os
import shutil
import subprocess
import sys
import tempfile

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SOURCE = "/root/reference"
TARGET = os.path.join(REPO, "baseline", "_ref")


def _newest(root):
    t = 0.0
    for d, _, files in os.walk(root):
        for f in files:
            if f.endswith(".py"):
                t = max(t, os.path.getmtime(os.path.join(d, f)))
    return t


def install(force=False):
    """Returns the target directory, or None where there is no source tree to install from
    (then whatever a previous build left in baseline/_ref is used)."""
    if not os.path.isdir(os.path.join(SOURCE, "kafka")):
        return None
    done = os.path.join(TARGET, "kafka", "linear_kf.py")
    if not force and os.path.exists(done) and os.path.getmtime(done) >= _newest(os.path.join(SOURCE, "kafka")):
        return TARGET
    tmp = tempfile.mkdtemp(prefix="kafka_ref_")
    try:
        src = os.path.join(tmp, "reference")
        shutil.copytree(SOURCE, src)
        shutil.rmtree(TARGET, ignore_errors=True)
        os.makedirs(TARGET)
        subprocess.check_call([sys.executable, "-m", "pip", "install", "--quiet", "--no-index", "--no-build-isolation",
                               "--no-deps", "--find-links", "/opt/wheelhouse", "--target", TARGET, src])
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    assert os.path.exists(done)
    return TARGET


if __name__ == "__main__":
    print(install(force="--force" in sys.argv))
