"""Materialise the unmodified reference package under ``baseline/_ref`` (git-ignored, shipped with the tree).

The reference is pure Python (``/root/reference/python/setup.py`` declares no extension modules), so the offline
install is a wheel build from a scratch copy of the source tree (the build writes ``build/`` and ``*.egg-info`` next
to ``setup.py`` and ``/root/reference`` is read-only).  Dependency resolution is skipped (``--no-deps``): several of
the reference's requirements (mpi4py, MNN, paho-mqtt, boto3, wget, h5py) have no wheel offline; the reference arm
stubs exactly those at import time.  When pip is unusable the package directory is copied as is, which yields the
same importable tree.

``ensure_reference()`` is idempotent and cheap when the package is already there; it is called by
``__graft_entry__.build()`` (so the directory exists before the tree is pushed to a GPU box) and by
``reference_arm.import_reference()`` (so a box that still has the reference sources can self-heal).
"""
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
SOURCES = ("/root/reference/python", os.environ.get("FEDML_REFERENCE_SRC", ""))


def _source_dir():
    for s in SOURCES:
        if s and os.path.isfile(os.path.join(s, "fedml", "__init__.py")):
            return s
    return None


def ensure_reference(verbose=False):
    """Return the path that must be put on ``sys.path`` to import the reference, installing it if needed.
    Raises ImportError with a one-line reason when neither an installed copy nor the sources exist."""
    marker = os.path.join(REF, "fedml", "__init__.py")
    if os.path.isfile(marker):
        return REF
    src = _source_dir()
    if src is None:
        raise ImportError("baseline/_ref/fedml missing and no reference sources at /root/reference/python "
                          "(run __graft_entry__.build() where the reference is mounted)")
    os.makedirs(REF, exist_ok=True)
    how = "pip"
    tmp = tempfile.mkdtemp(prefix="fedml_ref_")
    try:
        work = os.path.join(tmp, "src")
        shutil.copytree(src, work, symlinks=True)
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--quiet",
               "--find-links", "/opt/wheelhouse", "--target", REF, work]
        try:
            r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
            ok = r.returncode == 0 and os.path.isfile(marker)
            if verbose and not ok:
                print(r.stdout[-2000:])
        except Exception:  # noqa: BLE001
            ok = False
        if not ok:
            how = "copytree"
            if os.path.isdir(os.path.join(REF, "fedml")):
                shutil.rmtree(os.path.join(REF, "fedml"))
            shutil.copytree(os.path.join(src, "fedml"), os.path.join(REF, "fedml"), symlinks=True)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    with open(os.path.join(REF, "INSTALLED_BY"), "w") as f:
        f.write(f"{how} from {src}\n")
    if verbose:
        print(f"reference installed into {REF} ({how})")
    return REF


if __name__ == "__main__":
    print(ensure_reference(verbose=True))
