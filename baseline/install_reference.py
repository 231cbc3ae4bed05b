"""Offline install of the UNMODIFIED reference package into baseline/_ref (git-ignored; it travels to the GPU box with
the snapshot like the built .so files), so that ``bench.py --impl reference`` and bench.py's ``cuda_eager_baseline`` leg can
run the reference's own ``legged_gym/envs/pointfoot/point_foot.py`` there instead of the oracle port.

    python baseline/install_reference.py          (called by __graft_entry__.build() where /root/reference exists)

The reference is meant to be installed in editable mode (README.md:38-40 ``pip install -e .``): its ``setup.py`` calls
``find_packages()``, which skips every directory without an ``__init__.py`` -- envs/base, envs/pointfoot, scripts ... -- so a
plain ``pip install --target`` yields a package without point_foot.py.  The install therefore runs from a copy under /tmp
whose setup.py asks for ``find_namespace_packages(include=["legged_gym*"])`` instead; no source file of the package is
touched, and nothing but pip's output lands in baseline/_ref.  Dependencies (isaacgym: closed binary; rsl-rl: not in the
wheelhouse) are not installable: ``--no-deps``; oracle/ref_harness.py stubs them at import time.
"""
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TARGET = os.path.join(ROOT, "baseline", "_ref")
REFERENCE = os.environ.get("POINTFOOT_REFERENCE_SRC", "/root/reference")


def installed():
    return os.path.isfile(os.path.join(TARGET, "legged_gym", "envs", "pointfoot", "point_foot.py"))


def install(force=False):
    if installed() and not force:
        return True
    if not os.path.isfile(os.path.join(REFERENCE, "setup.py")):
        return False
    tmp = tempfile.mkdtemp(prefix="pointfoot_ref_")
    try:
        src = os.path.join(tmp, "reference")
        shutil.copytree(REFERENCE, src, ignore=shutil.ignore_patterns("resources", "read_bag", ".git", "logs"))
        setup = os.path.join(src, "setup.py")
        text = open(setup).read()
        text = text.replace("from setuptools import find_packages", "from setuptools import find_namespace_packages")
        text = text.replace("packages=find_packages()", 'packages=find_namespace_packages(include=["legged_gym*"])')
        open(setup, "w").write(text)
        shutil.rmtree(TARGET, ignore_errors=True)
        subprocess.check_call([sys.executable, "-m", "pip", "install", "-q", "--no-index", "--no-build-isolation", "--no-deps",
                               "--find-links", "/opt/wheelhouse", "--target", TARGET, src])
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return installed()


if __name__ == "__main__":
    ok = install(force="--force" in sys.argv)
    print("baseline/_ref:", "installed" if ok else "reference not available here")
