"""Offline install of the UNMODIFIED reference package into baseline/_ref (git-ignored; it travels to the GPU box with the
snapshot, /root/reference does not).

    python baseline/install_ref.py            (called by __graft_entry__.build() when /root/reference exists)

The reference declares poetry-core as its build backend, which is not in this image (`pip install /root/reference` stops with
"No module named 'poetry'").  The package is pure Python, so the install is done from a copy under /tmp (the source tree is
read-only) whose build metadata is replaced by a three-line setuptools `setup.py`; the `tianshou/` sources are installed
byte for byte:

    pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target baseline/_ref /tmp/<copy>

Its missing third-party imports (gymnasium, pettingzoo, overrides, sensai, h5py, deepdiff, matplotlib) are served by the
stand-ins in oracle/shims/ at run time (bench.py --impl reference), exactly as tests/golden/make_golden*.py do.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
TARGET = os.path.join(HERE, "_ref")


def installed() -> bool:
    return os.path.isfile(os.path.join(TARGET, "tianshou", "__init__.py"))


def install(force: bool = False) -> bool:
    """Returns True when baseline/_ref/tianshou is in place afterwards."""
    if installed() and not force:
        return True
    if not os.path.isdir(os.path.join(REF, "tianshou")):
        return False
    tmp = tempfile.mkdtemp(prefix="tsm_ref_")
    try:
        src = os.path.join(tmp, "src")
        os.makedirs(src)
        shutil.copytree(os.path.join(REF, "tianshou"), os.path.join(src, "tianshou"))
        with open(os.path.join(src, "setup.py"), "w") as f:
            f.write("from setuptools import setup, find_packages\n"
                    "setup(name='tianshou', version='2.0.0b1', packages=find_packages(include=['tianshou', 'tianshou.*']),\n"
                    "      include_package_data=True)\n")
        shutil.rmtree(TARGET, ignore_errors=True)
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--find-links", "/opt/wheelhouse",
               "--target", TARGET, src]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            print(r.stdout[-2000:], r.stderr[-2000:])
            return False
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return installed()


if __name__ == "__main__":
    ok = install(force="--force" in sys.argv)
    print("baseline/_ref:", "installed" if ok else "unavailable")
    sys.exit(0 if ok else 1)
