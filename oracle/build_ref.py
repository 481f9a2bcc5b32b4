"""Stage the UNTOUCHED reference package for timing on the GPU box (test / benchmark infrastructure, not product code).

    python oracle/build_ref.py            # build container only: needs /root/reference

Copies /root/reference/package/src/openflash (pure Python, no build step) byte for byte into oracle/_ref/openflash and
writes two import stubs beside it - oracle/_ref/matplotlib (the reference imports pyplot at module scope for plotting
only) and oracle/_ref/xarray (a ~20-line Dataset / DataArray stand-in: Results only needs named arrays) - because
neither package is installed in this image.  oracle/_ref/ is git-ignored (reference sources never enter
the history) but NOT gpurun-ignored, so it travels to the GPU box like the built .so files; `bench.py --impl reference`
and the cpu_baseline leg import it from there (kind "reference") and fall back to the oracle port (kind "port") when it
is absent.  Nothing under openflash_b200/ may import it.
"""
import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
SRC = "/root/reference/package/src/openflash"

XARRAY_STUB = '''"""Minimal stand-in for xarray (not installed): what openflash.results needs - named arrays with dims."""
import numpy as np


class _Var:
    def __init__(self, dims, data):
        self.dims = tuple(dims)
        self.data = np.asarray(data)
        self.values = self.data
        self.shape = self.data.shape


class DataArray(_Var):
    def __init__(self, data, dims=None, coords=None):
        super().__init__(dims or (), data)
        self.coords = coords or {}


class Dataset:
    def __init__(self, data_vars=None, coords=None, attrs=None):
        self.coords = dict(coords or {})
        self.attrs = attrs or {}
        self.data_vars = {}
        for k, v in (data_vars or {}).items():
            self[k] = v

    def __setitem__(self, k, v):
        self.data_vars[k] = _Var(*v) if isinstance(v, tuple) else v

    def __getitem__(self, k):
        return self.data_vars[k]

    def __contains__(self, k):
        return k in self.data_vars
'''

MPL_STUB = '''"""Stand-in for matplotlib (not installed): the reference imports pyplot at module scope, never on the hot path."""
from unittest.mock import MagicMock
import sys
pyplot = MagicMock(name="pyplot")
sys.modules[__name__ + ".pyplot"] = pyplot
'''


def build(verbose=True):
    if not os.path.isdir(SRC):
        if verbose:
            print("oracle/build_ref.py: /root/reference absent - keeping whatever oracle/_ref holds")
        return os.path.isdir(os.path.join(DST, "openflash"))
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    os.makedirs(DST)
    shutil.copytree(SRC, os.path.join(DST, "openflash"), ignore=shutil.ignore_patterns("__pycache__"))
    for mod, text in (("xarray", XARRAY_STUB), ("matplotlib", MPL_STUB)):
        os.makedirs(os.path.join(DST, mod))
        with open(os.path.join(DST, mod, "__init__.py"), "w") as f:
            f.write(text)
    digest = hashlib.sha256()
    for name in sorted(os.listdir(os.path.join(DST, "openflash"))):
        if name.endswith(".py"):
            with open(os.path.join(DST, "openflash", name), "rb") as f:
                digest.update(f.read())
    with open(os.path.join(DST, "SOURCE.txt"), "w") as f:
        f.write(f"copied from {SRC} (unmodified); sha256 of the concatenated *.py = {digest.hexdigest()}\n")
    if verbose:
        print("oracle/_ref staged:", digest.hexdigest()[:16])
    return True


if __name__ == "__main__":
    sys.exit(0 if build() else 1)
