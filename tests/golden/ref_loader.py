"""Import the UNTOUCHED reference (/root/reference) under two import stubs.

Build-container only: /root/reference does not exist on the GPU box, so nothing
that runs there may import this module.  It is used by make_golden.py (fixture
generation) and by the optional live cross-checks in tests/ (skipped when the
reference tree is absent).  matplotlib and xarray are not installed in this
image; the reference imports them at module scope for plotting / packaging
only, so a MagicMock and a ~20-line stand-in are enough.
"""
import os
import sys
import types
from unittest.mock import MagicMock

import numpy as np

REF_SRC = "/root/reference/package/src"


def available():
    return os.path.isdir(REF_SRC)


def load():
    if "openflash" in sys.modules:
        return sys.modules["openflash"]
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        plt = MagicMock(name="pyplot")
        mpl.pyplot = plt
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = plt
    if "xarray" not in sys.modules:
        xr = types.ModuleType("xarray")

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

        xr.Dataset, xr.DataArray = Dataset, DataArray
        sys.modules["xarray"] = xr
    sys.path.insert(0, REF_SRC)
    import openflash
    return openflash
