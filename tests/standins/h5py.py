"""TEST STAND-IN for h5py (absent from this image).  Not product code.

The reference's JRL demos import h5py at module level and use exactly: `h5py.File(name)` / `h5py.File(name, "w")`,
`f.create_dataset(key, data=array)`, `f[key]` (wrapped in `np.asarray`) and `f.close()`
(demo_2d_LBM_JRL_fluid_cylinder.py:266-268, demo_2d_LBM_JRL_C_cylinder.py:209-213, 251-254, 276, 409).  This module
provides those calls with an .npz container written AT THE REQUESTED PATH, so the demos' own `os.path.exists()`
tests on the `.h5` names behave as with real HDF5.  tests/conftest.py and oracle/run_demo.py put this directory on
sys.path only when the real h5py cannot be imported.  jrl_io.load_fields reads the same container."""
import os

import numpy as np

__jrl_standin__ = True


class File:
    def __init__(self, name, mode="r", **kw):
        self.filename = str(name)
        self.mode = mode
        self._data = {}
        if mode in ("r", "r+", "a") and os.path.exists(self.filename):
            with np.load(self.filename, allow_pickle=False) as z:
                self._data = {k: z[k] for k in z.files}
        elif mode in ("r", "r+"):
            raise FileNotFoundError(self.filename)
        self._dirty = False

    def create_dataset(self, name, data=None, **kw):
        self._data[name] = np.asarray(data)
        self._dirty = True
        return self._data[name]

    def __getitem__(self, name):
        return self._data[name]

    def __contains__(self, name):
        return name in self._data

    def keys(self):
        return self._data.keys()

    def close(self):
        if self._dirty:
            tmp = self.filename + ".tmp"
            with open(tmp, "wb") as fo:
                np.savez(fo, **self._data)
            os.replace(tmp, self.filename)
            self._dirty = False

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
