"""Test stand-in for ``h5py`` (absent from this image): the handful of calls ``search/search.py:51-63``,
``eval/eval.py`` and ``learnedmetricindex_b200.results`` make -- ``File(path, mode)`` as a context manager or with
``close()``, ``f.attrs[...]``, ``f.create_dataset(name, shape, dtype=...)[:] = a``, ``f[name]``, ``name in f`` --
over one ``numpy.savez`` archive stored AT the given path (whatever its suffix).  Test infrastructure only: it is
put on ``sys.path`` by ``tests/test_results_vs_reference_eval.py`` and by nothing else."""
import numpy as np

_ATTR = "__attr__"


class File:
    def __init__(self, path, mode="r"):
        self.path, self.mode = str(path), mode
        self.attrs, self._data, self._closed = {}, {}, False
        if mode.startswith("r"):
            with open(self.path, "rb") as fh:
                z = np.load(fh, allow_pickle=False)
                for key in z.files:
                    a = z[key]
                    if key.startswith(_ATTR):
                        self.attrs[key[len(_ATTR):]] = a.item() if a.ndim == 0 else a
                    else:
                        self._data[key] = a

    def create_dataset(self, name, shape, dtype=None):
        self._data[name] = np.empty(shape, dtype=dtype)
        return self._data[name]

    def __getitem__(self, name):
        return self._data[name]

    def __contains__(self, name):
        return name in self._data

    def close(self):
        if self._closed:
            return
        self._closed = True
        if self.mode.startswith("w"):
            payload = dict(self._data)
            payload.update({_ATTR + k: np.asarray(v) for k, v in self.attrs.items()})
            with open(self.path, "wb") as fh:
                np.savez(fh, **payload)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False
