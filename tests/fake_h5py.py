"""Minimal in-memory stand-in for h5py (absent from this image) so the HDF5 orchestration of the
drop-in (`load_*`, `save_hypergraph_to_h5`, `process_single_file`, `rebuild_*`) can be exercised.
Implements only what the reference / the drop-in touch: File(path, mode) context manager,
groups with `in`, `[]`, `[]=`, `del`, `create_group`, `require_group`, datasets with `[:]`, `.attrs`."""
import numpy as np

_FILES = {}


class _Dataset:
    def __init__(self, arr):
        self.arr = np.array(arr)

    def __getitem__(self, key):
        return self.arr[key]

    @property
    def shape(self):
        return self.arr.shape


class Group:
    def __init__(self):
        self.items = {}
        self.attrs = {}

    def __contains__(self, k):
        return k in self.items

    def __getitem__(self, k):
        return self.items[k]

    def __setitem__(self, k, v):
        if k in self.items:
            raise OSError(f"Unable to create dataset (name already exists): {k}")     # h5py behaviour
        self.items[k] = v if isinstance(v, (Group, _Dataset)) else _Dataset(v)

    def __delitem__(self, k):
        del self.items[k]

    def create_group(self, k):
        g = Group()
        self[k] = g
        return g

    def require_group(self, k):
        return self.items[k] if k in self.items else self.create_group(k)

    def keys(self):
        return self.items.keys()


class File(Group):
    def __new__(cls, path, mode="r"):
        if path not in _FILES:
            if mode == "r":
                raise FileNotFoundError(path)
            obj = super().__new__(cls)
            Group.__init__(obj)
            _FILES[path] = obj
        return _FILES[path]

    def __init__(self, path, mode="r"):
        pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def reset():
    _FILES.clear()
