"""Slide containers for the hypergraph pipeline.

The reference keeps every slide in one HDF5 file (``wsi/``, ``tma/``, and after preprocessing
``hypergraph/``; layout at reference preprocess_hypergraph.py:474-511, README.md:58-73) and talks to
it through h5py.  ``open_store`` hands the I/O functions of this package an object with that small
h5py surface -- groups with ``in`` / ``[]`` / ``[]=`` / ``del`` / ``require_group``, datasets with
``[:]``, ``.attrs`` -- for two back ends:

* any other path: ``h5py.File`` itself (h5py is imported lazily; the reference imports it at module
  scope, :15);
* ``*.npz``: ``NpzFile``, the same tree flattened to ``"group/sub/name"`` keys of one numpy archive
  (attributes as one JSON document under ``__attrs__``).  It needs nothing beyond numpy, is written
  atomically on close, and ``export_to_h5`` / ``import_from_h5`` convert losslessly in both directions
  when h5py is available.

Large slides: the reference stores the dense N x N kernel under ``hypergraph/similarity/wsi_internal``
(:501), 10 GB at N = 50k.  Beyond the dense limit this package stores the top-k lists of that kernel
(``similarity/wsi_topk_indices`` / ``wsi_topk_values``, k <= 32) and its streamed statistics instead.
"""
from __future__ import annotations

import json
import os
import tempfile
from typing import Dict, Optional

import numpy as np

ATTRS_KEY = "__attrs__"


class Dataset:
    """Read-only view of one stored array (what ``group[name]`` returns for a leaf).  Arrays of a file
    opened from disk are read on first access only: loading ``tma/features`` does not pull a stored
    N x N similarity matrix through memory."""

    def __init__(self, arr: Optional[np.ndarray] = None, source=None, key: Optional[str] = None):
        self._data = arr
        self._source, self._key = source, key          # lazily read member `key`.npy of an open archive

    @property
    def loaded(self) -> bool:
        return self._data is not None

    @property
    def _arr(self) -> np.ndarray:
        if self._data is None:
            self._data = self._source[self._key]
        return self._data

    def _header(self):
        if self._data is not None:
            return self._data.shape, self._data.dtype
        with self._source.zip.open(self._key + ".npy") as fh:
            version = np.lib.format.read_magic(fh)
            reader = np.lib.format.read_array_header_1_0 if version == (1, 0) else np.lib.format.read_array_header_2_0
            shape, _, dtype = reader(fh)
        return shape, dtype

    def __getitem__(self, key):
        return self._arr[key]

    def __array__(self, dtype=None, copy=None):
        return self._arr if dtype is None else self._arr.astype(dtype)

    def __len__(self):
        return self._header()[0][0]

    @property
    def shape(self):
        return self._header()[0]

    @property
    def dtype(self):
        return self._header()[1]


class Attrs(dict):
    """Attribute dictionary of one group; marks the file dirty on mutation."""

    def __init__(self, owner: "NpzFile", *a, **kw):
        super().__init__(*a, **kw)
        self._owner = owner

    def __setitem__(self, k, v):
        if isinstance(v, np.ndarray):
            v = v.tolist()
        elif isinstance(v, np.generic):
            v = v.item()
        json.dumps(v)                       # fail here, not at close, if it cannot be stored
        self._owner._touch()
        super().__setitem__(k, v)

    def __delitem__(self, k):
        self._owner._touch()
        super().__delitem__(k)


class Group:
    def __init__(self, root: "NpzFile", name: str):
        self._root = root
        self.name = name                    # "" for the file itself, else "a/b"
        self._items: Dict[str, object] = {}
        self.attrs = Attrs(root)

    # -- navigation -------------------------------------------------------------------------------
    def _split(self, key: str):
        parts = [p for p in key.split("/") if p]
        if not parts:
            raise KeyError(key)
        return parts

    def _walk(self, parts):
        node = self
        for p in parts:
            if not isinstance(node, Group) or p not in node._items:
                return None
            node = node._items[p]
        return node

    def __contains__(self, key) -> bool:
        return isinstance(key, str) and self._walk(self._split(key)) is not None

    def __getitem__(self, key: str):
        node = self._walk(self._split(key))
        if node is None:
            raise KeyError(f"Unable to open object (object '{key}' doesn't exist)")
        return node

    def __setitem__(self, key: str, value):
        self._root._require_writable()
        parts = self._split(key)
        parent = self
        for p in parts[:-1]:
            parent = parent.require_group(p)
        leaf = parts[-1]
        if leaf in parent._items:
            raise OSError(f"Unable to create dataset (name already exists): {leaf}")      # h5py behaviour
        if isinstance(value, (Group, Dataset)):
            raise TypeError("links between groups are not supported by the npz container")
        arr = np.asarray(value)
        if arr.dtype == object:
            raise TypeError("object arrays cannot be stored")
        parent._items[leaf] = Dataset(arr)
        self._root._touch()

    def __delitem__(self, key: str):
        self._root._require_writable()
        parts = self._split(key)
        parent = self._walk(parts[:-1]) if len(parts) > 1 else self
        if parent is None or parts[-1] not in parent._items:
            raise KeyError(key)
        del parent._items[parts[-1]]
        self._root._touch()

    def create_dataset(self, name: str, data=None, **_ignored):
        self[name] = data
        return self[name]

    def create_group(self, name: str) -> "Group":
        self._root._require_writable()
        parts = self._split(name)
        parent = self
        for p in parts[:-1]:
            parent = parent.require_group(p)
        if parts[-1] in parent._items:
            raise ValueError(f"Unable to create group (name already exists): {name}")
        g = Group(self._root, (parent.name + "/" if parent.name else "") + parts[-1])
        parent._items[parts[-1]] = g
        self._root._touch()
        return g

    def require_group(self, name: str) -> "Group":
        node = self._walk(self._split(name))
        if node is None:
            return self.create_group(name)
        if not isinstance(node, Group):
            raise TypeError(f"Incompatible object (Dataset) already exists: {name}")
        return node

    def keys(self):
        return self._items.keys()

    def items(self):
        return self._items.items()

    def __iter__(self):
        return iter(self._items)

    def __len__(self):
        return len(self._items)

    # -- flattening -------------------------------------------------------------------------------
    def _flatten(self, arrays: dict, attrs: dict, groups: list):
        if self.attrs:
            attrs[self.name] = dict(self.attrs)
        for k, v in self._items.items():
            path = (self.name + "/" if self.name else "") + k
            if isinstance(v, Group):
                groups.append(path)
                v._flatten(arrays, attrs, groups)
            else:
                arrays[path] = v


class NpzFile(Group):
    """``h5py.File``-shaped container persisted as one ``.npz`` archive.

    mode ``r`` (must exist), ``a`` (read/write, created if missing), ``w`` (start empty),
    ``r+`` (read/write, must exist).  Changes are written on ``close()`` / leaving the ``with``
    block, to a temporary file in the same directory that then replaces the target."""

    def __init__(self, path: str, mode: str = "r"):
        if mode not in ("r", "r+", "a", "w"):
            raise ValueError(f"Invalid mode; must be one of r, r+, w, a (got {mode!r})")
        self.path = os.fspath(path)
        self.mode = mode
        self._dirty = False
        self._closed = False
        self._loading = True
        self._zip = None
        super().__init__(self, "")
        exists = os.path.exists(self.path)
        if mode in ("r", "r+") and not exists:
            raise FileNotFoundError(f"Unable to open file (no such file: {self.path})")
        if exists and mode != "w":
            self._load()
        self._loading = False
        if mode == "w":
            self._dirty = True

    def _require_writable(self):
        if self._closed:
            raise ValueError("Invalid file (file is closed)")
        if self.mode == "r" and not self._loading:
            raise OSError("Unable to modify (file is opened read-only)")

    def _touch(self):
        if not self._loading:
            self._require_writable()
            self._dirty = True

    def _load(self):
        self._zip = np.load(self.path, allow_pickle=False)          # members are read on demand
        z = self._zip
        meta = json.loads(str(z[ATTRS_KEY][()])) if ATTRS_KEY in z.files else {"attrs": {}, "groups": []}
        for g in meta.get("groups", []):
            self.require_group(g)
        for key in z.files:
            if key == ATTRS_KEY:
                continue
            parts = [p for p in key.split("/") if p]
            parent = self
            for p in parts[:-1]:
                parent = parent.require_group(p)
            parent._items[parts[-1]] = Dataset(source=z, key=key)
        for gname, kv in meta.get("attrs", {}).items():
            node = self if gname == "" else self.require_group(gname)
            for k, v in kv.items():
                node.attrs[k] = v

    def flush(self):
        if not self._dirty or self.mode == "r":
            return
        import shutil
        import zipfile
        datasets, attrs, groups = {}, {}, []
        self._flatten(datasets, attrs, groups)
        directory = os.path.dirname(os.path.abspath(self.path))
        fd, tmp = tempfile.mkstemp(prefix=".hg_b200_", suffix=".npz", dir=directory)
        try:
            with os.fdopen(fd, "wb") as raw, zipfile.ZipFile(raw, "w", zipfile.ZIP_STORED, allowZip64=True) as out:
                def put(name, arr):
                    with out.open(name + ".npy", "w", force_zip64=True) as fh:
                        np.lib.format.write_array(fh, np.asanyarray(arr), allow_pickle=False)
                for name, ds in datasets.items():
                    if ds.loaded:
                        put(name, ds._arr)
                    else:                                          # untouched member: copy its bytes as they are
                        with ds._source.zip.open(ds._key + ".npy") as fi, \
                                out.open(name + ".npy", "w", force_zip64=True) as fo:
                            shutil.copyfileobj(fi, fo, 1 << 24)
                put(ATTRS_KEY, np.array(json.dumps({"attrs": attrs, "groups": groups})))
            os.replace(tmp, self.path)
        except BaseException:
            if os.path.exists(tmp):
                os.unlink(tmp)
            raise
        self._dirty = False

    def _release(self):
        if self._zip is not None:
            # arrays handed out stay valid (they are copies); unread members can no longer be read
            self._zip.close()
            self._zip = None
        self._closed = True

    def close(self):
        if not self._closed:
            self.flush()
            self._release()

    def __enter__(self):
        return self

    def __exit__(self, exc_type, exc, tb):
        if exc_type is None:
            self.close()
        else:                               # leave the file on disk as it was
            self._release()
        return False


def is_npz(path) -> bool:
    return os.fspath(path).lower().endswith(".npz")


def _h5py():
    try:
        import h5py
    except ModuleNotFoundError as e:       # same failure mode as the reference's module import (:15)
        raise ModuleNotFoundError("h5py is required for the HDF5 I/O functions of hg_b200 "
                                  "(use a .npz slide container where h5py is not installed)") from e
    return h5py


def open_store(path, mode: str = "r"):
    """The slide container at `path`: NpzFile for ``*.npz``, h5py.File otherwise."""
    if is_npz(path):
        return NpzFile(path, mode)
    return _h5py().File(path, mode)


def _copy_tree(src, dst, is_group):
    for k, v in src.attrs.items():
        dst.attrs[k] = v
    for name in src.keys():
        node = src[name]
        if is_group(node):
            _copy_tree(node, dst.require_group(name), is_group)
        else:
            if name in dst:
                del dst[name]
            dst[name] = np.asarray(node[()] if getattr(node, "shape", None) == () else node[:])


def export_to_h5(npz_path: str, h5_path: str) -> None:
    """Write the content of an .npz slide container into an HDF5 file with the reference's layout."""
    h5py = _h5py()
    with NpzFile(npz_path, "r") as src, h5py.File(h5_path, "a") as dst:
        _copy_tree(src, dst, lambda n: isinstance(n, Group))


def import_from_h5(h5_path: str, npz_path: str, groups: Optional[list] = None) -> None:
    """Copy an HDF5 slide file (or just the named top-level groups) into an .npz container."""
    h5py = _h5py()
    with h5py.File(h5_path, "r") as src, NpzFile(npz_path, "a") as dst:
        if groups is None:
            _copy_tree(src, dst, lambda n: hasattr(n, "keys"))
        else:
            for g in groups:
                if g in src:
                    _copy_tree(src[g], dst.require_group(g), lambda n: hasattr(n, "keys"))
