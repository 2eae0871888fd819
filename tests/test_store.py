"""The .npz slide container (hg_b200.store): h5py-shaped tree persisted with numpy only.
CPU tests; the h5py converters are exercised against the in-memory h5py stand-in."""
import json
import os
import sys

import numpy as np
import pytest

import fake_h5py
from hg_b200 import store


def _fill(f):
    w = f.create_group("wsi")
    w["features"] = np.arange(12, dtype=np.float32).reshape(4, 3)
    w["positions"] = np.arange(8, dtype=np.int64).reshape(4, 2)
    hg = f.require_group("hypergraph")
    hg.require_group("wsi_super")["features"] = np.ones((2, 3), dtype=np.float32)
    hg["edge_index"] = np.array([[0, 1], [1, 2]], dtype=np.int64)
    hg.require_group("similarity").attrs["wsi_shape"] = [4, 4]
    hg.attrs["stats"] = json.dumps({"k": 5})
    f.require_group("empty_group")


def test_round_trip_on_disk(tmp_path):
    path = str(tmp_path / "slide.npz")
    with store.open_store(path, "w") as f:
        assert isinstance(f, store.NpzFile)
        _fill(f)
    assert os.path.exists(path) and not [p for p in os.listdir(tmp_path) if p.startswith(".hg_b200_")]
    with store.open_store(path, "r") as f:
        assert "wsi" in f and "features" in f["wsi"] and "wsi/features" in f and "nope" not in f and "wsi/nope" not in f
        assert f["wsi"]["features"][:].dtype == np.float32 and f["wsi"]["features"].shape == (4, 3)
        assert np.array_equal(f["wsi/positions"][:], np.arange(8).reshape(4, 2)) and f["wsi/positions"][:].dtype == np.int64
        assert np.array_equal(f["hypergraph"]["edge_index"][1], [1, 2])
        assert f["hypergraph"]["similarity"].attrs["wsi_shape"] == [4, 4]
        assert json.loads(f["hypergraph"].attrs["stats"]) == {"k": 5}
        assert "empty_group" in f and len(f["empty_group"]) == 0
        assert sorted(f.keys()) == ["empty_group", "hypergraph", "wsi"]
        with pytest.raises(KeyError):
            f["hypergraph"]["missing"]
        with pytest.raises(OSError):
            f["x"] = np.zeros(1)                                   # read-only


def test_append_replace_and_h5py_collision_semantics(tmp_path):
    path = str(tmp_path / "slide.npz")
    with store.NpzFile(path, "a") as f:
        _fill(f)
    with store.NpzFile(path, "a") as f:
        with pytest.raises(OSError, match="name already exists"):   # what h5py does on plain re-assignment
            f["hypergraph"]["edge_index"] = np.zeros((2, 1), dtype=np.int64)
        del f["hypergraph"]["edge_index"]
        f["hypergraph"]["edge_index"] = np.zeros((2, 5), dtype=np.int64)
        f["hypergraph"].attrs["stats"] = json.dumps({"k": 7})
        with pytest.raises(TypeError):
            f.require_group("wsi/features")                        # a dataset is in the way
    with store.NpzFile(path, "r") as f:
        assert f["hypergraph"]["edge_index"].shape == (2, 5)
        assert json.loads(f["hypergraph"].attrs["stats"]) == {"k": 7}
        assert np.array_equal(f["wsi"]["features"][:], np.arange(12, dtype=np.float32).reshape(4, 3))   # untouched


def test_failed_block_leaves_file_untouched(tmp_path):
    path = str(tmp_path / "slide.npz")
    with store.NpzFile(path, "w") as f:
        f["a"] = np.arange(3)
    before = open(path, "rb").read()
    with pytest.raises(RuntimeError):
        with store.NpzFile(path, "a") as f:
            f["b"] = np.arange(4)
            raise RuntimeError("boom")
    assert open(path, "rb").read() == before
    with pytest.raises(FileNotFoundError):
        store.NpzFile(str(tmp_path / "missing.npz"), "r")
    with pytest.raises(ValueError):
        store.NpzFile(path, "x")
    with pytest.raises(TypeError):
        with store.NpzFile(path, "a") as f:
            f["obj"] = np.array([{"a": 1}], dtype=object)


def test_open_store_dispatch_and_missing_h5py(tmp_path, monkeypatch):
    monkeypatch.setitem(sys.modules, "h5py", None)                 # import h5py -> ModuleNotFoundError / ImportError
    with pytest.raises(ImportError, match="h5py is required"):
        store.open_store(str(tmp_path / "slide.h5"), "r")
    fake_h5py.reset()
    monkeypatch.setitem(sys.modules, "h5py", fake_h5py)
    with store.open_store(str(tmp_path / "slide.h5"), "w") as f:
        assert isinstance(f, fake_h5py.File)


def test_export_and_import_h5(tmp_path, monkeypatch):
    fake_h5py.reset()
    monkeypatch.setitem(sys.modules, "h5py", fake_h5py)
    npz, h5, back = str(tmp_path / "s.npz"), str(tmp_path / "s.h5"), str(tmp_path / "back.npz")
    with store.NpzFile(npz, "w") as f:
        _fill(f)
    store.export_to_h5(npz, h5)
    with fake_h5py.File(h5, "r") as f:
        assert np.array_equal(f["hypergraph"]["wsi_super"]["features"][:], np.ones((2, 3), dtype=np.float32))
        assert f["hypergraph"]["similarity"].attrs["wsi_shape"] == [4, 4]
    store.export_to_h5(npz, h5)                                    # exporting twice replaces, does not collide
    store.import_from_h5(h5, back)
    with store.NpzFile(npz, "r") as a, store.NpzFile(back, "r") as b:
        for key in ("wsi/features", "wsi/positions", "hypergraph/edge_index", "hypergraph/wsi_super/features"):
            assert np.array_equal(a[key][:], b[key][:]) and a[key].dtype == b[key].dtype
        assert b["hypergraph"].attrs["stats"] == a["hypergraph"].attrs["stats"]
    only = str(tmp_path / "only_wsi.npz")
    store.import_from_h5(h5, only, groups=["wsi"])
    with store.NpzFile(only, "r") as f:
        assert sorted(f.keys()) == ["wsi"]


def test_members_are_read_on_demand_and_copied_untouched(tmp_path):
    path = str(tmp_path / "slide.npz")
    big = np.random.default_rng(0).standard_normal((300, 300)).astype(np.float32)
    with store.NpzFile(path, "w") as f:
        f.require_group("hypergraph/similarity")["wsi_internal"] = big
        f.require_group("tma")["features"] = np.ones((3, 4), dtype=np.float32)
    with store.NpzFile(path, "r") as f:
        node = f["hypergraph/similarity/wsi_internal"]
        assert node.shape == (300, 300) and node.dtype == np.float32 and len(node) == 300   # header only
        assert not node.loaded
        assert f["tma"]["features"][:].sum() == 12 and not node.loaded
    with store.NpzFile(path, "a") as f:                       # replace a small dataset: the big one is never parsed
        del f["tma"]["features"]
        f["tma"]["features"] = np.zeros((2, 2), dtype=np.float32)
        node = f["hypergraph/similarity/wsi_internal"]
    assert not node.loaded
    with pytest.raises(Exception):
        node[:]                                               # the file is closed, like an h5py dataset
    with store.NpzFile(path, "r") as f:
        assert np.array_equal(f["hypergraph/similarity/wsi_internal"][:], big)
        assert f["tma/features"].shape == (2, 2)
    with np.load(path) as z:                                  # still a plain numpy archive
        assert np.array_equal(z["hypergraph/similarity/wsi_internal"], big)
