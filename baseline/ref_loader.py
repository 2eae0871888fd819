"""Import the unmodified reference package from baseline/_ref (see install_ref.py).

Test / benchmark infrastructure only: the product package never imports this module.

The reference imports ``h5py`` at module scope (preprocess_hypergraph.py:15) although only its
file I/O uses it; h5py is not in this image, so an empty stand-in module is registered when the
real one is missing.  Nothing of the reference itself is modified.  The package is loaded under
the private name ``_hg_reference`` so that it cannot shadow (or be shadowed by) the drop-in alias
package ``build_hypergraph`` of this repository.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref", "build_hypergraph")
_NAME = "_hg_reference"


def available() -> bool:
    return os.path.isfile(os.path.join(REF_DIR, "preprocess_hypergraph.py"))


def load():
    """Returns (package, preprocess_hypergraph module, similarity_kernel module)."""
    if _NAME in sys.modules:
        pkg = sys.modules[_NAME]
        return pkg, sys.modules[_NAME + ".preprocess_hypergraph"], sys.modules[_NAME + ".similarity_kernel"]
    if not available():
        raise FileNotFoundError("baseline/_ref is empty: run `python baseline/install_ref.py` in the build container")
    if "h5py" not in sys.modules:
        try:
            import h5py  # noqa: F401
        except ModuleNotFoundError:
            sys.modules["h5py"] = types.ModuleType("h5py")
    spec = importlib.util.spec_from_file_location(_NAME, os.path.join(REF_DIR, "__init__.py"),
                                                  submodule_search_locations=[REF_DIR])
    pkg = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = pkg
    spec.loader.exec_module(pkg)                      # its relative imports load the two submodules
    return pkg, sys.modules[_NAME + ".preprocess_hypergraph"], sys.modules[_NAME + ".similarity_kernel"]


class NoCliqueKMeans:
    """Stand-in for sklearn.cluster.KMeans that puts every row in a cluster id >= num_hyperedges, so
    the reference's clique stage (preprocess_hypergraph.py:391-400) adds nothing: the kNN + de-dup +
    weights part of `build_hypergraph_knn_kmeans` alone.  Installed with
    ``preprocess_hypergraph.KMeans = NoCliqueKMeans``; the reference's code is not edited."""

    def __init__(self, n_clusters=8, **kw):
        self.n_clusters = n_clusters

    def fit_predict(self, x):
        import numpy as np
        return np.full(x.shape[0], 10 ** 6, dtype=np.int64)
