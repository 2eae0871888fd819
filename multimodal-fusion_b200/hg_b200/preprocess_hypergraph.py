"""Drop-in for the reference's ``build_hypergraph/preprocess_hypergraph.py``.

Per-slide pipeline (reference :514-603): load h5 -> KMeans super-patches (+ N x N
kernel statistics) -> WSI x TMA Gaussian similarity -> KMeans grouping -> kNN +
clique hyperedges, de-dup, cosine weights -> save h5.  Names, argument order,
defaults, return arity and the h5 layout follow the reference; the arithmetic on
the path (similarity, kNN, edge assembly, weights, statistics) runs in sm_100a
CUDA through the C ABI.  KMeans stays scikit-learn on the host exactly as in the
reference (labels are an input of the edge stage).

Differences a user can observe, all documented in DESIGN.md:
  * edge_index comes out sorted by (min, max) (reference: CPython set order);
  * tensors are returned on the CUDA device;
  * re-saving into an h5 that already holds a hypergraph overwrites the datasets
    (the reference would raise on the second save).
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib as L
from . import ops
from .similarity_kernel import compute_combined_similarity
from .store import _h5py, open_store  # noqa: F401  (_h5py kept for callers of the old name)

MAX_DENSE_N = int(os.environ.get("HG_B200_MAX_DENSE_N", "32768"))   # N x N FP32 = 4 GiB at 32k
STORED_TOPK = int(os.environ.get("HG_B200_STORED_TOPK", "32"))      # per-row entries of K_wsi kept beyond that


KMEANS_BACKEND = os.environ.get("HG_B200_KMEANS", "sklearn")        # "sklearn" (reference behaviour) | "cuda"


def _kmeans(n_clusters: int):
    """KMeans(n_clusters, random_state=42, n_init=10) as at reference :150, :299, :391.  The default
    is scikit-learn on the host, label-for-label what the reference computes; HG_B200_KMEANS=cuda
    (or KMEANS_BACKEND = "cuda") switches to the GPU Lloyd of hg_b200.kmeans (statistical parity)."""
    if KMEANS_BACKEND == "cuda":
        from .kmeans import KMeans
    else:
        from sklearn.cluster import KMeans     # reference :20
    return KMeans(n_clusters=n_clusters, random_state=42, n_init=10)


def _cluster(x: torch.Tensor, n_clusters: int) -> np.ndarray:
    est = _kmeans(n_clusters)
    return est.fit_predict(x if KMEANS_BACKEND == "cuda" else x.detach().cpu().numpy())


def _device(device, *tensors) -> torch.device:
    if device is not None:
        return L.require_cuda(device)
    for t in tensors:
        if t is not None and t.is_cuda:
            return t.device
    return L.require_cuda(None)


# ------------------------------------------------------------------------------ I/O
def load_wsi_data(h5_path: str) -> Tuple[torch.Tensor, torch.Tensor]:
    """wsi/features [N, D] and wsi/positions [N, 2|3] as float32 CPU tensors (reference :31-62)."""
    with open_store(h5_path, "r") as f:
        if "wsi" not in f or "features" not in f["wsi"]:
            raise ValueError(f"WSI features not found in {h5_path}")
        feats = torch.from_numpy(np.asarray(f["wsi"]["features"][:])).float()
        if "positions" in f["wsi"]:
            pos = torch.from_numpy(np.asarray(f["wsi"]["positions"][:])).float()
        else:
            pos = torch.zeros(feats.shape[0], 2, dtype=torch.float32)
            print("⚠️  WSI positions not found, using dummy positions")
    return feats, pos


def load_tma_data(h5_path: str) -> Optional[torch.Tensor]:
    """tma/features [N_tma, D] or None (reference :65-84)."""
    with open_store(h5_path, "r") as f:
        if "tma" in f and "features" in f["tma"]:
            return torch.from_numpy(np.asarray(f["tma"]["features"][:])).float()
    return None


def load_similarity_matrices(h5_path: str):
    """(hypergraph/similarity/wsi_internal, .../wsi_tma) or None each (reference :726-755)."""
    wsi_sim = wsi_tma = None
    with open_store(h5_path, "r") as f:
        if "hypergraph" in f and "similarity" in f["hypergraph"]:
            g = f["hypergraph"]["similarity"]
            if "wsi_internal" in g:
                wsi_sim = torch.from_numpy(np.asarray(g["wsi_internal"][:])).float()
            if "wsi_tma" in g:
                wsi_tma = torch.from_numpy(np.asarray(g["wsi_tma"][:])).float()
    return wsi_sim, wsi_tma


def load_similarity_topk(h5_path: str):
    """(indices int64 [N, k], values float32 [N, k]) of hypergraph/similarity/wsi_topk_*, or None."""
    with open_store(h5_path, "r") as f:
        if "hypergraph" in f and "similarity" in f["hypergraph"]:
            g = f["hypergraph"]["similarity"]
            if "wsi_topk_indices" in g and "wsi_topk_values" in g:
                return (torch.from_numpy(np.asarray(g["wsi_topk_indices"][:])).long(),
                        torch.from_numpy(np.asarray(g["wsi_topk_values"][:])).float())
    return None


def _put(group, name, array):
    if name in group:
        del group[name]
    group[name] = array


def save_hypergraph_to_h5(h5_path: str, wsi_super_features: torch.Tensor, wsi_super_positions: torch.Tensor,
                          tma_features: torch.Tensor, edge_index: torch.Tensor, edge_weights: torch.Tensor,
                          group_labels: np.ndarray, stats: Dict,
                          wsi_similarity_matrix: Optional[torch.Tensor] = None,
                          wsi_tma_similarity_matrix: Optional[torch.Tensor] = None,
                          wsi_similarity_topk: Optional[Tuple[torch.Tensor, torch.Tensor]] = None):
    """Write group ``hypergraph/`` (layout of reference :474-511, README.md:58-73).  Existing
    datasets are replaced (the reference's plain assignment fails on a file that was already
    processed).  `wsi_similarity_topk` = (indices [N, k], values [N, k]) is what large
    slides store in place of the dense ``similarity/wsi_internal``.  `h5_path` may be a ``.npz``
    container (hg_b200.store)."""
    with open_store(h5_path, "a") as f:
        hg = f.require_group("hypergraph")
        sup = hg.require_group("wsi_super")
        _put(sup, "features", wsi_super_features.cpu().numpy())
        _put(sup, "positions", wsi_super_positions.cpu().numpy())
        _put(hg.require_group("tma"), "features", tma_features.cpu().numpy())
        _put(hg, "edge_index", edge_index.cpu().numpy())
        _put(hg, "edge_weights", edge_weights.cpu().numpy())
        _put(hg, "group_labels", np.asarray(group_labels))
        if wsi_similarity_matrix is not None:
            sim = hg.require_group("similarity")
            _put(sim, "wsi_internal", wsi_similarity_matrix.cpu().numpy())
            sim.attrs["wsi_shape"] = list(wsi_similarity_matrix.shape)
        if wsi_tma_similarity_matrix is not None:
            sim = hg.require_group("similarity")
            _put(sim, "wsi_tma", wsi_tma_similarity_matrix.cpu().numpy())
            sim.attrs["wsi_tma_shape"] = list(wsi_tma_similarity_matrix.shape)
        if wsi_similarity_topk is not None:
            sim = hg.require_group("similarity")
            _put(sim, "wsi_topk_indices", wsi_similarity_topk[0].cpu().numpy())
            _put(sim, "wsi_topk_values", wsi_similarity_topk[1].cpu().numpy())
        hg.attrs["stats"] = json.dumps(stats, default=float)


# ---------------------------------------------------------------------- pipeline stages
def aggregate_wsi_super_patches(wsi_features: torch.Tensor, wsi_positions: torch.Tensor, num_super_patches: int,
                                lambda_h: float = 1.0, lambda_g: float = 1.0,
                                device: Optional[torch.device] = None,
                                wsi_similarity_matrix: Optional[torch.Tensor] = None
                                ) -> Tuple[torch.Tensor, torch.Tensor, Dict, torch.Tensor]:
    """KMeans super-patches + kernel statistics (reference :87-199).

    Returns (super_features [S, D], super_positions [S, 2|3], stats, K_wsi).  K_wsi is the
    dense N x N kernel when N <= HG_B200_MAX_DENSE_N, else None (the reference cannot run
    there at all) and its statistics are streamed through the tensor-core kernel (A9)."""
    dev = _device(device, wsi_features)
    feats = L.f32c(wsi_features, dev)
    pos = L.f32c(wsi_positions, dev)
    n = feats.shape[0]
    if wsi_similarity_matrix is not None:
        K_wsi = L.f32c(wsi_similarity_matrix, dev)
    elif n <= MAX_DENSE_N:
        K_wsi = compute_combined_similarity(feats, pos, lambda_h, lambda_g)
    else:
        K_wsi = None
    labels_np = _cluster(feats, num_super_patches)                              # reference :147-151
    # segmented means (:157-170) and intra-cluster similarity means (:175-184): one device call
    agg = ops.aggregate_super_patches(feats, pos, labels_np, num_super_patches, lambda_h, lambda_g,
                                      dense_k=K_wsi, device=dev)
    avg_intra, _, first_empty, _ = agg["summary"].tolist()
    if first_empty >= 0:
        raise ValueError(f"Cluster {int(first_empty)} is empty")                # reference :164
    super_f, super_p = agg["features"], agg["positions"]
    if K_wsi is not None:
        k_stats = ops.similarity_stats(K_wsi)
    else:
        k_stats = ops.similarity_stats_streaming(feats, pos, lambda_h, lambda_g, device=dev)
    stats = {
        "num_original_patches": n,
        "num_super_patches": num_super_patches,
        "avg_intra_cluster_similarity": float(avg_intra),
        "wsi_similarity_matrix_stats": k_stats,
    }
    return super_f, super_p, stats, K_wsi


def compute_wsi_tma_similarity(wsi_features: torch.Tensor, wsi_positions: torch.Tensor, tma_features: torch.Tensor,
                               lambda_h: float = 1.0, lambda_g: float = 1.0,
                               device: Optional[torch.device] = None) -> Tuple[torch.Tensor, Dict]:
    """[N_super, N_tma] Gaussian kernel on features only, direct-difference form, + statistics
    (reference :202-267; positions / lambda_g are unused there as well, :245-246)."""
    dev = _device(device, wsi_features)
    sim = ops.similarity_dense(wsi_features, tma_features, lambda_h=lambda_h, lambda_g=0.0,
                               form=L.FORM_DIRECT, device=dev)
    return sim, ops.similarity_stats(sim)


def group_by_similarity(similarity_matrix: torch.Tensor, num_groups: int, method: str = "kmeans"
                        ) -> Tuple[np.ndarray, Dict]:
    """Group WSI super-patches by their similarity rows (reference :270-332)."""
    sim_np = similarity_matrix.detach().cpu().numpy()
    if method == "kmeans":
        labels = _cluster(similarity_matrix, num_groups)
        stats = {"method": "kmeans", "num_groups": num_groups}
    elif method == "knn":
        # unused by the pipeline (process_single_file passes 'kmeans'); kept on scikit-learn
        # exactly as the reference does it (:308-321)
        from sklearn.neighbors import NearestNeighbors
        k = min(num_groups, sim_np.shape[1])
        knn = NearestNeighbors(n_neighbors=k, metric="cosine").fit(sim_np.T)
        nearest = knn.kneighbors(sim_np)[1][:, 0]
        labels = _kmeans(num_groups).fit_predict(nearest.reshape(-1, 1))
        stats = {"method": "knn", "num_groups": num_groups, "k": k}
    else:
        raise ValueError(f"Unknown grouping method: {method}")
    stats["group_sizes"] = [int(np.sum(labels == i)) for i in range(num_groups)]
    return labels, stats


def build_hypergraph_knn_kmeans(wsi_features: torch.Tensor, tma_features: torch.Tensor, group_labels: np.ndarray,
                                k: int = 5, num_hyperedges: int = 10, device: Optional[torch.device] = None,
                                *, hyperedge_labels=None, return_csr: bool = False
                                ) -> Tuple[torch.Tensor, torch.Tensor, Dict]:
    """kNN + KMeans-clique hyperedges (reference :335-433) -- THE hot function.

    Euclidean kNN over cat([wsi_super; tma]) (self excluded by index), all pairs inside each
    of `num_hyperedges` KMeans clusters, undirected de-dup, weights max(0, cos).  Keyword-only
    extras: `hyperedge_labels` injects the cluster labels (skips the host KMeans; labels
    outside [0, num_hyperedges) form no clique), `return_csr` adds the CSR incidence to stats.
    """
    dev = _device(device, wsi_features)
    allf = torch.cat([L.f32c(wsi_features, dev), L.f32c(tma_features, dev)], 0)
    n = allf.shape[0]
    if hyperedge_labels is None:
        hyperedge_labels = _cluster(allf, num_hyperedges)                            # reference :391-392
    if k <= 32:
        # kNN (reference :379-388) + cliques + de-dup (:394-410) + weights (:413-422): one C call
        built = ops.build_knn_hypergraph(allf, k, "euclidean", labels=torch.as_tensor(np.asarray(hyperedge_labels)),
                                         num_hyperedges=num_hyperedges, return_csr=return_csr, device=dev)
        edge_index, edge_weights = built["edge_index"], built["edge_weights"]
        csr = (built["csr_rowptr"], built["csr_col"], built["csr_eid"]) if return_csr else None
    else:
        idx, _ = ops.knn_feature(allf, k, "euclidean", precision="fp32", device=dev)
        edge_index = ops.build_edges([idx], n, torch.as_tensor(np.asarray(hyperedge_labels)), num_hyperedges, device=dev)
        edge_weights = ops.edge_weights_cosine(allf, edge_index)
        csr = ops.csr_from_coo(edge_index, n) if return_csr else None
    stats = {
        "num_nodes": n,
        "num_wsi_super_patches": int(wsi_features.shape[0]),
        "num_tma_patches": int(tma_features.shape[0]),
        "num_edges": int(edge_index.shape[1]),
        "num_hyperedges": num_hyperedges,
        "k": k,
    }
    if return_csr:
        stats["csr"] = csr
    return edge_index, edge_weights, stats


def _load_inputs(h5_path: str):
    """(wsi_features, wsi_positions, tma_features | None) of one slide file -- the host half of a slide that the
    dataset loop reads ahead while the previous slide is on the GPU."""
    wsi_features, wsi_positions = load_wsi_data(h5_path)
    return wsi_features, wsi_positions, load_tma_data(h5_path)


def process_single_file(h5_path: str, num_wsi_super_patches: int = 100, num_groups: int = 10,
                        hypergraph_k: int = 5, num_hyperedges: int = 10, lambda_h: float = 1.0,
                        lambda_g: float = 1.0, device: Optional[torch.device] = None, *,
                        inputs=None, writer=None) -> Dict:
    """One slide end to end (reference :514-603).  Keyword-only extras for the dataset loop: `inputs` =
    what `_load_inputs` returned for this file (read ahead), `writer` = an executor that takes the save;
    the returned stats then carry the pending write under "_pending_save"."""
    dev = _device(device)
    print(f"📂 Processing: {h5_path}")
    wsi_features, wsi_positions, tma_features = inputs if inputs is not None else _load_inputs(h5_path)
    if tma_features is None:
        print("⚠️  TMA features not found, skipping hypergraph construction")
        return {"status": "skipped", "reason": "no_tma"}
    print(f"  🔄 Aggregating WSI patches into {num_wsi_super_patches} super patches...")
    sup_f, sup_p, wsi_stats, wsi_sim = aggregate_wsi_super_patches(
        wsi_features, wsi_positions, num_wsi_super_patches, lambda_h, lambda_g, dev)
    print("  🔄 Computing WSI-TMA similarity...")
    sim, sim_stats = compute_wsi_tma_similarity(sup_f, sup_p, tma_features, lambda_h, lambda_g, dev)
    print(f"  🔄 Grouping into {num_groups} groups...")
    group_labels, group_stats = group_by_similarity(sim, num_groups, method="kmeans")
    print(f"  🔄 Building hypergraph (KNN k={hypergraph_k}, KMeans {num_hyperedges} hyperedges)...")
    edge_index, edge_weights, hg_stats = build_hypergraph_knn_kmeans(
        sup_f, tma_features, group_labels, hypergraph_k, num_hyperedges, dev)
    print("  💾 Saving to h5 (including similarity matrices for fast tuning)...")
    all_stats = {"wsi_aggregation": wsi_stats, "similarity": sim_stats, "grouping": group_stats,
                 "hypergraph": hg_stats}
    topk = None
    if wsi_sim is None and wsi_features.shape[0] > 1:
        # beyond the dense limit: the k strongest entries of every row of K_wsi instead of N x N
        kk = min(STORED_TOPK, wsi_features.shape[0] - 1)
        topk = ops.knn_feature(wsi_features, kk, "gaussian", positions=wsi_positions, lambda_h=lambda_h,
                               lambda_g=lambda_g, device=dev)

    def save():
        save_hypergraph_to_h5(h5_path, sup_f, sup_p, tma_features, edge_index, edge_weights, group_labels, all_stats,
                              wsi_similarity_matrix=wsi_sim, wsi_tma_similarity_matrix=sim, wsi_similarity_topk=topk)
        print(f"  ✅ Done! ({os.path.basename(h5_path)})" if writer is not None else "  ✅ Done!")

    if writer is None:
        save()
        return all_stats
    out = dict(all_stats)
    out["_pending_save"] = writer.submit(save)          # D2H + file write run behind the next slide's kernels
    return out


IO_OVERLAP = os.environ.get("HG_B200_IO_OVERLAP", "1") != "0"
MAX_PENDING_SAVES = max(1, int(os.environ.get("HG_B200_MAX_PENDING_SAVES", "2")))


class _BoundedWriter:
    """One writer thread with at most `max_pending` saves queued or running: submit() blocks the slide loop
    when the writer falls behind.  A queued save keeps its slide's device tensors alive (up to the dense
    N x N K_wsi, 4 GiB at MAX_DENSE_N), so an unbounded queue could run the GPU out of memory whenever the
    D2H + file write are slower than the GPU stages; the reference's sequential loop (:651-670) never holds
    more than one slide."""

    def __init__(self, max_pending: int = MAX_PENDING_SAVES):
        import threading
        from concurrent.futures import ThreadPoolExecutor
        self._pool = ThreadPoolExecutor(1)
        self._slots = threading.BoundedSemaphore(max_pending)

    def submit(self, fn, *args, **kw):
        self._slots.acquire()

        def run():
            try:
                return fn(*args, **kw)
            finally:
                self._slots.release()
        try:
            return self._pool.submit(run)
        except BaseException:
            self._slots.release()
            raise

    def shutdown(self, wait: bool = True):
        self._pool.shutdown(wait=wait)


def _run_over_csv(csv_path: str, data_root_dir: str, desc: str, worker, output_stats_path: Optional[str],
                  loader=None):
    """Per-slide loop with the reference's error convention: a failing slide is reported and skipped
    (reference :651-670, :963-982).  The reference loop is strictly sequential (load -> build -> save).
    Here, when `loader` is given (HG_B200_IO_OVERLAP=0 turns it off), the next slide's file is read by a
    helper thread while this one is on the GPU, and saves run on a writer thread: `worker(path, inputs=...,
    writer=...)`.  Under torch.distributed (world size > 1) every rank takes the slides of its LPT share
    (multigpu.shard_slides on the file sizes, a proxy for N); rank 0 gathers and writes the statistics."""
    import pandas as pd
    from tqdm import tqdm
    import torch.distributed as dist
    from concurrent.futures import ThreadPoolExecutor
    from .multigpu import shard_slides
    df = pd.read_csv(csv_path)
    if "h5_file_path" not in df.columns:
        raise ValueError("CSV must contain 'h5_file_path' column")
    todo = []
    for idx, row in df.iterrows():
        rel = row["h5_file_path"]
        h5_path = os.path.join(data_root_dir, rel)
        if not os.path.exists(h5_path):
            print(f"⚠️  File not found: {h5_path}")
            continue
        todo.append((idx, row, rel, h5_path))
    world, rank = 1, 0
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        world, rank = dist.get_world_size(), dist.get_rank()
        sizes = [os.path.getsize(t[3]) if os.path.isfile(t[3]) else 1 for t in todo]
        todo = [todo[i] for i in shard_slides(sizes, world)[rank]]
    overlap = loader is not None and IO_OVERLAP
    reader = ThreadPoolExecutor(1) if overlap else None
    writer = _BoundedWriter() if overlap else None
    done = []                                           # (csv index, stats) of the slides that went through
    try:
        ahead = reader.submit(loader, todo[0][3]) if overlap and todo else None
        for pos, (idx, row, rel, h5_path) in enumerate(tqdm(todo, total=len(todo), desc=desc)):
            mine, ahead = ahead, None
            if overlap and pos + 1 < len(todo):
                ahead = reader.submit(loader, todo[pos + 1][3])
            try:
                stats = worker(h5_path, inputs=mine.result(), writer=writer) if overlap else worker(h5_path)
                stats["case_id"] = row.get("case_id", f"case_{idx}")
                stats["h5_path"] = rel
                done.append((idx, stats))
            except Exception as e:                      # noqa: BLE001 -- reference behaviour
                print(f"❌ Error processing {h5_path}: {e}")
                import traceback
                traceback.print_exc()
        all_stats = []
        for idx, stats in done:                         # a save that failed drops its slide, like any other error
            pending = stats.pop("_pending_save", None)
            try:
                if pending is not None:
                    pending.result()
                all_stats.append((idx, stats))
            except Exception as e:                      # noqa: BLE001
                print(f"❌ Error saving {stats.get('h5_path')}: {e}")
    finally:
        for ex in (reader, writer):
            if ex is not None:
                ex.shutdown(wait=True)
    if world > 1:
        parts = [None] * world
        dist.all_gather_object(parts, all_stats)
        all_stats = sorted((item for part in parts for item in part), key=lambda t: t[0])
    all_stats = [s for _, s in all_stats]
    if output_stats_path and rank == 0:
        with open(output_stats_path, "w") as f:
            json.dump(all_stats, f, indent=2, default=float)
        print(f"📊 Statistics saved to: {output_stats_path}")
    return all_stats


def process_dataset(csv_path: str, data_root_dir: str, num_wsi_super_patches: int = 100, num_groups: int = 10,
                    hypergraph_k: int = 5, num_hyperedges: int = 10, lambda_h: float = 1.0, lambda_g: float = 1.0,
                    output_stats_path: Optional[str] = None, device: Optional[torch.device] = None):
    """All slides of a CSV (reference :606-678)."""
    return _run_over_csv(
        csv_path, data_root_dir, "Processing files",
        lambda p, **io: process_single_file(p, num_wsi_super_patches, num_groups, hypergraph_k, num_hyperedges,
                                            lambda_h, lambda_g, device, **io),
        output_stats_path, loader=_load_inputs)


def rebuild_hypergraph_from_similarity(h5_path: str, num_wsi_super_patches: int = None, num_groups: int = None,
                                       hypergraph_k: int = 5, num_hyperedges: int = 10,
                                       threshold_median_ratio: float = None,
                                       device: Optional[torch.device] = None) -> Dict:
    """Re-run the stages with new counts, reusing what the h5 already stores (reference :758-916;
    lambda is hard-coded to 1.0 there, :812-830)."""
    dev = _device(device)
    print(f"🔄 Rebuilding hypergraph from stored similarity matrices: {h5_path}")
    wsi_features, wsi_positions = load_wsi_data(h5_path)
    tma_features = load_tma_data(h5_path)
    if tma_features is None:
        raise ValueError("TMA features not found")
    wsi_sim, wsi_tma_sim = load_similarity_matrices(h5_path)
    if wsi_sim is None and wsi_features.shape[0] <= MAX_DENSE_N:
        print("⚠️  WSI similarity matrix not found, recomputing...")
        wsi_sim = compute_combined_similarity(wsi_features.to(dev), wsi_positions.to(dev), lambda_h=1.0, lambda_g=1.0)
    wsi_stats = {}
    if num_wsi_super_patches is not None:
        sup_f, sup_p, wsi_stats, _ = aggregate_wsi_super_patches(
            wsi_features, wsi_positions, num_wsi_super_patches, lambda_h=1.0, lambda_g=1.0, device=dev,
            wsi_similarity_matrix=wsi_sim)
        sim, sim_stats = compute_wsi_tma_similarity(sup_f, sup_p, tma_features, 1.0, 1.0, dev)
    else:
        with open_store(h5_path, "r") as f:
            if "hypergraph" not in f or "wsi_super" not in f["hypergraph"]:
                raise ValueError("WSI super patches not found and num_wsi_super_patches not specified")
            sup_f = torch.from_numpy(np.asarray(f["hypergraph"]["wsi_super"]["features"][:])).float().to(dev)
            sup_p = torch.from_numpy(np.asarray(f["hypergraph"]["wsi_super"]["positions"][:])).float().to(dev)
        if wsi_tma_sim is not None and wsi_tma_sim.shape[0] == sup_f.shape[0]:
            sim = wsi_tma_sim.to(dev)
            sim_stats = ops.similarity_stats(sim)
        else:
            sim, sim_stats = compute_wsi_tma_similarity(sup_f, sup_p, tma_features, 1.0, 1.0, dev)
    if num_groups is not None:
        group_labels, group_stats = group_by_similarity(sim, num_groups, method="kmeans")
    else:
        with open_store(h5_path, "r") as f:
            if "hypergraph" not in f or "group_labels" not in f["hypergraph"]:
                raise ValueError("Group labels not found and num_groups not specified")
            group_labels = np.asarray(f["hypergraph"]["group_labels"][:])
        group_stats = {"method": "existing", "num_groups": int(len(np.unique(group_labels)))}
    edge_index, edge_weights, hg_stats = build_hypergraph_knn_kmeans(
        sup_f, tma_features, group_labels, hypergraph_k, num_hyperedges, dev)
    if threshold_median_ratio is not None and edge_weights.numel():
        threshold = float(edge_weights.median()) * threshold_median_ratio       # reference :890-897
        mask = edge_weights >= threshold
        edge_index, edge_weights = edge_index[:, mask], edge_weights[mask]
        hg_stats.update(num_edges_after_threshold=int(edge_weights.shape[0]), threshold=threshold,
                        threshold_ratio=threshold_median_ratio)
    all_stats = {"wsi_aggregation": wsi_stats if num_wsi_super_patches is not None else {},
                 "similarity": sim_stats, "grouping": group_stats, "hypergraph": hg_stats}
    save_hypergraph_to_h5(h5_path, sup_f, sup_p, tma_features, edge_index, edge_weights, group_labels, all_stats,
                          wsi_similarity_matrix=wsi_sim, wsi_tma_similarity_matrix=sim)
    print("  ✅ Done!")
    return all_stats


def batch_rebuild_hypergraph(csv_path: str, data_root_dir: str, num_wsi_super_patches: int = None,
                             num_groups: int = None, hypergraph_k: int = 5, num_hyperedges: int = 10,
                             threshold_median_ratio: float = None, output_stats_path: Optional[str] = None,
                             device: Optional[torch.device] = None):
    """Rebuild every slide of a CSV (reference :919-990)."""
    return _run_over_csv(
        csv_path, data_root_dir, "Rebuilding hypergraphs",
        lambda p: rebuild_hypergraph_from_similarity(p, num_wsi_super_patches, num_groups, hypergraph_k,
                                                     num_hyperedges, threshold_median_ratio, device),
        output_stats_path)


def main(argv=None):
    """CLI with the reference's flags (reference :681-723)."""
    import argparse
    ap = argparse.ArgumentParser(description="Preprocess hypergraph data (B200-native)")
    ap.add_argument("--csv_path", type=str, required=True, help="CSV with an h5_file_path column")
    ap.add_argument("--data_root_dir", type=str, required=True, help="root directory of the h5 files")
    ap.add_argument("--num_wsi_super_patches", type=int, default=100)
    ap.add_argument("--num_groups", type=int, default=10)
    ap.add_argument("--hypergraph_k", type=int, default=5)
    ap.add_argument("--num_hyperedges", type=int, default=10)
    ap.add_argument("--lambda_h", type=float, default=1.0)
    ap.add_argument("--lambda_g", type=float, default=1.0)
    ap.add_argument("--output_stats", type=str, default=None)
    ap.add_argument("--device", type=str, default="auto", help="auto or cuda[:i]; there is no CPU path")
    args = ap.parse_args(argv)
    device = L.require_cuda(None if args.device == "auto" else args.device)
    print(f"🚀 Starting preprocessing with device: {device}")
    process_dataset(args.csv_path, args.data_root_dir, args.num_wsi_super_patches, args.num_groups,
                    args.hypergraph_k, args.num_hyperedges, args.lambda_h, args.lambda_g, args.output_stats, device)


if __name__ == "__main__":
    main()
