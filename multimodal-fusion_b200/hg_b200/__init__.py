"""hg_b200 -- B200-native hypergraph construction (drop-in for the reference's
``build_hypergraph`` package: same 17 exported names, reference __init__.py:28-46)."""
from .similarity_kernel import (
    compute_morphological_similarity,
    compute_spatial_similarity,
    compute_combined_similarity,
    build_weighted_hypergraph,
    mean_pool_with_similarity,
    build_hypergraph_data,
    knn_hyperedges,
)
from .preprocess_hypergraph import (
    process_single_file,
    process_dataset,
    load_wsi_data,
    load_tma_data,
    aggregate_wsi_super_patches,
    compute_wsi_tma_similarity,
    group_by_similarity,
    build_hypergraph_knn_kmeans,
    save_hypergraph_to_h5,
    load_similarity_matrices,
    rebuild_hypergraph_from_similarity,
    batch_rebuild_hypergraph,
    load_similarity_topk,
)
from .store import open_store
from .ops import hypergraph_propagate
from .pipeline import SlidePipeline, build_batch, build_many

__all__ = [
    "compute_morphological_similarity", "compute_spatial_similarity", "compute_combined_similarity",
    "build_weighted_hypergraph", "mean_pool_with_similarity",
    "process_single_file", "process_dataset", "load_wsi_data", "load_tma_data",
    "aggregate_wsi_super_patches", "compute_wsi_tma_similarity", "group_by_similarity",
    "build_hypergraph_knn_kmeans", "save_hypergraph_to_h5", "load_similarity_matrices",
    "rebuild_hypergraph_from_similarity", "batch_rebuild_hypergraph",
    # beyond the reference surface
    "build_hypergraph_data", "knn_hyperedges", "build_batch", "build_many", "SlidePipeline", "hypergraph_propagate",
    "load_similarity_topk", "open_store",
]
