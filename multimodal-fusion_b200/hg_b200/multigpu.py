"""Multi-GPU plumbing for the two ways the path shards.

* slide-per-GPU: slides are independent units (the reference's `process_dataset` loop,
  preprocess_hypergraph.py:651-670, has no cross-slide state) -> longest-processing-time
  assignment of slides to ranks, NO data-path collective;
* row-partitioned single slide: rank r owns rows [r*N/G, (r+1)*N/G); ONE all-gather of the
  feature row blocks (NCCL over NVLink), each rank runs the fused kernel on its row range
  against all N columns, ONE small all-gather of the [N/G, k] lists, assembly replicated.

One process per GPU (`torchrun`); `torch.distributed` is plumbing only.  The host-side logic
is backend-agnostic so world_size-2 gloo tests cover it on CPU with the kNN stage injected.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_slides(sizes: Sequence[int], world_size: int) -> List[List[int]]:
    """LPT schedule: slides sorted by cost ~ N^2 descending, each to the least-loaded rank.
    Returns, per rank, the list of slide indices (in processing order)."""
    order = sorted(range(len(sizes)), key=lambda i: (-int(sizes[i]), i))
    load = [0] * world_size
    plan: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda q: (load[q], q))
        plan[r].append(i)
        load[r] += int(sizes[i]) ** 2
    return plan


def row_range(n: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous, near-equal row blocks; the first n % G ranks get one extra row."""
    base, extra = divmod(n, world_size)
    r0 = rank * base + min(rank, extra)
    return r0, r0 + base + (1 if rank < extra else 0)


def all_gather_rows(local: torch.Tensor, n: int, group=None) -> torch.Tensor:
    """Concatenate per-rank row blocks (sizes given by row_range) into [n, ...] on every rank."""
    world = dist.get_world_size(group)
    shape = tuple(local.shape[1:])
    sizes = [row_range(n, r, world)[1] - row_range(n, r, world)[0] for r in range(world)]
    if len(set(sizes)) == 1:
        out = torch.empty((n,) + shape, dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    # ragged blocks: pad to the largest block (at most one extra row), gather, drop the padding
    m = max(sizes)
    padded = torch.zeros((m,) + shape, dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    gathered = torch.empty((world * m,) + shape, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    return torch.cat([gathered[r * m: r * m + sizes[r]] for r in range(world)], 0)


def knn_hyperedges_row_partitioned(local_features: torch.Tensor, n: int, *, k: int = 16, kernel: str = "gaussian",
                                   lambda_h: float = 1.0, lambda_g: float = 0.0,
                                   positions: Optional[torch.Tensor] = None, spatial_k: int = 0,
                                   return_csr: bool = False, group=None,
                                   knn_fn: Optional[Callable] = None, assemble_fn: Optional[Callable] = None) -> dict:
    """Config C5: `local_features` is this rank's row block [r1-r0, D]; `positions` (replicated,
    [n, 2|3]) is small.  Returns the same dict as `knn_hyperedges` on every rank."""
    from . import ops
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    r0, r1 = row_range(n, rank, world)
    assert local_features.shape[0] == r1 - r0, "local block does not match row_range()"
    feats = all_gather_rows(local_features, n, group)                   # the one big collective
    pos = positions if (positions is not None) else None
    if knn_fn is None:
        knn_fn = lambda x, rows: ops.knn_feature(x, k, kernel, positions=pos if lambda_g else None,   # noqa: E731
                                                 lambda_h=lambda_h, lambda_g=lambda_g, rows=rows, return_cos=True)
    local = knn_fn(feats, (r0, r1))
    idx_local, val_local = local[0], local[1]
    out = {"knn_idx": all_gather_rows(idx_local, n, group), "knn_val": all_gather_rows(val_local, n, group)}
    # cos(f_i, f_j) of every list entry travels with the lists: the edge weights are then looked up
    # instead of gathering two feature rows per edge
    knn_cos = all_gather_rows(local[2], n, group) if len(local) > 2 else None
    lists = [out["knn_idx"]]
    if spatial_k and pos is not None:
        sidx, sd2 = ops.knn_spatial(pos, spatial_k, rows=(r0, r1))
        out["spatial_idx"] = all_gather_rows(sidx, n, group)
        out["spatial_d2"] = all_gather_rows(sd2, n, group)
        lists.append(out["spatial_idx"])
    if assemble_fn is None:
        def assemble_fn(lists, feats):
            ei = ops.build_edges(lists, n)
            w = (ops.edge_weights_from_knn(feats, out["knn_idx"], knn_cos, ei) if knn_cos is not None
                 else ops.edge_weights_cosine(feats, ei))
            res = {"edge_index": ei, "edge_weights": w}
            if return_csr:
                res["csr_rowptr"], res["csr_col"], res["csr_eid"] = ops.csr_from_coo(ei, n)
            return res
    out.update(assemble_fn(lists, feats))
    return out
