"""Multi-GPU plumbing for the two ways the path shards.

* slide-per-GPU: slides are independent units (the reference's `process_dataset` loop,
  preprocess_hypergraph.py:651-670, has no cross-slide state) -> longest-processing-time
  assignment of slides to ranks, NO data-path collective;
* row-partitioned single slide: rank r owns rows [r*N/G, (r+1)*N/G); ONE all-gather of the
  feature row blocks (NCCL over NVLink), each rank runs the fused kernel on its row range
  against all N columns, ONE small all-gather of the [N/G, k] lists, assembly replicated.
  With 256-aligned row blocks (`row_range_aligned`) the ranks run the SYMMETRIC pass instead
  (hg_knn_sympart_*): they share the checkerboard of tile pairs and push the transposed survivors
  into the owning rank's buckets over NVLink; the result is bit-identical to the single-GPU build.
  (Round 1's "overlapped exchange" -- peers' blocks pulled by the copy engines while the fused
  kernel swept the blocks already here, hg_knn_staged_* -- was never faster than gather-then-compute
  on B200 (57.1 vs 56.2 ms at C5: the kernel is power-bound) and was removed in round 2.)

One process per GPU (`torchrun`); `torch.distributed` is plumbing only.  The host-side logic
is backend-agnostic so world_size-2 gloo tests cover it on CPU with the kNN stage injected.
"""
from __future__ import annotations

import os
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


PANEL_ALIGN = 256     # column tile of the fused kernel: aligned row blocks start on multiples of it


def row_range_aligned(n: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Row blocks whose boundaries are multiples of 256 (the last block takes what is left, trailing
    ranks may own nothing): the partition the symmetric pass needs."""
    chunk = -(-(-(-n // world_size)) // PANEL_ALIGN) * PANEL_ALIGN
    r0 = min(n, rank * chunk)
    return r0, min(n, r0 + chunk)


def all_gather_rows(local: torch.Tensor, n: int, group=None, ranges=row_range) -> torch.Tensor:
    """Concatenate per-rank row blocks (sizes given by `ranges`) into [n, ...] on every rank."""
    world = dist.get_world_size(group)
    shape = tuple(local.shape[1:])
    sizes = [ranges(n, r, world)[1] - ranges(n, r, world)[0] for r in range(world)]
    if len(set(sizes)) == 1:
        out = torch.empty((n,) + shape, dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    # ragged blocks: pad to the largest block, gather, drop the padding
    m = max(sizes)
    padded = torch.zeros((m,) + shape, dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    gathered = torch.empty((world * m,) + shape, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    return torch.cat([gathered[r * m: r * m + sizes[r]] for r in range(world)], 0)


class _SymPartState:
    """Per-(group, chunk) buffers of the partitioned symmetric pass: the own bucket arrays in ONE allocation, exported
    once over CUDA IPC and mapped by every peer FOR ITS OWN DEVICE (collective constructor; hg_ipc_open -- torch's
    own tensor sharing maps with the exporting device current, which serves copies but not kernels running on
    another GPU), and the flag word.  `peer_cnt` / `peer_bkt` = raw device pointers per rank, None when the mapping
    is not possible on this system (all ranks agree)."""

    BKT_CAP = 512                                      # TC_BKT_CAP (knn_tc.cuh)

    def __init__(self, chunk: int, device, group=None):
        import ctypes as C
        from . import _lib as L
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        cnt_bytes = -(-chunk * 4 // 256) * 256
        self.flags = torch.zeros(2, dtype=torch.int32, device=device)
        self.peer_cnt = self.peer_bkt = None
        self.why = ""                                  # what kept the mapping from working, for diagnostics
        self._mapped = []
        self.own = None                                # raw device pointer of the own arrays (hg_peer_alloc)
        lib = L.load_library()
        ok, handle = 1, None
        with torch.cuda.device(device):
            ptr, hbuf = C.c_void_p(0), C.create_string_buffer(64)
            if lib.hg_peer_alloc(cnt_bytes + chunk * self.BKT_CAP * 8, C.byref(ptr), hbuf) == 0 and ptr.value:
                self.own = ptr.value
                handle = (hbuf.raw, 0)
            else:
                self.why = f"export failed: {L.last_error()}"
                ok = 0
        handles = [None] * world
        dist.all_gather_object(handles, handle, group=group)
        bases = [None] * world
        if ok and all(h is not None for h in handles):
            with torch.cuda.device(device):
                for r, (hbytes, off) in enumerate(handles):
                    if r == rank:
                        bases[r] = self.own
                        continue
                    base = C.c_void_p(0)
                    if lib.hg_ipc_open(hbytes, C.byref(base)) != 0 or not base.value:
                        self.why = L.last_error()
                        ok = 0
                        break
                    self._mapped.append(base.value)
                    bases[r] = base.value + off
        else:
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 1:
            self.peer_cnt = bases
            self.peer_bkt = [b + cnt_bytes for b in bases]
            self.cnt_ptr, self.bkt_ptr = self.own, self.own + cnt_bytes


_SYMPART = {}


def _sympart_state(chunk: int, device, group=None) -> _SymPartState:
    key = (id(group) if group is not None else None, int(chunk), torch.device(device).index)
    if key not in _SYMPART:
        _SYMPART[key] = _SymPartState(chunk, device, group)
    return _SYMPART[key]


def _knn_sympart(feats: torch.Tensor, n: int, k: int, kernel: str, lambda_h: float, lambda_g: float,
                 pos: Optional[torch.Tensor], group):
    """Feature kNN of this rank's aligned row block in SYMMETRIC mode (hg_knn_sympart_*): the ranks share the
    checkerboard of tile pairs, transposed survivors travel into the owning rank's buckets over NVLink.  `feats` =
    all n rows (after the all-gather).  Returns (idx, val, cos) of the own rows, or None when peer memory cannot be
    mapped (all ranks agree)."""
    import ctypes as C
    from . import _lib as L
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    dev = feats.device
    chunk = row_range_aligned(n, 0, world)[1]
    r0, r1 = row_range_aligned(n, rank, world)
    st = _sympart_state(chunk, dev, group)
    if st.peer_bkt is None:
        if os.environ.get("HG_B200_SYMPART_DEBUG"):
            print(f"[hg_b200] rank {rank}: symmetric row partition unavailable ({st.why or 'a peer could not map'})", flush=True)
        return None
    lib = L.load_library()
    d = feats.shape[1]
    use_pos = pos is not None and kernel == "gaussian" and lambda_g != 0.0
    p = L.f32c(pos, dev) if use_pos else None
    desc = L.KnnPart(n=n, d=d, row_begin=r0, row_end=r1, reserved0=0, pos=L.ptr(p),
                       pos_dim=0 if p is None else p.shape[1], metric=L.METRIC[kernel], k=k, exclude_self=1,
                       reserved1=0, reserved2=0, lambda_h=float(lambda_h), lambda_g=float(lambda_g if use_pos else 0.0))
    n_pad = -(-n // 256) * 256
    with torch.cuda.device(dev):
        nbytes = C.c_size_t(0)
        L.check(lib.hg_knn_sympart_workspace_bytes(C.byref(desc), C.byref(nbytes)))
        ws = L.Workspace.get(dev, nbytes.value, "sympart")
        stream = L.stream_ptr(dev)
        thr = torch.full((n_pad,), float("nan"), dtype=torch.float32, device=dev)
        L.check(lib.hg_knn_sympart_prepass(C.byref(desc), L.ptr(feats), L.ptr(thr), st.cnt_ptr, L.ptr(st.flags),
                                           L.ptr(ws), nbytes.value, stream))
        # every rank needs every column's bound; behind this collective all bucket counters are cleared
        thr[:n] = all_gather_rows(thr[r0:r1].clone(), n, group, row_range_aligned)
        arr_cnt = (C.c_void_p * world)(*st.peer_cnt)
        arr_bkt = (C.c_void_p * world)(*st.peer_bkt)
        L.check(lib.hg_knn_sympart_sweep(C.byref(desc), L.ptr(thr), chunk, world, arr_cnt, arr_bkt, L.ptr(st.flags),
                                         L.ptr(ws), nbytes.value, stream))
        # behind this collective all pushes have landed, and every rank knows whether any bucket overflowed
        dist.all_reduce(st.flags, op=dist.ReduceOp.MAX, group=group)
        rows = r1 - r0
        out_idx = torch.empty((rows, k), dtype=torch.int32, device=dev)
        out_val = torch.empty((rows, k), dtype=torch.float32, device=dev)
        out_cos = torch.empty((rows, k), dtype=torch.float32, device=dev)
        L.check(lib.hg_knn_sympart_finish(C.byref(desc), L.ptr(feats), L.ptr(st.flags), st.cnt_ptr, st.bkt_ptr,
                                          L.ptr(out_idx), L.ptr(out_val), L.ptr(out_cos), L.ptr(ws), nbytes.value, stream))
    return out_idx, out_val, out_cos


def knn_hyperedges_row_partitioned(local_features: torch.Tensor, n: int, *, k: int = 16, kernel: str = "gaussian",
                                   lambda_h: float = 1.0, lambda_g: float = 0.0,
                                   positions: Optional[torch.Tensor] = None, spatial_k: int = 0,
                                   return_csr: bool = False, group=None, aligned: bool = False,
                                   symmetric: Optional[bool] = None,
                                   knn_fn: Optional[Callable] = None, assemble_fn: Optional[Callable] = None,
                                   spatial_fn: Optional[Callable] = None) -> dict:
    """Config C5: `local_features` is this rank's row block [r1-r0, D] (`row_range`, or
    `row_range_aligned` with aligned=True); `positions` (replicated, [n, 2|3]) is small.  Returns
    the same dict as `knn_hyperedges` on every rank.  With aligned=True the feature kNN runs in symmetric mode
    over the row blocks (see `_knn_sympart`; symmetric=False or HG_B200_SYMPART=0: classic pass); "mode" in the dict
    says which pass ran."""
    from . import ops
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    ranges = row_range_aligned if aligned else row_range
    r0, r1 = ranges(n, rank, world)
    assert local_features.shape[0] == r1 - r0, "local block does not match the row partition"
    pos = positions if (positions is not None) else None
    local = feats = None
    all_have_rows = all(ranges(n, r, world)[1] > ranges(n, r, world)[0] for r in range(world))
    # symmetric mode over aligned row blocks (default on CUDA; HG_B200_SYMPART=0 or symmetric=False: classic pass):
    # the ranks share the checkerboard of tile pairs instead of each computing every tile of its rows
    mode = "classic row blocks"
    if symmetric is None:
        symmetric = os.environ.get("HG_B200_SYMPART", "1") != "0"
    stride = 16 if n >= 32768 else 8
    symmetric = (symmetric and local is None and aligned and knn_fn is None and local_features.is_cuda and world > 1
                 and world <= 8 and all_have_rows and k <= 32 and n // stride >= 4 * (k + 8))
    if symmetric:
        feats = all_gather_rows(local_features, n, group, ranges)           # the one big collective
        local = _knn_sympart(feats, n, k, kernel, lambda_h, lambda_g, pos, group)
        if local is not None:
            mode = "symmetric tile pairs (buckets over NVLink)"
    if local is None:
        if feats is None:
            feats = all_gather_rows(local_features, n, group, ranges)       # the one big collective
        with_cos = knn_fn is None          # an injected knn_fn returns (idx, val)
        if knn_fn is None:
            knn_fn = lambda x, rows: ops.knn_feature(x, k, kernel, positions=pos if lambda_g else None,   # noqa: E731
                                                     lambda_h=lambda_h, lambda_g=lambda_g, rows=rows, return_cos=True)
        if r1 > r0:
            local = knn_fn(feats, (r0, r1))
        else:                              # more ranks than row blocks: nothing to rank here
            local = (torch.empty((0, k), dtype=torch.int32, device=feats.device),
                     torch.empty((0, k), dtype=torch.float32, device=feats.device))
            if with_cos:
                local += (torch.empty((0, k), dtype=torch.float32, device=feats.device),)
    idx_local, val_local = local[0], local[1]
    out = {"knn_idx": all_gather_rows(idx_local, n, group, ranges),
           "knn_val": all_gather_rows(val_local, n, group, ranges), "mode": mode}
    # cos(f_i, f_j) of every list entry travels with the lists: the edge weights are then looked up
    # instead of gathering two feature rows per edge
    knn_cos = all_gather_rows(local[2], n, group, ranges) if len(local) > 2 else None
    lists = [out["knn_idx"]]
    if spatial_k and pos is not None:
        if r1 > r0:
            sidx, sd2 = (spatial_fn or ops.knn_spatial)(pos, spatial_k, rows=(r0, r1))
        else:                              # a rank without rows still takes part in the gathers below
            sidx = torch.empty((0, spatial_k), dtype=torch.int32, device=feats.device)
            sd2 = torch.empty((0, spatial_k), dtype=torch.float32, device=feats.device)
        out["spatial_idx"] = all_gather_rows(sidx, n, group, ranges)
        out["spatial_d2"] = all_gather_rows(sd2, n, group, ranges)
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
