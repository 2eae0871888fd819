"""Run under torchrun on >= 2 GPUs: row-partitioned build (NCCL all-gather of feature row blocks,
per-rank row range through the fused kernel, all-gather of the lists) must equal the single-GPU
build bit for bit.  Launched by tests/test_multigpu_nccl.py."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "multimodal-fusion_b200"))
import hg_b200                                   # noqa: E402
from hg_b200 import multigpu, synth              # noqa: E402


def _lib_err():
    from hg_b200 import _lib
    return _lib.last_error()


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    n, d, k = 20_001, 256, 16                    # odd n: ragged row blocks
    feats, pos = synth.make_slide(n, d, 99, "gaussian")
    lam_g = 1.0 / 128
    x, p = torch.from_numpy(feats).to(dev), torch.from_numpy(pos).to(dev)
    r0, r1 = multigpu.row_range(n, rank, world)
    out = multigpu.knn_hyperedges_row_partitioned(x[r0:r1].contiguous(), n, k=k, kernel="gaussian", lambda_h=1.0,
                                                  lambda_g=lam_g, positions=p, spatial_k=8, return_csr=True)
    ref = hg_b200.knn_hyperedges(x, p, k=k, kernel="gaussian", lambda_h=1.0, lambda_g=lam_g, spatial_k=8,
                                 return_csr=True)
    for key in ("knn_idx", "knn_val", "spatial_idx", "spatial_d2", "edge_index", "edge_weights", "csr_rowptr",
                "csr_col", "csr_eid"):
        assert torch.equal(out[key], ref[key]), f"rank {rank}: {key} differs"
    # 256-aligned row blocks: classic pass, then the symmetric pass (tile pairs shared by the ranks, buckets over
    # NVLink) twice in a row on the same exchange buffers; same result bit for bit
    a0, a1 = multigpu.row_range_aligned(n, rank, world)
    modes = []
    for symmetric in (False, True, True):
        out = multigpu.knn_hyperedges_row_partitioned(x[a0:a1].contiguous(), n, k=k, kernel="gaussian", lambda_h=1.0,
                                                      lambda_g=lam_g, positions=p, spatial_k=8, return_csr=True,
                                                      aligned=True, symmetric=symmetric)
        for key in ("knn_idx", "knn_val", "spatial_idx", "spatial_d2", "edge_index", "edge_weights", "csr_rowptr",
                    "csr_col", "csr_eid"):
            assert torch.equal(out[key], ref[key]), f"rank {rank}: {key} differs (aligned, symmetric={symmetric})"
        modes.append(out["mode"])
        torch.cuda.synchronize()
        if rank == 0 and os.environ.get("HG_NCCL_CHECK_VERBOSE"):
            print(f"phase ok: symmetric={symmetric} mode={out['mode']}", flush=True)
    assert modes[0].startswith("classic") and modes[1].startswith("symmetric") and modes[2].startswith("symmetric"), (modes, _lib_err())
    E_small = int(ref["edge_index"].shape[1])
    dist.barrier()
    # the same at C5's width and beyond C3's row count: n = 100,003 x 1536 (ragged blocks), every rank generates
    # the whole slide from the seed, ranks > 0 drop everything but their block before the build
    big = None
    if os.environ.get("HG_NCCL_CHECK_BIG", "1") == "1":
        del x, p, out, ref
        nb, db = 100_003, 1536
        fb = torch.from_numpy(synth.make_features(nb, db, 1240, "gaussian")).to(dev)
        b0, b1 = multigpu.row_range(nb, rank, world)
        refb = hg_b200.knn_hyperedges(fb, k=k, kernel="gaussian", lambda_h=1.0, return_csr=True)     # symmetric single-GPU pass
        outb = multigpu.knn_hyperedges_row_partitioned(fb[b0:b1].contiguous(), nb, k=k, kernel="gaussian", lambda_h=1.0,
                                                       return_csr=True)                                # ragged blocks: classic
        c0, c1 = multigpu.row_range_aligned(nb, rank, world)
        outs = multigpu.knn_hyperedges_row_partitioned(fb[c0:c1].contiguous(), nb, k=k, kernel="gaussian", lambda_h=1.0,
                                                       return_csr=True, aligned=True)                  # aligned: symmetric
        assert outs["mode"].startswith("symmetric"), outs["mode"]
        for key in ("knn_idx", "knn_val", "edge_index", "edge_weights", "csr_rowptr", "csr_col", "csr_eid"):
            assert torch.equal(outb[key], refb[key]), f"rank {rank}: {key} differs at n = {nb}"
            assert torch.equal(outs[key], refb[key]), f"rank {rank}: {key} differs at n = {nb} (symmetric tile pairs)"
        big = (nb, db, int(refb["edge_index"].shape[1]))
        dist.barrier()
    if rank == 0:
        print(f"NCCL_ROWPART_OK world={world} n={n} E={E_small} modes={modes} big={big}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
