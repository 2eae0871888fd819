"""float64 checker for the kNN lists, evaluated with torch ON THE GPU (test infrastructure only).

The numpy oracle (oracle/oracle_np.py) needs minutes per thousand rows at N = 50k..250k; a B200
evaluates the same float64 scores for every row of C2 in seconds, so the full-size parity tests can
check ALL rows instead of a 1 % sample.  Formulas follow the oracle's (and sklearn's) expansion form
d2 = |x_i|^2 + |x_j|^2 - 2 x_i.x_j in float64, clamped at 0 (sklearn _argkmin / reference
preprocess_hypergraph.py:379-382); Gaussian scores add lambda_g |p_i - p_j|^2
(similarity_kernel.py:79-84, 122); cosine scores are -cos.  Self is excluded by index.

check_knn_fp64 returns a report: every row whose neighbour SET differs from the float64 set is
listed with the gap between the swapped candidates; a row is a near-tie when that gap is at most
1e-5 relative (north_star tolerance), anything else is wrong.
"""
from __future__ import annotations

import numpy as np
import torch

RTOL = 1e-5


def _scores(x64, sq, rows, metric, lambda_h, lambda_g, p64):
    dot = x64[rows] @ x64.T
    if metric == "cosine":
        nrm = sq.sqrt().clamp_min(1e-8)
        s = -(dot / nrm[rows, None]) / nrm[None, :]
    else:
        s = (sq[rows, None] + sq[None, :] - 2.0 * dot).clamp_min_(0.0)
        if metric == "gaussian":
            s *= lambda_h
            if p64 is not None and lambda_g:
                for c in range(p64.shape[1]):
                    df = p64[rows, c][:, None] - p64[None, :, c]
                    s += lambda_g * df * df
    s[torch.arange(rows.numel(), device=s.device), rows] = float("inf")
    return s


def check_knn_fp64(x: torch.Tensor, idx: torch.Tensor, val: torch.Tensor, rows: torch.Tensor, metric: str,
                   lambda_h: float = 1.0, lambda_g: float = 0.0, pos: torch.Tensor = None, chunk: int = 2048) -> dict:
    """x [N, D] float32 CUDA; idx / val [R, k] = the lists of global rows `rows` [R] (CUDA long)."""
    dev = x.device
    k = idx.shape[1]
    x64 = x.double()
    sq = (x64 * x64).sum(1)
    p64 = pos.double() if pos is not None else None
    near, wrong = [], []
    max_rel = 0.0
    not_sorted = 0
    for a in range(0, rows.numel(), chunk):
        rr = rows[a:a + chunk]
        s = _scores(x64, sq, rr, metric, lambda_h, lambda_g, p64)
        ref_s, ref_i = torch.topk(s, k, dim=1, largest=False)
        got_i = idx[a:a + chunk].long()
        got_s = torch.gather(s, 1, got_i.clamp_min(0))
        assert bool((got_i >= 0).all()) and bool((got_i != rr[:, None]).all()), "missing / self entries in the lists"
        # values against float64 (no absolute term): Gaussian K, cosine, or squared distance
        ref_v = {"gaussian": torch.exp(-got_s), "cosine": -got_s, "euclidean": got_s}[metric]
        rel = ((val[a:a + chunk].double() - ref_v).abs() / ref_v.abs().clamp_min(1e-300))
        max_rel = max(max_rel, float(rel.max()))
        # the returned order is by exact FP32 value: in float64 it may only be violated inside the tolerance
        not_sorted += int(((got_s[:, 1:] - got_s[:, :-1]) < -RTOL * got_s.abs().amax(1, keepdim=True).clamp_min(1.0)).any(1).sum())
        diff = (torch.sort(got_i, 1).values != torch.sort(ref_i, 1).values).any(1)
        for r in torch.nonzero(diff).flatten().tolist():
            g, o = set(got_i[r].tolist()), set(ref_i[r].tolist())
            only_g = torch.tensor(sorted(g - o), device=dev)
            only_o = torch.tensor(sorted(o - g), device=dev)
            sg, so = s[r, only_g], s[r, only_o]
            scale = max(1.0, float(so.abs().max()))
            gap = float((sg.max() - so.min()).abs()) / scale
            rec = (int(rr[r]), gap, only_g.tolist(), only_o.tolist())
            (near if gap <= RTOL else wrong).append(rec)
        del s
    return {"rows": int(rows.numel()), "near_ties": near, "wrong": wrong, "max_rel_val": max_rel, "unsorted_rows": not_sorted}


def assert_knn_fp64(report: dict, label: str, max_near_tie_frac: float = 2e-3) -> None:
    near, wrong = report["near_ties"], report["wrong"]
    print(f"[fp64 check] {label}: {report['rows']} rows, {len(near)} near-tie rows, {len(wrong)} wrong, "
          f"max rel value error {report['max_rel_val']:.2e}")
    for row, gap, g, o in near[:20]:
        print(f"    near-tie row {row}: gap {gap:.2e} relative, lists hold {g} where float64 has {o}")
    assert not wrong, f"{label}: rows beyond the near-tie tolerance: {wrong[:5]}"
    assert len(near) <= max(2, int(max_near_tie_frac * report["rows"])), f"{label}: {len(near)} near-tie rows"
    assert report["max_rel_val"] <= RTOL, f"{label}: value error {report['max_rel_val']:.2e} > 1e-5 relative"
    assert report["unsorted_rows"] == 0, f"{label}: {report['unsorted_rows']} rows not ordered best-first"
