"""Stream-ordered build (hg_build_knn_hypergraph_enqueue): same outputs as the synchronising entry
point bit for bit, also with many slides in flight on several streams (build_batch, pipeline lanes)."""
import numpy as np
import pytest
import torch

import hg_b200
from hg_b200 import ops, synth
from hg_b200.pipeline import SlidePipeline, build_batch, build_many
from test_gpu_parity import _cuda

pytestmark = pytest.mark.gpu
KEYS = ("knn_idx", "knn_val", "edge_index", "edge_weights", "csr_rowptr", "csr_col", "csr_eid")


def _equal(a: dict, b: dict, keys=KEYS):
    for name in keys:
        if name in a or name in b:
            ta, tb = a[name].cpu(), b[name].cpu()
            assert ta.shape == tb.shape, name
            assert torch.equal(ta, tb) or torch.allclose(ta.float(), tb.float(), rtol=0, atol=0, equal_nan=True), name


@pytest.mark.parametrize("n,d,k,metric,spatial_k,cliques", [
    (3000, 256, 8, "gaussian", 0, False),
    (5000, 128, 16, "cosine", 0, False),
    (9000, 64, 8, "gaussian", 8, False),          # spatial grid lists join the edge set
    (2500, 96, 5, "euclidean", 0, True),          # cluster cliques
    (14000, 128, 16, "gaussian", 0, False),       # symmetric tensor-core pass
])
def test_enqueue_equals_synchronous_build(n, d, k, metric, spatial_k, cliques):
    feats, pos = synth.make_slide(n, d, 31 + n, "cosine" if metric == "cosine" else "gaussian")
    x, p = _cuda(feats), (_cuda(pos) if spatial_k else None)
    kw = dict(positions=p, lambda_h=1.0, lambda_g=(1.0 / 128 if spatial_k else 0.0), spatial_k=spatial_k, return_csr=True)
    if cliques:
        labels = torch.from_numpy(np.random.default_rng(3).integers(0, 40, n).astype(np.int32))
        kw.update(labels=labels, num_hyperedges=12)
    ref = ops.build_knn_hypergraph(x, k, metric, **kw)
    pend = ops.build_knn_hypergraph(x, k, metric, enqueue_only=True, **kw)
    assert isinstance(pend, ops.PendingBuild)
    out = pend.result()
    _equal(out, ref, KEYS + (("spatial_idx", "spatial_d2") if spatial_k else ()))
    assert out["edge_index"].shape[1] * 2 == out["csr_col"].numel() == int(out["csr_rowptr"][-1])


def test_build_batch_many_small_slides_on_four_streams():
    sizes = synth.slide_sizes(24, seed=5, lo=300, hi=5000)
    slides = [_cuda(synth.make_features(int(s), 128, 100 + i, "gaussian")) for i, s in enumerate(sizes)]
    kw = dict(k=8, kernel="gaussian", lambda_h=1.0, return_csr=True)
    outs = build_batch(slides, lanes=4, **kw)
    torch.cuda.synchronize()
    assert len(outs) == len(slides)
    for x, out in zip(slides, outs):
        _equal(out, hg_b200.knn_hyperedges(x, **kw))
    again = build_batch(slides, lanes=3, **kw)      # streams and workspaces reused
    for a, b in zip(outs, again):
        _equal(a, b)


@pytest.mark.parametrize("group_rows,with_pos", [(4096, False), (128 * 1024, False), (16384, True)])
def test_build_batch_grouped_front_half_equals_per_slide_builds(group_rows, with_pos):
    """hg_build_batch_grouped_enqueue: the slides of a group share one operand buffer and ONE launch of the fused
    kernel (unit table); small groups force many groups, alternating front workspaces and single-slide groups; the
    9,500-row slide takes the per-slide (symmetric) path in the middle of the batch.  Bit-equal to the ungrouped
    batch and to one-by-one builds."""
    sizes = [300, 5000, 2049, 9500, 256, 257, 4096, 8192, 1000, 3000, 2, 700, 6000, 2304]
    slides = []
    for i, s in enumerate(sizes):
        f, p = synth.make_slide(int(s), 128, 300 + i, "gaussian")
        slides.append((_cuda(f), _cuda(p)) if with_pos else _cuda(f))
    kw = dict(k=16, kernel="gaussian", lambda_h=1.0, return_csr=True)
    keys = KEYS
    if with_pos:
        kw.update(lambda_g=1.0 / 128, spatial_k=4)
        keys = KEYS + ("spatial_idx", "spatial_d2")
    grouped = build_batch(slides, lanes=3, grouped=True, group_rows=group_rows, **kw)
    from hg_b200 import pipeline
    assert pipeline.LAST_BATCH_STATS["grouped"] is True
    plain = build_batch(slides, lanes=3, grouped=False, **kw)
    torch.cuda.synchronize()
    for a, b in zip(grouped, plain):
        _equal(a, b, keys)
    for i in (0, 3, 7, 10):
        x, p = slides[i] if with_pos else (slides[i], None)
        _equal(grouped[i], hg_b200.knn_hyperedges(x, p, **kw), keys)
    again = build_batch(slides, lanes=2, grouped=True, group_rows=group_rows, **kw)     # workspaces reused
    for a, b in zip(grouped, again):
        _equal(a, b, keys)


@pytest.mark.parametrize("lanes", [1, 3])
def test_pipeline_lanes_from_host_memory(lanes):
    sizes = [700, 4100, 900, 2600, 3300, 512, 6000]
    hosts = [torch.from_numpy(synth.make_features(s, 64, 7 + s, "gaussian")).pin_memory() for s in sizes]
    kw = dict(k=8, kernel="gaussian", lambda_h=1.0, return_csr=True)
    pipe = SlidePipeline(lanes=lanes, keys=KEYS, **kw)
    got = []
    for res in build_many(iter(hosts), pipeline=pipe):
        got.append({name: t.clone() for name, t in res.items()})      # the pinned slots rotate
    assert len(got) == len(hosts)
    for h, res in zip(hosts, got):
        _equal(res, hg_b200.knn_hyperedges(h.cuda(), **kw))
    assert pipe.h2d_bytes == sum(s * 64 * 4 for s in sizes) and pipe.d2h_bytes > 0


def test_enqueue_edge_cases():
    x = _cuda(synth.make_features(2, 32, 1, "gaussian"))                 # two nodes: one edge
    out = ops.build_knn_hypergraph(x, 4, "euclidean", return_csr=True, enqueue_only=True).result()
    assert out["edge_index"].tolist() == [[0], [1]]
    assert out["csr_rowptr"].tolist() == [0, 1, 2] and out["csr_col"].tolist() == [1, 0]
    with pytest.raises(ValueError):
        ops.build_knn_hypergraph(x[:1], 4, "euclidean", enqueue_only=True)


def test_build_batch_with_positions_and_threads():
    sizes = [1500, 9000, 2200, 400]
    slides = []
    for i, s in enumerate(sizes):
        f, p = synth.make_slide(s, 96, 50 + i, "gaussian")
        slides.append((_cuda(f), _cuda(p)))
    kw = dict(k=8, kernel="gaussian", lambda_h=1.0, lambda_g=1.0 / 128, spatial_k=8, return_csr=True)
    outs = build_batch(slides, lanes=2, threads=2, **kw)
    keys = KEYS + ("spatial_idx", "spatial_d2")
    for (x, p), out in zip(slides, outs):
        _equal(out, hg_b200.knn_hyperedges(x, p, **kw), keys)
    with pytest.raises(ValueError):
        build_batch([slides[0][0], _cuda(synth.make_features(100, 32, 1))], k=4)      # mixed feature dimensions
    assert len(build_batch([], k=4)) == 0
