"""Host-side multi-GPU logic on CPU: world_size-2 gloo processes (no GPU needed).
The CUDA stages are injected with the oracle so only the sharding / gather plumbing is under test."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_np as oracle
from hg_b200 import multigpu, synth


def test_shard_slides_lpt_balance():
    sizes = synth.slide_sizes(512)
    assert sizes.min() >= 2048 and sizes.max() <= 8192
    for world in (1, 2, 4, 8):
        plan = multigpu.shard_slides(sizes, world)
        flat = sorted(i for p in plan for i in p)
        assert flat == list(range(512))                       # every slide exactly once
        load = [sum(int(sizes[i]) ** 2 for i in p) for p in plan]
        assert max(load) / (sum(load) / world) < 1.01         # LPT: < 1 % imbalance on 512 slides


def test_row_range_partition():
    for n, world in [(250_000, 8), (10, 3), (7, 8), (1, 2)]:
        edges = [multigpu.row_range(n, r, world) for r in range(world)]
        assert edges[0][0] == 0 and edges[-1][1] == n
        assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
        lens = [b - a for a, b in edges]
        assert max(lens) - min(lens) <= 1


def test_row_range_aligned_partition():
    for n, world in [(250_000, 8), (20_001, 4), (300, 4), (1, 2), (512, 2), (257, 3)]:
        edges = [multigpu.row_range_aligned(n, r, world) for r in range(world)]
        assert edges[0][0] == 0 and edges[-1][1] == n
        assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
        assert all(a % 256 == 0 or a == n for a, _ in edges)          # panel boundaries of the fused kernel
        assert all(b % 256 == 0 or b == n for _, b in edges)
        lens = [b - a for a, b in edges]
        assert max(lens) - (-(-n // world)) < 256                      # at most one tile more than an even share


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, d, k, q, aligned=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        feats = synth.make_features(n, d, 77, "gaussian")
        r0, r1 = (multigpu.row_range_aligned if aligned else multigpu.row_range)(n, rank, world)

        def knn_fn(x, rows):                                   # stands in for ops.knn_feature on this rank's rows
            idx, val = oracle.knn_feature(x.numpy(), k, "gaussian", rows=np.arange(rows[0], rows[1]))
            return torch.from_numpy(idx.astype(np.int32)), torch.from_numpy(val.astype(np.float32))

        def assemble_fn(lists, x):
            ei = oracle.edges_from_knn(lists[0].numpy())
            return {"edge_index": torch.from_numpy(ei)}

        out = multigpu.knn_hyperedges_row_partitioned(torch.from_numpy(feats[r0:r1]), n, k=k, knn_fn=knn_fn,
                                                      assemble_fn=assemble_fn, aligned=aligned)
        assert out["mode"] == "classic row blocks"             # no CUDA here: gather-then-compute
        q.put((rank, out["knn_idx"].numpy(), out["edge_index"].numpy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(240)
@pytest.mark.parametrize("n,aligned", [(301, False), (301, True), (200, True)])
def test_row_partitioned_equals_single_process(n, aligned):
    # odd n: uneven row blocks -> all_gather of ragged chunks; aligned: blocks of 256 rows, with n = 200
    # the second rank owns nothing
    d, k, world = 32, 4, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, d, k, q, aligned)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=100) for _ in range(world)]
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    feats = synth.make_features(n, d, 77, "gaussian")
    ref_idx, _ = oracle.knn_feature(feats, k, "gaussian")
    ref_ei = oracle.edges_from_knn(ref_idx)
    for _, idx, ei in results:
        assert (idx == ref_idx).all()
        assert (ei == ref_ei).all()


def _worker_spatial(rank, world, port, n, d, k, ks, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        feats, pos = synth.make_slide(n, d, 78, "gaussian")
        r0, r1 = multigpu.row_range_aligned(n, rank, world)

        def knn_fn(x, rows):
            idx, val = oracle.knn_feature(x.numpy(), k, "gaussian", rows=np.arange(rows[0], rows[1]))
            return torch.from_numpy(idx.astype(np.int32)), torch.from_numpy(val.astype(np.float32))

        def spatial_fn(p, kk, rows):                          # stands in for ops.knn_spatial: needs a non-empty range
            assert rows[0] < rows[1]
            idx, d2 = oracle.knn_spatial(p.numpy(), kk, rows=np.arange(rows[0], rows[1]))
            return torch.from_numpy(idx.astype(np.int32)), torch.from_numpy(d2.astype(np.float32))

        def assemble_fn(lists, x):
            return {"edge_index": torch.from_numpy(oracle.edges_from_knn(np.concatenate([t.numpy() for t in lists], 1)))}

        out = multigpu.knn_hyperedges_row_partitioned(torch.from_numpy(feats[r0:r1]), n, k=k, knn_fn=knn_fn,
                                                      positions=torch.from_numpy(pos), spatial_k=ks, spatial_fn=spatial_fn,
                                                      assemble_fn=assemble_fn, aligned=True)
        q.put((rank, out["spatial_idx"].numpy(), out["edge_index"].numpy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_rank_without_rows_still_joins_the_spatial_gathers():
    """n = 200 in 256-row blocks on 2 ranks: rank 1 owns nothing.  With spatial_k > 0 it used to call the spatial
    kNN on an empty row range (ValueError) while rank 0 waited in the all-gather."""
    n, d, k, ks, world = 200, 16, 4, 3, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_spatial, args=(r, world, port, n, d, k, ks, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=100) for _ in range(world)]
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    feats, pos = synth.make_slide(n, d, 78, "gaussian")
    ref_idx, _ = oracle.knn_feature(feats, k, "gaussian")
    ref_s, _ = oracle.knn_spatial(pos, ks)
    ref_ei = oracle.edges_from_knn(np.concatenate([ref_idx, ref_s], 1))
    for _, sidx, ei in results:
        assert (sidx == ref_s).all() and (ei == ref_ei).all()


# ---- dataset loop (preprocess_hypergraph._run_over_csv): read-ahead / writer threads, rank sharding --------
def _dataset_files(root, sizes):
    names = []
    for i, sz in enumerate(sizes):
        name = f"slide_{i}.bin"
        with open(os.path.join(root, name), "wb") as f:
            f.write(b"x" * sz)
        names.append(name)
    csv = os.path.join(root, "cases.csv")
    with open(csv, "w") as f:
        f.write("case_id,h5_file_path\n" + "\n".join(f"c{i},{n}" for i, n in enumerate(names + ["missing.bin"])) + "\n")
    return csv, names


def _fake_worker(log):
    def worker(path, inputs=None, writer=None):
        if path.endswith("slide_2.bin"):
            raise RuntimeError("boom")                              # a failing slide is reported and skipped
        stats = {"bytes": len(inputs) if inputs is not None else os.path.getsize(path)}
        log.append(os.path.basename(path))
        if writer is not None:
            def save():
                if path.endswith("slide_4.bin"):
                    raise OSError("disk full")                      # so is a slide whose save fails
            stats["_pending_save"] = writer.submit(save)
        return stats
    return worker


def test_dataset_loop_reads_ahead_and_keeps_the_error_convention(tmp_path):
    import json
    from hg_b200 import preprocess_hypergraph as pp
    csv, names = _dataset_files(str(tmp_path), [100, 400, 50, 300, 200, 10])
    for overlap in (True, False):
        log = []
        out = str(tmp_path / f"stats_{overlap}.json")
        loader = (lambda p: open(p, "rb").read()) if overlap else None
        res = pp._run_over_csv(csv, str(tmp_path), "test", _fake_worker(log), out, loader=loader)
        want = ["c0", "c1", "c3", "c5"] if overlap else ["c0", "c1", "c3", "c4", "c5"]   # c4 fails only in its save
        assert [r["case_id"] for r in res] == want
        assert [r["bytes"] for r in res][:2] == [100, 400]
        assert all("_pending_save" not in r for r in res)
        assert len(json.load(open(out))) == len(want)
        assert log == [n for n in names if n != "slide_2.bin"]      # CSV order, every present file once


def test_writer_never_holds_more_than_two_pending_saves():
    """saves slower than the slide loop: submit() blocks instead of queueing device tensors without bound"""
    import threading
    import time
    from hg_b200 import preprocess_hypergraph as pp
    w = pp._BoundedWriter(max_pending=2)
    live, peak, lock = [0], [0], threading.Lock()

    def slow_save():
        time.sleep(0.02)
        with lock:
            live[0] -= 1

    futs = []
    for _ in range(12):
        with lock:                                   # "a slide's tensors are alive" from submit until its save has run
            live[0] += 1
            peak[0] = max(peak[0], live[0])
        futs.append(w.submit(slow_save))
    for f in futs:
        f.result()
    w.shutdown()
    assert peak[0] <= 3                              # two pending + the one being handed over


def _dataset_rank(rank, world, port, root, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from hg_b200 import preprocess_hypergraph as pp
        log = []
        res = pp._run_over_csv(os.path.join(root, "cases.csv"), root, "test", _fake_worker(log),
                               os.path.join(root, "stats.json"), loader=lambda p: open(p, "rb").read())
        q.put((rank, log, [r["case_id"] for r in res]))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_dataset_loop_shards_slides_over_ranks(tmp_path):
    import json
    sizes = [100, 400, 50, 300, 200, 10, 350]
    csv, names = _dataset_files(str(tmp_path), sizes)
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dataset_rank, args=(r, world, port, str(tmp_path), q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted(q.get(timeout=100) for _ in range(world))
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    logs = [set(r[1]) for r in results]
    assert not (logs[0] & logs[1])                                   # every slide on exactly one rank
    assert logs[0] | logs[1] == {n for n in names if n != "slide_2.bin"}
    plan = multigpu.shard_slides(sizes, world)
    assert logs[0] == {names[i] for i in plan[0] if i != 2}          # the LPT share of rank 0
    want = ["c0", "c1", "c3", "c5", "c6"]                            # c2 fails in the worker, c4 in its save
    assert all(r[2] == want for r in results)                        # every rank returns the gathered statistics
    assert [s["case_id"] for s in json.load(open(tmp_path / "stats.json"))] == want
