"""world_size-2 gloo test (CPU) of the view-sharding + flat gradient all-reduce host logic:
the reduced gradient must equal the single-process gradient of the MEAN loss over the same views
(reference semantics: src/training/trainer.py:458-500)."""
import importlib
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _loss(params, view):
    a, b = params
    return ((a * view["w"]).sum() + (b ** 2 * view["s"]).sum()) * view["k"]


def _views(n):
    g = torch.Generator().manual_seed(0)
    return [{"w": torch.randn(7, 3, generator=g), "s": torch.randn(5, generator=g), "k": float(i + 1)} for i in range(n)]


def _worker(rank, world, port, n_views, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg = importlib.import_module("3d-gaussian-splatting_b200")
    torch.manual_seed(1)
    params = [torch.randn(7, 3, requires_grad=True), torch.randn(5, requires_grad=True)]
    views = _views(n_views)
    assert pkg.shard_views(n_views, rank, world) == list(range(rank, n_views, world))
    step = pkg.ViewShardedStep(params)
    mean_loss = step(views, lambda v: _loss(params, v))
    out[rank] = (mean_loss, [p.grad.clone() for p in params])
    dist.destroy_process_group()


def test_view_sharded_gradients_equal_single_process_mean(tmp_path):
    world, n_views = 2, 5  # uneven shard: rank 0 gets 3 views, rank 1 gets 2
    port = 29500 + (os.getpid() % 500)
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, n_views, out), nprocs=world, join=True)
    torch.manual_seed(1)
    params = [torch.randn(7, 3, requires_grad=True), torch.randn(5, requires_grad=True)]
    views = _views(n_views)
    loss = sum(_loss(params, v) for v in views) / n_views
    loss.backward()
    for r in range(world):
        ml, grads = out[r]
        assert abs(ml - float(loss)) < 1e-5 * max(1.0, abs(float(loss)))
        for g, p in zip(grads, params):
            torch.testing.assert_close(g, p.grad, rtol=1e-5, atol=1e-6)
    # replicas identical
    for g0, g1 in zip(out[0][1], out[1][1]):
        assert torch.equal(g0, g1)


def test_shard_views_partition():
    sys.path.insert(0, ROOT)
    pkg = importlib.import_module("3d-gaussian-splatting_b200")
    for n in (0, 1, 7, 64):
        for w in (1, 2, 4, 8):
            allv = sorted(i for r in range(w) for i in pkg.shard_views(n, r, w))
            assert allv == list(range(n))
            sizes = [len(pkg.shard_views(n, r, w)) for r in range(w)]
            assert max(sizes) - min(sizes) <= 1


def test_balance_views_is_deterministic_capped_and_balanced():
    import importlib
    pkg = importlib.import_module("3d-gaussian-splatting_b200")
    costs = sorted([19.5, 17.8, 17.7, 18.9, 18.1, 19.9, 17.2, 18.4, 19.1, 17.9, 18.8, 18.0, 19.3, 17.5, 18.6, 18.2], reverse=True)
    parts = pkg.balance_views(costs, 4, per_rank=4)
    assert sorted(i for p in parts for i in p) == list(range(16)) and all(len(p) == 4 for p in parts)
    assert parts == pkg.balance_views(list(costs), 4, per_rank=4)
    loads = [sum(costs[i] for i in p) for p in parts]
    plain = [sum(costs[r * 4: r * 4 + 4]) for r in range(4)]
    assert max(loads) - min(loads) <= max(plain) - min(plain) and max(loads) - min(loads) < 1.0
    # uncapped: any split; over-subscribed cap: an error
    parts = pkg.balance_views([5, 1, 1, 1, 1, 1], 2)
    assert sorted(sum(([i] for p in parts for i in p), [])) == list(range(6)) and {len(p) for p in parts} == {1, 5}
    import pytest as _pt
    with _pt.raises(ValueError):
        pkg.balance_views([1, 2, 3], 2, per_rank=1)


def _worker_flat(rank, world, port, n_views, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg = importlib.import_module("3d-gaussian-splatting_b200")
    torch.manual_seed(1)
    params = [torch.randn(7, 3, requires_grad=True), torch.randn(5, requires_grad=True)]
    views = _views(n_views)
    fg = pkg.FlatGradients(params)
    grads = []
    for _ in range(2):  # two steps: begin_step must really restart from zero
        fg.begin_step()
        for i in pkg.shard_views(n_views, rank, world):
            _loss(params, views[i]).backward()
        fg.allreduce(n_views)
        assert all(p.grad.data_ptr() == v.data_ptr() for p, v in zip(params, fg.views))
        grads.append([p.grad.clone() for p in params])
    out[rank] = grads
    dist.destroy_process_group()


def test_flat_gradients_one_collective_equals_single_process_mean():
    world, n_views = 2, 5
    port = 29000 + (os.getpid() % 500)
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_flat, args=(world, port, n_views, out), nprocs=world, join=True)
    torch.manual_seed(1)
    params = [torch.randn(7, 3, requires_grad=True), torch.randn(5, requires_grad=True)]
    (sum(_loss(params, v) for v in _views(n_views)) / n_views).backward()
    for r in range(world):
        for step in range(2):
            for g, p in zip(out[r][step], params):
                torch.testing.assert_close(g, p.grad, rtol=1e-5, atol=1e-6)
