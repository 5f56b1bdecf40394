"""World-size-2 test of the data-parallel plumbing on CPU (gloo): bucketed gradient all-reduce overlapped with
backward gives every rank the average gradient, and the epoch-end confusion-matrix reduction sums exactly."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from pytorch_segmentation_b200.parallel import GradBucketer
    from pytorch_segmentation_b200.utils.utils import counts_from_confusion
    torch.manual_seed(0)                                      # same weights on every rank
    model = torch.nn.Sequential(torch.nn.Linear(64, 512), torch.nn.ReLU(), torch.nn.Linear(512, 512), torch.nn.ReLU(),
                                torch.nn.Linear(512, 8))
    bucketer = GradBucketer(model.parameters(), bucket_mb=0.5, first_bucket_mb=0.01)
    g = torch.Generator().manual_seed(100 + rank)             # per-rank shard of the batch
    xs = [torch.randn(16, 64, generator=g) for _ in range(2)]
    ys = [torch.randint(0, 8, (16,), generator=g) for _ in range(2)]
    res = {}
    for it in range(2):                                       # second step: grads are already bucket views
        model.zero_grad(set_to_none=(it == 0))
        if it == 1:
            bucketer.zero_grad()
        loss = torch.nn.functional.cross_entropy(model(xs[it]), ys[it])
        loss.backward()
        bucketer.finish()
        res['grads%d' % it] = [p.grad.clone() for p in model.parameters()]
    res['nb'] = bucketer.num_buckets
    # eval-side collective: one all-reduce of the int64 confusion matrix (replaces test.py:55-57)
    nc = 5
    pred = torch.randint(0, nc, (1000,), generator=g)
    tgt = torch.randint(0, nc, (1000,), generator=g)
    cm = torch.bincount(tgt * nc + pred, minlength=nc * nc).reshape(nc, nc)
    local = cm.clone()
    dist.all_reduce(cm, op=dist.ReduceOp.SUM)
    res['cm_local'], res['cm'] = local, cm
    res['xs'], res['ys'] = xs, ys
    torch.save(res, os.path.join(out, 'r%d.pt' % rank))
    dist.destroy_process_group()


def test_bucketed_allreduce_world2(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(os.path.join(str(tmp_path), 'r%d.pt' % i)) for i in range(world)]
    assert r[0]['nb'] >= 3
    # reference: single process, average of the two shards' gradients
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(64, 512), torch.nn.ReLU(), torch.nn.Linear(512, 512), torch.nn.ReLU(),
                                torch.nn.Linear(512, 8))
    for it in range(2):
        want = None
        for rank in range(world):
            model.zero_grad()
            torch.nn.functional.cross_entropy(model(r[rank]['xs'][it]), r[rank]['ys'][it]).backward()
            gs = [p.grad.clone() for p in model.parameters()]
            want = gs if want is None else [a + b for a, b in zip(want, gs)]
        want = [g / world for g in want]
        for rank in range(world):
            for a, b in zip(r[rank]['grads%d' % it], want):
                assert torch.allclose(a, b, rtol=1e-5, atol=1e-7)
    assert torch.equal(r[0]['cm'], r[0]['cm_local'] + r[1]['cm_local'])
    assert torch.equal(r[0]['cm'], r[1]['cm'])
