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


class _TwoLayerNode(torch.autograd.Function):
    """A hand-sequenced node like the ResNet-50 backbone (one autograd node, explicit backward): it hands its parameter
    gradients to engine.deliver_grads stage by stage and returns whatever the sink left in the dict."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2):
        h = torch.relu(x @ w1.t() + b1)
        ctx.save_for_backward(x, h, w1, w2)
        ctx.params = (w1, b1, w2, b2)
        return h @ w2.t() + b2

    @staticmethod
    def backward(ctx, dy):
        from pytorch_segmentation_b200 import engine as E
        x, h, w1, w2 = ctx.saved_tensors
        p1, pb1, p2, pb2 = ctx.params
        grads = {p2: dy.t() @ h, pb2: dy.sum(0)}
        E.deliver_grads(grads, [p2, pb2])                       # second layer done: its all-reduce may start
        dh = (dy @ w2) * (h > 0)
        grads[p1], grads[pb1] = dh.t() @ x, dh.sum(0)
        E.deliver_grads(grads, [p1, pb1])
        return (dh @ w1,) + tuple(grads[p] for p in (p1, pb1, p2, pb2))


def _early_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from pytorch_segmentation_b200.parallel import GradBucketer
    torch.manual_seed(0)
    head = torch.nn.Linear(32, 4)
    w1, b1 = torch.nn.Parameter(torch.randn(256, 64) * 0.1), torch.nn.Parameter(torch.zeros(256))
    w2, b2 = torch.nn.Parameter(torch.randn(32, 256) * 0.1), torch.nn.Parameter(torch.zeros(32))
    params = [w1, b1, w2, b2] + list(head.parameters())
    bucketer = GradBucketer(params, bucket_mb=0.02, first_bucket_mb=0.001).attach()
    g = torch.Generator().manual_seed(200 + rank)
    res = {'grads': [], 'early': []}
    for it in range(2):
        x, y = torch.randn(8, 64, generator=g), torch.randint(0, 4, (8,), generator=g)
        for p in params:
            p.grad = None
        torch.nn.functional.cross_entropy(head(_TwoLayerNode.apply(x, w1, b1, w2, b2)), y).backward()
        res['early'].append(len(bucketer._early))
        bucketer.finish()
        res['grads'].append([p.grad.clone() for p in params])
        res.setdefault('xy', []).append((x, y))
    bucketer.remove()
    torch.save(res, os.path.join(out, 'e%d.pt' % rank))
    dist.destroy_process_group()


def test_early_gradient_delivery_world2(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_early_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(os.path.join(str(tmp_path), 'e%d.pt' % i)) for i in range(world)]
    assert r[0]['early'] == [4, 4]                              # the node's four parameters bypassed autograd's hooks
    torch.manual_seed(0)
    head = torch.nn.Linear(32, 4)
    w1, b1 = torch.nn.Parameter(torch.randn(256, 64) * 0.1), torch.nn.Parameter(torch.zeros(256))
    w2, b2 = torch.nn.Parameter(torch.randn(32, 256) * 0.1), torch.nn.Parameter(torch.zeros(32))
    params = [w1, b1, w2, b2] + list(head.parameters())
    for it in range(2):
        want = None
        for rank in range(world):
            x, y = r[rank]['xy'][it]
            for p in params:
                p.grad = None
            hmid = torch.relu(x @ w1.t() + b1) @ w2.t() + b2       # plain autograd reference
            torch.nn.functional.cross_entropy(head(hmid), y).backward()
            gs = [p.grad.clone() for p in params]
            want = gs if want is None else [a + b for a, b in zip(want, gs)]
        want = [g / world for g in want]
        for rank in range(world):
            for a, b in zip(r[rank]['grads'][it], want):
                assert torch.allclose(a, b, rtol=1e-5, atol=1e-6)


def _sync_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from pytorch_segmentation_b200.parallel import GradBucketer
    torch.manual_seed(1000 + rank)                            # DIFFERENT weights and statistics on every rank
    model = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3), torch.nn.BatchNorm2d(8), torch.nn.Conv2d(8, 4, 1))
    model[1].running_mean.normal_()
    model[1].num_batches_tracked.fill_(rank + 5)
    model[2].weight.requires_grad_(False)                     # a frozen parameter is not in the bucketer but must sync too
    before = {k: v.clone() for k, v in model.state_dict().items()}
    GradBucketer(model.parameters(), bucket_mb=0.01).sync_initial_state(model)
    torch.save({'before': before, 'after': {k: v.clone() for k, v in model.state_dict().items()}},
               os.path.join(out, 's%d.pt' % rank))
    dist.destroy_process_group()


def test_sync_initial_state_world2(tmp_path):
    """GradBucketer.sync_initial_state = DDP's construction-time broadcast: every rank ends up with rank 0's parameters
    AND buffers (BN running statistics, num_batches_tracked), whatever it was initialised with."""
    world, port = 2, _free_port()
    mp.spawn(_sync_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(os.path.join(str(tmp_path), 's%d.pt' % i)) for i in range(world)]
    assert any(not torch.equal(r[0]['before'][k], r[1]['before'][k]) for k in r[0]['before'])
    for k, v in r[0]['before'].items():
        assert torch.equal(r[0]['after'][k], v), k            # rank 0 keeps its state
        assert torch.equal(r[1]['after'][k], v), k            # rank 1 receives it
