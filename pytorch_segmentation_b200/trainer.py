"""One optimisation step of the hot path, optionally replayed from a CUDA graph.

The reference's step is ``outputs = model(inputs); loss = compute_loss(outputs, targets, model); loss.backward();
optimizer.step(); optimizer.zero_grad()`` (train.py:61-72, driven by the external ``Trainer``).  ``TrainStep`` is that
sequence for the new path.  One DeepLabV3+ step is ~600 kernel launches of 5-300 us each, so the Python/ctypes enqueue
cost is of the same order as the GPU time; with ``graph=True`` the whole step (forward, loss, backward, bucketed
all-reduce, optimizer update, weight re-packing) is captured once into a CUDA graph on static input buffers and
replayed with one launch per step — streams and graphs, no tracing compiler.

Every kernel in the graph is the same libb200seg launch the eager path makes (the capture simply records them);
buffers come from the graph's private pool of the PyTorch caching allocator.

Gradient accumulation (``--accumulate``, train.py:94, default 2, implemented by the reference's external ``Trainer``):
``TrainStep(accumulate=k)`` runs forward + backward on every call and the gradient exchange + optimizer step on every
k-th; the data-parallel all-reduce is skipped on the other k-1 (``GradBucketer.enabled``), as DDP's ``no_sync`` would.
Each micro-batch's loss is divided by k before backward (the step then descends the mean over the k micro-batches; the
external Trainer's own scaling is unknown — ``average_accumulated=False`` leaves the losses unscaled).  Under a CUDA
graph the k micro-steps and the update are ONE captured graph over k static input slots: calls 1..k-1 only stage their
micro-batch (and return None), call k replays.

What a captured step freezes, and what it does not: shapes, the set of parameters with gradients, momentum / nesterov
on-off and the accumulate count are structural — changing one needs a new TrainStep.  Learning rate, momentum value,
weight decay, Adam's betas / step count are read by the optimizer kernels from a device vector that is refreshed from
``optimizer.param_groups`` before every replay (optim._HyperMixin), so LR schedulers work unchanged.
"""
import torch

from . import _lib
from . import engine as E
from .utils.utils import compute_loss


class TrainStep:
    def __init__(self, model, optimizer, bucketer=None, graph=False, loss_fn=compute_loss, warmup=3, accumulate=1,
                 average_accumulated=True):
        self.model, self.optimizer, self.bucketer, self.loss_fn = model, optimizer, bucketer, loss_fn
        self.use_graph = bool(graph)
        self.accumulate = int(accumulate)
        if self.accumulate < 1:
            raise ValueError('TrainStep(accumulate=%r): need >= 1' % (accumulate,))
        if self.use_graph and warmup < 1:
            # a step captured before any eager one would record the optimizer's first-step path (momentum buffers
            # initialised from the gradient on EVERY replay) and, with freshly packed weights, no weight re-packing
            raise ValueError('TrainStep(graph=True) needs warmup >= 1 eager step before the capture')
        self.warmup = warmup
        self.loss_scale = 1.0 / self.accumulate if average_accumulated else 1.0
        self._graph = None
        self._x = self._t = self._loss = self._side = None
        self._calls = 0
        self._micro = 0                        # micro-batches taken since the last optimizer step
        self._xs = self._ts = None             # graph mode with accumulate > 1: the k static input slots
        self._stage = None
        self._staged = False
        self.launches_per_step = None          # libb200seg launches recorded in the captured step

    # the eager step ------------------------------------------------------------------------------------------
    def _micro_step(self, x, t, first, last):
        """forward + loss + backward of one micro-batch; zeroes the gradients before the first, exchanges and applies
        them after the last."""
        if first:
            self.optimizer.zero_grad(set_to_none=True)
        if self.bucketer is not None:
            self.bucketer.enabled = last           # accumulate locally, all-reduce the sum once (DDP no_sync)
        # the step's logits are consumed by compute_loss alone, which takes loss and gradient from the head's class map:
        # let the head skip writing them (engine.DEFER_LOGITS; compute_loss materialises them if it cannot)
        defer = self.loss_fn is compute_loss and tuple(t.shape[1:]) == tuple(x.shape[2:])
        E.DEFER_LOGITS = defer
        try:
            out = self.model(x)
        finally:
            E.DEFER_LOGITS = False
        loss = self.loss_fn(out, t, self.model)
        (loss * self.loss_scale if self.loss_scale != 1.0 else loss).backward()
        if last:
            if self.bucketer is not None:
                self.bucketer.finish()
            self.optimizer.step()
        return loss.detach()

    def _step(self, x, t):
        first, last = self._micro == 0, self._micro == self.accumulate - 1
        loss = self._micro_step(x, t, first, last)
        self._micro = 0 if last else self._micro + 1
        return loss

    # capture ----------------------------------------------------------------------------------------------------
    def _set_inputs(self, x, t):
        if self._x is None:
            self._x, self._t = torch.empty_like(x), torch.empty_like(t)
            self._side = torch.cuda.Stream()
        if x.shape != self._x.shape or t.shape != self._t.shape:
            raise ValueError('TrainStep(graph=True) is bound to inputs %s / %s, got %s / %s'
                             % (tuple(self._x.shape), tuple(self._t.shape), tuple(x.shape), tuple(t.shape)))
        if x.data_ptr() != self._x.data_ptr():
            self._x.copy_(x, non_blocking=True)
        if t.data_ptr() != self._t.data_ptr():
            self._t.copy_(t, non_blocking=True)

    def _capture(self):
        torch.cuda.synchronize()
        # every packed bf16 weight copy counts as stale from here, so the capture records the re-packing launch whatever
        # the caches held (a graph without it would replay on frozen weights)
        E.WEIGHT_EPOCH += 1
        g = torch.cuda.CUDAGraph()
        l0 = _lib.launches()
        k = self.accumulate
        with torch.cuda.graph(g):
            if k == 1:
                self._loss = self._micro_step(self._x, self._t, True, True)
            else:
                losses = [self._micro_step(self._xs[i], self._ts[i], i == 0, i == k - 1) for i in range(k)]
                self._loss = torch.stack(losses).mean()
        self.launches_per_step = _lib.launches() - l0
        self._graph = g
        self._fresh_capture = True

    def __call__(self, x, t):
        """x: N x 3 x H x W images, t: N x H x W int64 masks, both on the model's device.  Returns the (detached)
        0-dim loss tensor; in graph mode it is a static buffer overwritten by the next call.

        Graph mode: the first ``warmup`` calls run eagerly on a side stream (allocator, lazy state and launch-plan
        memos settle; they are real optimisation steps on the data passed in), the next call captures the step, and
        that and every later call replays it."""
        if not self.use_graph:
            return self._step(x, t)
        if self.accumulate > 1 and (self._graph is not None or self._calls >= self.warmup * self.accumulate):
            return self._graph_accumulate(x, t)
        self._set_inputs(x, t)
        if self._graph is None and self._calls < self.warmup * self.accumulate:
            self._calls += 1
            self._side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._side):
                loss = self._step(self._x, self._t)
            torch.cuda.current_stream().wait_stream(self._side)
            return loss
        if self._graph is None:
            self._capture()
        return self._replay()

    def _replay(self):
        # the optimizer kernels read their hyper-parameters from a device vector whose pinned host mirror the graph copies
        # in on every replay: bring the mirror up to date with param_groups (LR schedule, Adam step count).  The capture
        # call's own step() has just done that.
        if getattr(self, '_fresh_capture', False):
            self._fresh_capture = False
        elif hasattr(self.optimizer, 'refresh_hyper'):
            self.optimizer.refresh_hyper()
        self._graph.replay()
        # a replay updates the parameters without bumping their Python-side version counters: tell the packed-weight
        # caches that whatever they hold is stale for any later eager call (eval after training)
        E.WEIGHT_EPOCH += 1
        return self._loss

    def _graph_accumulate(self, x, t):
        """accumulate > 1 under a graph: stage micro-batch i into its static slot; the k-th call replays the one graph
        that holds all k forward/backward passes and the update.  Returns None on staging calls, the mean loss of the k
        micro-batches on the k-th."""
        k = self.accumulate
        if self._xs is None:
            self._xs = [torch.empty_like(x) for _ in range(k)]
            self._ts = [torch.empty_like(t) for _ in range(k)]
        if x.shape != self._xs[0].shape or t.shape != self._ts[0].shape:
            raise ValueError('TrainStep(graph=True) is bound to inputs %s / %s, got %s / %s'
                             % (tuple(self._xs[0].shape), tuple(self._ts[0].shape), tuple(x.shape), tuple(t.shape)))
        self._xs[self._micro].copy_(x, non_blocking=True)
        self._ts[self._micro].copy_(t, non_blocking=True)
        if self._micro < k - 1:
            self._micro += 1
            return None
        self._micro = 0
        if self._graph is None:
            self._capture()
        return self._replay()

    # host-side input pipeline -------------------------------------------------------------------------------
    def prefetch(self, x_host, t_host):
        """Starts the host -> device copy of a batch (pinned host tensors) on a copy stream; it overlaps whatever the
        compute stream is doing.  ``step_prefetched()`` consumes it.  This is the DataLoader(pin_memory=True) +
        ``.to(device, non_blocking=True)`` hand-off of train.py:28-31 made explicit and double-buffered."""
        dev = next(self.model.parameters()).device
        if self._stage is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._stage = (torch.empty(x_host.shape, dtype=x_host.dtype, device=dev),
                           torch.empty(t_host.shape, dtype=t_host.dtype, device=dev))
            self._stage_free = torch.cuda.Event()
            self._copy_done = torch.cuda.Event()
            self._stage_free.record(torch.cuda.current_stream())
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(self._stage_free)      # the previous batch has left the staging buffers
            self._stage[0].copy_(x_host, non_blocking=True)
            self._stage[1].copy_(t_host, non_blocking=True)
            self._copy_done.record(self._copy_stream)
        self._staged = True

    def step_prefetched(self, next_batch=None):
        """One step on the batch staged by ``prefetch``; ``next_batch=(x_host, t_host)`` starts the copy of the following
        batch before this step's kernels are enqueued, so the copy runs under the step."""
        if not getattr(self, '_staged', False):
            raise RuntimeError('TrainStep.step_prefetched(): no batch staged — call prefetch(x_host, t_host) first')
        cur = torch.cuda.current_stream()
        cur.wait_event(self._copy_done)
        if self._x is None:
            self._x, self._t = torch.empty_like(self._stage[0]), torch.empty_like(self._stage[1])
            self._side = torch.cuda.Stream()
        self._x.copy_(self._stage[0], non_blocking=True)          # device -> device, ~30 us: frees the staging pair
        self._t.copy_(self._stage[1], non_blocking=True)
        self._stage_free.record(cur)
        self._staged = False
        if next_batch is not None:
            self.prefetch(*next_batch)
        return self(self._x, self._t)

    @property
    def static_inputs(self):
        return self._x, self._t
