"""One optimisation step of the hot path, optionally replayed from a CUDA graph.

The reference's step is ``outputs = model(inputs); loss = compute_loss(outputs, targets, model); loss.backward();
optimizer.step(); optimizer.zero_grad()`` (train.py:61-72, driven by the external ``Trainer``).  ``TrainStep`` is that
sequence for the new path.  One DeepLabV3+ step is ~600 kernel launches of 5-300 us each, so the Python/ctypes enqueue
cost is of the same order as the GPU time; with ``graph=True`` the whole step (forward, loss, backward, bucketed
all-reduce, optimizer update, weight re-packing) is captured once into a CUDA graph on static input buffers and
replayed with one launch per step — streams and graphs, no tracing compiler.

Every kernel in the graph is the same libb200seg launch the eager path makes (the capture simply records them);
buffers come from the graph's private pool of the PyTorch caching allocator.
"""
import torch

from . import _lib
from . import engine as E
from .utils.utils import compute_loss


class TrainStep:
    def __init__(self, model, optimizer, bucketer=None, graph=False, loss_fn=compute_loss, warmup=3):
        self.model, self.optimizer, self.bucketer, self.loss_fn = model, optimizer, bucketer, loss_fn
        self.use_graph = bool(graph)
        self.warmup = warmup
        self._graph = None
        self._x = self._t = self._loss = self._side = None
        self._calls = 0
        self._stage = None
        self._staged = False
        self.launches_per_step = None          # libb200seg launches recorded in the captured step

    # the eager step ------------------------------------------------------------------------------------------
    def _step(self, x, t):
        self.optimizer.zero_grad(set_to_none=True)
        out = self.model(x)
        loss = self.loss_fn(out, t, self.model)
        loss.backward()
        if self.bucketer is not None:
            self.bucketer.finish()
        self.optimizer.step()
        return loss.detach()

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
        g = torch.cuda.CUDAGraph()
        l0 = _lib.launches()
        with torch.cuda.graph(g):
            self._loss = self._step(self._x, self._t)
        self.launches_per_step = _lib.launches() - l0
        self._graph = g

    def __call__(self, x, t):
        """x: N x 3 x H x W images, t: N x H x W int64 masks, both on the model's device.  Returns the (detached)
        0-dim loss tensor; in graph mode it is a static buffer overwritten by the next call.

        Graph mode: the first ``warmup`` calls run eagerly on a side stream (allocator, lazy state and launch-plan
        memos settle; they are real optimisation steps on the data passed in), the next call captures the step, and
        that and every later call replays it."""
        if not self.use_graph:
            return self._step(x, t)
        self._set_inputs(x, t)
        if self._graph is None and self._calls < self.warmup:
            self._calls += 1
            self._side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._side):
                loss = self._step(self._x, self._t)
            torch.cuda.current_stream().wait_stream(self._side)
            return loss
        if self._graph is None:
            self._capture()
        self._graph.replay()
        # a replay updates the parameters without bumping their Python-side version counters: tell the packed-weight
        # caches that whatever they hold is stale for any later eager call (eval after training)
        E.WEIGHT_EPOCH += 1
        return self._loss

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
