"""Data-parallel gradient exchange: bucketed all-reduce overlapped with backward.

The reference trains with one process per GPU (train.py:108-119) and leaves the gradient all-reduce to
the external ``Trainer``'s DDP wrap.  This is the equivalent for the new path, without wrapping the
model: parameters are grouped into ~25 MB buckets in reverse registration order (the order backward
produces them); a post-accumulate-grad hook copies each gradient into its slot of the bucket's flat
buffer, re-points ``param.grad`` at that slot, and — when the bucket is full — launches one
asynchronous ``all_reduce`` (NCCL over NVLink/NVSwitch on GPUs, gloo in the CPU tests) while backward
keeps running.  ``finish()`` waits for the outstanding buckets and averages.

Every rank must produce gradients for the SAME set of parameters in every step (as with DDP without
``find_unused_parameters``): ``finish()`` launches the partially filled buckets, and ranks that disagree on which
buckets those are would wait on each other forever.  ``sync_initial_state()`` is DDP's construction-time broadcast.

There is no collective on the data path itself: images are independent through every kernel
(train-mode BatchNorm statistics stay per-GPU, as in the reference, which shows no SyncBN).
"""
import os

import torch
import torch.distributed as dist

# diagnosis only (bench A/B): bucket the gradients but skip the collective, to separate the bucketing's own cost from the
# all-reduce's contention with backward — the ranks then train independently
_SKIP_ALLREDUCE = os.environ.get('B200SEG_SKIP_ALLREDUCE', '0') == '1'


class GradBucketer:
    def __init__(self, params, bucket_mb=25.0, first_bucket_mb=1.0, process_group=None, average=True, last_bucket_mb=4.0):
        self.params = [p for p in params if p.requires_grad]
        self.group = process_group
        self.average = average
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self._buckets = []          # (flat buffer, [(param, offset, numel)])
        self._slot = {}             # id(param) -> (bucket index, offset, numel)
        self._pending = []
        self._ready = []
        self._handles = []
        self._build(bucket_mb, first_bucket_mb, last_bucket_mb)
        self._hooks = [p.register_post_accumulate_grad_hook(self._on_grad) for p in self.params]
        self.enabled = True         # set False on non-final micro-steps of gradient accumulation
        self._early = set()         # parameters whose gradient reached its slot before autograd delivered it

    def sync_initial_state(self, module=None, src=0):
        """Broadcast every parameter (and, given the module, every buffer: BatchNorm running statistics) from rank ``src``
        — what DistributedDataParallel does at construction.  Without it each rank trains from its own random
        initialisation unless the caller seeded them identically, and rank 0's checkpoint is not the collectively
        trained model.  Tensors travel flattened per dtype (a handful of collectives).  Returns self."""
        if self.world <= 1:
            return self
        tensors = list(self.params)
        if module is not None:
            seen = {id(p) for p in tensors}
            tensors += [p for p in module.parameters() if id(p) not in seen]
            tensors += [b for b in module.buffers()]
        by_dtype = {}
        for t in tensors:
            by_dtype.setdefault((t.dtype, t.device), []).append(t)
        with torch.no_grad():
            for group in by_dtype.values():
                flat = torch.cat([t.detach().reshape(-1) for t in group])
                dist.broadcast(flat, src=src, group=self.group)
                off = 0
                for t in group:
                    t.copy_(flat[off:off + t.numel()].view_as(t))
                    off += t.numel()
        return self

    def attach(self):
        """Become the engine's early-gradient sink (engine.deliver_grads): hand-sequenced nodes such as the ResNet-50
        backbone then feed their buckets stage by stage instead of all at once when their backward ends."""
        from . import engine as E
        E.GRAD_SINK = self._deliver
        return self

    def detach(self):
        from . import engine as E
        if E.GRAD_SINK == self._deliver:
            E.GRAD_SINK = None

    def _deliver(self, grads, params):
        if not self.enabled:
            return
        for p in params:
            slot = self._slot.get(id(p))
            g = grads[p]
            # only a fresh gradient can bypass AccumulateGrad's add; anything else takes the ordinary hook path
            if (slot is None or p.grad is not None or id(p) in self._early or g.dtype != p.dtype
                    or g.numel() != p.numel()):
                continue
            bi, off, n = slot
            view = self._buckets[bi][0][off:off + n].view_as(p)
            if g.is_cuda and g.dtype == torch.float32:
                self._moves[bi].append((view, g.contiguous()))
            else:
                view.copy_(g.reshape(p.shape))
            # the node returns None for this parameter: .grad is set here, AccumulateGrad and its hook never run
            p.grad = view
            grads[p] = None
            self._early.add(id(p))
            self._ready[bi] += 1
            if self._ready[bi] == self._pending[bi]:
                self._launch(bi)

    def _build(self, bucket_mb, first_bucket_mb, last_bucket_mb=0.0):
        cap = int(first_bucket_mb * (1 << 20))
        cur, cur_bytes = [], 0
        groups = []
        for p in reversed(self.params):
            nbytes = p.numel() * p.element_size()
            if cur and (cur_bytes + nbytes > cap or cur[0].dtype != p.dtype or cur[0].device != p.device):
                groups.append(cur)
                cur, cur_bytes = [], 0
                cap = int(bucket_mb * (1 << 20))
            cur.append(p)
            cur_bytes += nbytes
        if cur:
            groups.append(cur)
        # The bucket that completes LAST (the first-registered parameters: the stem and the first backbone stage) starts
        # its all-reduce when backward is over, i.e. fully exposed: keep it small, so what is exposed is a few MB and the
        # rest of that tail travels with the previous bucket, under the last stage's backward.
        if last_bucket_mb and groups:
            tail, cap, nbytes = groups[-1], int(last_bucket_mb * (1 << 20)), 0
            cut = len(tail)
            while cut > 1 and nbytes + tail[cut - 1].numel() * tail[cut - 1].element_size() <= cap:
                cut -= 1
                nbytes += tail[cut].numel() * tail[cut].element_size()
            if 0 < cut < len(tail):
                groups[-1:] = [tail[:cut], tail[cut:]]
        for bi, g in enumerate(groups):
            total = sum(p.numel() for p in g)
            flat = torch.zeros((total,), dtype=g[0].dtype, device=g[0].device)
            slots, off = [], 0
            for p in g:
                slots.append((p, off, p.numel()))
                self._slot[id(p)] = (bi, off, p.numel())
                off += p.numel()
            self._buckets.append((flat, slots))
            self._pending.append(len(slots))
        self._ready = [0] * len(self._buckets)
        self._moves = [[] for _ in self._buckets]

    @property
    def num_buckets(self):
        return len(self._buckets)

    def bucket_sizes(self):
        return [b[0].numel() for b in self._buckets]

    def _on_grad(self, p):
        if not self.enabled:
            return
        bi, off, n = self._slot[id(p)]
        flat = self._buckets[bi][0]
        view = flat[off:off + n].view_as(p)
        if id(p) in self._early:
            # delivered by its node ahead of autograd (which still runs this hook, with an undefined gradient, when
            # the node returns None): the slot is filled and counted, .grad must simply still be that slot
            if p.grad is None or p.grad.data_ptr() != view.data_ptr():
                raise RuntimeError('GradBucketer: a gradient delivered early was replaced before the step finished')
            return
        if p.grad.data_ptr() != view.data_ptr():
            g = p.grad
            if g.is_cuda and g.dtype == torch.float32 and g.is_contiguous():
                self._moves[bi].append((view, g))       # moved by ONE multi-tensor launch when the bucket is complete
            else:
                view.copy_(g)
            p.grad = view
        self._ready[bi] += 1
        if self._ready[bi] == self._pending[bi]:
            self._launch(bi)

    def _flush_moves(self, bi):
        moves = self._moves[bi]
        if not moves:
            return
        from . import _lib, engine as _engine
        arr = (_lib.SgdDesc * len(moves))()
        for i, (dst, src) in enumerate(moves):
            arr[i] = _lib.SgdDesc(dst.data_ptr(), src.data_ptr(), None, src.numel())
        _lib.check(_lib.load().b200seg_copy_multi(arr, len(moves), _engine._stream()),
                   'copy_multi')
        _lib.count_launch((len(moves) + 95) // 96)
        self._moves[bi] = []

    def _launch(self, bi):
        self._flush_moves(bi)
        if self.world > 1 and not _SKIP_ALLREDUCE:
            self._handles.append((bi, dist.all_reduce(self._buckets[bi][0], op=dist.ReduceOp.SUM, group=self.group,
                                                      async_op=True)))

    def finish(self):
        """Call after backward (before the optimizer step): flushes partially filled buckets (parameters
        that received no gradient contribute zeros), waits for the collectives and averages."""
        launched = {bi for bi, _ in self._handles}
        for bi in range(len(self._buckets)):
            if self._ready[bi] != self._pending[bi] and self._ready[bi] >= 0 and self.enabled:
                if bi not in launched and self._ready[bi] > 0:
                    self._launch(bi)
        for bi in range(len(self._buckets)):
            self._flush_moves(bi)
        prof = getattr(self, 'profile', None)
        if prof is not None:
            # exposed collective time: how long the compute stream stalls here for all-reduces still in flight
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        for bi, h in self._handles:
            h.wait()
            if self.average and self.world > 1:
                self._buckets[bi][0].div_(self.world)
        if prof is not None:
            e1.record()
            prof.append((e0, e1))
        self._handles = []
        self._ready = [0] * len(self._buckets)
        self._early.clear()

    def zero_grad(self):
        """Zero the flat buffers in place (gradients stay views of them)."""
        for flat, _ in self._buckets:
            flat.zero_()

    def remove(self):
        self.detach()
        for h in self._hooks:
            h.remove()
        self._hooks = []
