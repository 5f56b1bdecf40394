#!/usr/bin/env python
"""bench.py — the BASELINE.json configurations of the pytorch_segmentation hot path on N B200s.

  python bench.py [--config train|infer|aspp|hrnet|unet|unet_cpu] [--gpus N] [--steps K] [--warmup W]
                  [--impl b200|reference] [--batch B]

  --config train (default; BASELINE configs[3], the headline): DeepLabV3+ 21-class 512x512 bf16 training.  One step =
      forward + compute_loss + backward (+ bucketed gradient all-reduce for N > 1) + SGD update on one synthetic batch
      of B images per GPU (random-init weights; per-rank seeds).  metric: train images/sec.
  --config infer  (configs[1]): DeepLabV3+ 21-class 512x512 bf16 inference batch 16 incl. the mIoU confusion histogram
      of test.py:28-46 (fused evaluation from the class map).
  --config aspp   (configs[2]): the ASPP block isolated (2048-ch input, rates 6/12/18), forward + backward; the
      fprop / dgrad / wgrad launches are listed separately in ``kernels``.
  --config hrnet  (configs[4]): HRNet 19-class 1024x512 bf16 training, per-GPU batch 8.
  --config unet   (configs[0] on the GPU): UNet 2-class 320x320 batch 4 training step.
  --config unet_cpu (configs[0] as stated: CPU): the oracle's UNet 2-class 320x320 batch 4 fwd+bwd on the host cores.

Launched by the driver as one process per GPU (torch.distributed.run) for N > 1.  Rank 0 prints ONE JSON line.

  value     whole-job images/sec with the step's inputs already resident in HBM
  e2e       the same through the public API with HOST buffers: pinned-host -> device copies of the uint8 images and
            masks (utils.post_fetch_fn drop-in: 1-byte pixels and labels) and a device -> host read of the loss inside the
            timed region, every step
  roofline  the dominant kernel (tcgen05 implicit-GEMM conv, all fprop/dgrad launches of the step): nominal dense
            FLOPs (2*N*P*Q*K*C*R*S, padding taps counted — SURVEY.md §8d) / CUDA-event time, vs the measured bf16
            peak in MEASURED_PEAKS.json; ``flops_executed`` / ``frac_executed`` leave out the all-padding (tap, tile)
            pairs the kernel skips; ``roofline_bn`` is the same for the HBM-bound BatchNorm family
  cpu_baseline / --impl reference
            the CPU oracle (oracle/restatement.py: the reference's own PyTorch calls, fp32) on the host cores on a
            bounded sample of the same workload
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

NC, H, W = 21, 512, 512
METRIC = 'train images/sec @512x512 (DeepLabV3+ 21-class, bf16)'
CPU_SAMPLE_BATCH = 2


def peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        p = json.load(open(path))
        return {'bf16_tflops': p.get('bf16_tflops'), 'bf16_tflops_sustained': p.get('bf16_tflops_sustained'),
                'hbm_gbs': p.get('hbm_gbs'), 'source': 'measured (MEASURED_PEAKS.json)'}
    return {'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0, 'hbm_gbs': 6650.0,
            'source': 'fallback (B200_PROFILING.md)'}


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.check_output(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                               '--format=csv,noheader,nounits'], timeout=5).decode().strip()
                self.rows.append([v.strip() for v in out.split(',')])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace('.', '').isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace('.', '').isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower().startswith('active')
                                                          for r in self.rows)]
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': reasons, 'samples': len(self.rows)}


# -----------------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle on the host cores
# -----------------------------------------------------------------------------------------------------------
def cpu_reference_run(steps, warmup, batch=CPU_SAMPLE_BATCH):
    """DeepLabV3+ 21-class 512x512 train steps (fwd + compute_loss + bwd + SGD) through the oracle, fp32, all
    host threads.  Returns (images/sec, seconds per step, threads)."""
    from oracle import restatement as R
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    model = R.OracleModel('deeplabv3plus', NC, seed=0).train()
    params = [p for p in model.backbone.parameters()] + [v for v in model.sd.values() if v.requires_grad]
    opt = torch.optim.SGD(params, lr=1e-3, momentum=0.9)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(batch, 3, H, W, generator=g)
    t = torch.randint(0, NC, (batch, H, W), generator=g)

    def step():
        opt.zero_grad(set_to_none=True)
        loss = R.compute_loss(model(x), t)
        loss.backward()
        opt.step()
        return float(loss)

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / max(steps, 1)
    return batch / dt, dt, threads


def cpu_model():
    try:
        for line in open('/proc/cpuinfo'):
            if line.startswith('model name'):
                return line.split(':', 1)[1].strip()
    except OSError:
        pass
    return 'unknown'


def cpu_unet_run(steps, warmup=1):
    """BASELINE.md §4 "config 1 exactly": UNet(2).train(), randn(4,3,320,320) (seed 0), randint(0,2,(4,320,320)) (seed 1),
    random init (seed 2), fwd + compute_loss + bwd, fp32, all host threads.  Returns (best s/step, median s/step, threads)."""
    from oracle import restatement as R
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    model = R.OracleModel('unet', 2, seed=2).train()
    x = torch.randn(4, 3, 320, 320, generator=torch.Generator().manual_seed(0))
    t = torch.randint(0, 2, (4, 320, 320), generator=torch.Generator().manual_seed(1))
    times = []
    for i in range(warmup + steps):
        model.zero_grad()
        t0 = time.perf_counter()
        loss = R.compute_loss(model(x), t)
        loss.backward()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    times.sort()
    return times[0], times[len(times) // 2], threads


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    if args.config in ('unet', 'unet_cpu'):
        steps = max(5, min(args.steps, 20))
        best, med, threads = cpu_unet_run(steps, 1)
        ips = 4 / med
        line = {'impl': 'reference', 'metric': 'UNet 2-class 320x320 batch 4 fwd+bwd images/sec', 'value': ips,
                'unit': 'images/sec', 'n_gpus': args.gpus, 'steps': steps, 'warmup': 1, 'ms_per_step': med * 1e3,
                'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
                'config': {'workload': 'UNet 2-class 320x320 batch 4 fwd+compute_loss+bwd (BASELINE configs[0]); CPU oracle '
                                       'port of the reference model code, fp32', 'per_step_batch': 4},
                'cpu_baseline': {'value': ips, 'unit': 'images/sec', 'cores': threads, 'kind': 'port', 'cpu': cpu_model(),
                                 'best_s_per_step': best, 'median_s_per_step': med,
                                 'sample': 'config 1 exactly (BASELINE.md §4), 1 warm-up + %d timed steps; value = median' % steps},
                'e2e': {'value': ips, 'unit': 'images/sec', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
                'gpu_launches': 0}
        print(json.dumps(line), flush=True)
        return
    if args.config in ('infer', 'aspp', 'hrnet'):
        ips, dt, threads, batch, what = cpu_other_run(args.config, max(2, min(args.steps, 10)))
        metric = OTHER[args.config][6]
        line = {'impl': 'reference', 'metric': metric, 'value': ips, 'unit': 'images/sec', 'n_gpus': args.gpus,
                'steps': max(2, min(args.steps, 10)), 'warmup': 1, 'ms_per_step': dt * 1e3, 'higher_is_better': True,
                'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
                'config': {'workload': what, 'per_step_batch': batch},
                'cpu_baseline': {'value': ips, 'unit': 'images/sec', 'cores': threads, 'kind': 'port', 'cpu': cpu_model(),
                                 'sample': 'batch %d per step (bounded sample), 1 warm-up' % batch},
                'e2e': {'value': ips, 'unit': 'images/sec', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
                'gpu_launches': 0}
        print(json.dumps(line), flush=True)
        return
    steps, warmup = max(1, min(args.steps, 20)), max(1, min(args.warmup, 2))
    ips, dt, threads = cpu_reference_run(steps, warmup)
    sample = 'batch %d per step, %d timed steps (bounded sample of the per-GPU batch-%d step)' % (
        CPU_SAMPLE_BATCH, steps, args.batch)
    line = {'impl': 'reference', 'metric': METRIC, 'value': ips, 'unit': 'images/sec', 'n_gpus': args.gpus,
            'steps': steps, 'warmup': warmup, 'ms_per_step': dt * 1e3, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': {'workload': 'DeepLabV3+ 21-class 512x512 training (configs[3]); CPU oracle port of the '
                                   'reference model code, fp32', 'per_step_batch': CPU_SAMPLE_BATCH},
            'cpu_baseline': {'value': ips, 'unit': 'images/sec', 'cores': threads, 'kind': 'port', 'sample': sample,
                             'cpu': cpu_model()},
            'e2e': {'value': ips, 'unit': 'images/sec', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


def cpu_other_run(config, steps):
    """The oracle's CPU leg of the configurations other than training / UNet: a bounded batch of the same workload.
    Returns (images/sec, seconds per step, threads, batch, workload text)."""
    from oracle import restatement as R
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(0)
    if config == 'infer':
        batch = 2
        model = R.OracleModel('deeplabv3plus', NC, seed=0).eval()
        x, t = torch.randn(batch, 3, H, W, generator=g), torch.randint(0, NC, (batch, H, W), generator=g)

        def step():
            with torch.no_grad():
                y = model(x)
                return R.confusion_matrix(y.max(1)[1], t, NC)
        what = 'DeepLabV3+ 21-class 512x512 inference + confusion histogram (configs[1]); CPU oracle, fp32'
    elif config == 'aspp':
        batch = 2
        sd = {}
        full = R.OracleModel('deeplabv3plus', NC, seed=0)
        sd = {k: v for k, v in full.sd.items() if k.startswith('aspp.')}
        x = torch.randn(batch, 2048, 32, 32, generator=g).requires_grad_(True)
        gy = torch.randn(batch, 256, 32, 32, generator=g)

        def step():
            x.grad = None
            for v in sd.values():
                v.grad = None
            y = R.aspp(sd, 'aspp', x, R.DEEPLAB_RATES, True)
            y.backward(gy)
            return y
        what = 'ASPP(2048, 256, [6, 12, 18]) forward + backward (configs[2]); CPU oracle, fp32'
    else:
        from oracle import hrnet_restatement as HR
        batch = 1
        sd = HR.make_state(19, seed=0)
        for k, v in sd.items():
            if v.is_floating_point() and 'running_' not in k:
                v.requires_grad_(True)
        x, t = torch.randn(batch, 3, 512, 1024, generator=g), torch.randint(0, 19, (batch, 512, 1024), generator=g)

        def step():
            for v in sd.values():
                v.grad = None
            loss = R.compute_loss(HR.hrnet(sd, x, training=True), t)
            loss.backward()
            return loss
        what = 'HRNet 19-class 1024x512 training step fwd+loss+bwd (configs[4]); CPU oracle, fp32'
    step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return batch / dt, dt, threads, batch, what


# -----------------------------------------------------------------------------------------------------------
# the B200 arm
# -----------------------------------------------------------------------------------------------------------
def run_b200(args):
    import pytorch_segmentation_b200 as P
    from pytorch_segmentation_b200 import _lib, engine as E
    from pytorch_segmentation_b200.parallel import GradBucketer

    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py: no CUDA device — the B200 arm has no CPU fallback (use --impl reference)')
    if world > 1:
        # the gradient all-reduce runs under backward: give NCCL a few SMs of its own instead of letting it push the
        # one-CTA-per-SM conv kernels into a second wave (csrc/common.cu gemm_sm_count)
        if args.sm_reserve:
            os.environ.setdefault('B200SEG_SM_RESERVE', str(args.sm_reserve))
        if args.nccl_ctas:
            os.environ.setdefault('NCCL_MAX_CTAS', str(args.nccl_ctas))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
    _lib.load()

    B = args.batch
    torch.manual_seed(1234)                       # identical initial weights on every rank
    model = P.DeepLabV3Plus(NC).to(dev).train()
    # one multi-tensor launch (b200seg_sgd_multi); the 1/world of the gradient average is folded into the update
    opt = P.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9, grad_scale=1.0 / world)
    bucketer = (GradBucketer(model.parameters(), bucket_mb=args.bucket_mb, average=False, last_bucket_mb=args.last_bucket_mb).attach().sync_initial_state(model)
                if world > 1 else None)

    # inputs as the reference's dataset yields them (utils/datasets.py:161-194): uint8 images, uint8 masks.  The
    # normalisation of post_fetch_fn (utils/datasets.py:199-205) runs on the device inside the stem's space-to-depth
    # pass, the loss reads the 1-byte masks directly (utils.post_fetch_fn drop-in)
    g = torch.Generator(device='cpu').manual_seed(100 + rank)
    x_host = torch.randint(0, 256, (B, 3, H, W), generator=g, dtype=torch.uint8).pin_memory()
    t_host = torch.randint(0, NC, (B, H, W), generator=g, dtype=torch.uint8).pin_memory()
    x_dev, t_dev = x_host.to(dev), t_host.to(dev)
    loss_host = torch.zeros((), dtype=torch.float32).pin_memory()

    use_graph = args.graph in ('on', 'auto')
    eager = P.TrainStep(model, opt, bucketer, graph=False)
    step = P.TrainStep(model, opt, bucketer, graph=True) if use_graph else eager

    def step_e2e():
        # the public call with HOST buffers, every step: pinned host -> device copies of this step's images and masks
        # (issued one step ahead on a copy stream, so they run under the previous step's kernels), the step, and the
        # loss read back to the host
        loss = step.step_prefetched(next_batch=(x_host, t_host))
        loss_host.copy_(loss, non_blocking=True)
        torch.cuda.current_stream().synchronize()      # the user reads the loss every step (Trainer prints it)
        return loss_host

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / 1e3

    for _ in range(args.warmup + (4 if use_graph else 0)):      # graph mode: 3 eager calls + the capture come first
        step(x_dev, t_dev)
    # host cost of enqueueing one eager step (Python + ctypes + allocator), for the record: it is what the graph removes
    # (the first eager call after graph replays re-packs every weight and grows the default memory pool by the ~9 GB of
    # activations the graph keeps in its private pool — ~200 ms of cudaMalloc — so the second call is the one timed)
    eager(x_dev, t_dev)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    l0 = _lib.launches()
    eager(x_dev, t_dev)
    host_enqueue_ms = (time.perf_counter() - t0) * 1e3
    eager_launches = _lib.launches() - l0
    torch.cuda.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = _lib.launches()
    secs = timed(lambda: step(x_dev, t_dev), args.steps)
    launches = (step.launches_per_step * args.steps) if use_graph else (_lib.launches() - l0)
    clocks = sampler.stop() if sampler else None
    step.prefetch(x_host, t_host)                      # the first batch's copy; every later one overlaps a step
    for _ in range(min(args.warmup, 2)):
        step_e2e()
    secs_e2e = timed(step_e2e, args.steps)
    torch.cuda.synchronize()
    # the same step called eagerly, as the reference's external Trainer would (no graph): host-bound by the Python +
    # ctypes enqueue cost the graph removes
    eager_ips = None
    if use_graph:
        eager(x_dev, t_dev)
        k_eager = max(3, min(args.steps, 10))
        eager(x_dev, t_dev)                             # the pool has regrown by now: a steady-state eager step
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        eager(x_dev, t_dev)
        host_enqueue_ms = (time.perf_counter() - t0) * 1e3
        eager_ips = world * B * k_eager / timed(lambda: eager(x_dev, t_dev), k_eager)

    # per-launch CUDA-event timing of the tcgen05 conv kernels over a few more steps (the events sit on the
    # launching stream; kept out of the throughput loop above so the headline has no instrumentation in it)
    # per-launch timing runs with the weight-gradient lane off: concurrent kernels would share SMs with the launch
    # being timed and understate it
    lane_was, E._LANE_ON = E._LANE_ON, False
    E.PROFILE = []
    E.OP_PROFILE = []
    if bucketer is not None:
        bucketer.profile = []
    prof_steps = min(args.steps, 3)
    for _ in range(prof_steps):
        # the host needs about as long to enqueue an eager step as the GPU to run it; a ~40 ms spin kernel ahead of each
        # profiled step lets the host run ahead, so the event pairs bracket back-to-back kernels, not launch latency
        torch.cuda._sleep(80_000_000)
        eager(x_dev, t_dev)
    torch.cuda.synchronize()
    rec, E.PROFILE = E.PROFILE, None
    op_rec, E.OP_PROFILE = E.OP_PROFILE, None
    E._LANE_ON = lane_was
    # data parallel: every rank must hold the same parameters after the same number of averaged updates
    sync_diff = exposed_ms = None
    if bucketer is not None:
        ev, bucketer.profile = bucketer.profile, None
        exposed_ms = sum(a.elapsed_time(b) for a, b in ev) / max(len(ev), 1)
    if world > 1:
        chk = torch.stack([p.detach().double().sum() for p in model.parameters()])
        hi, lo = chk.clone(), chk.clone()
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        sync_diff = float((hi - lo).abs().max())
    agg = {}
    for tag, flops, e0, e1, info in rec:
        a = agg.setdefault(tag, [0.0, 0.0, 0, 0.0, 0.0])
        a[0] += flops
        a[1] += e0.elapsed_time(e1) * 1e-3
        a[2] += 1
        a[3] += flops * info['run_share']           # FLOPs of the (tap, tile) pairs actually run
        a[4] += info['bytes']
    pk = peaks()
    igemm = [(t, v) for t, v in agg.items() if t.startswith('conv_fprop') or t.startswith('conv_dgrad')]
    ig_flops, ig_secs, ig_n, ig_exec, ig_bytes = (sum(v[i] for _, v in igemm) for i in range(5))
    peak_tf = pk['bf16_tflops_sustained'] or pk['bf16_tflops']
    achieved = ig_flops / ig_secs / 1e12 if ig_secs > 0 else 0.0
    detail = sorted(({'launch': t, 'tflops': round(v[0] / v[1] / 1e12, 1), 'us': round(v[1] / v[2] * 1e6, 1),
                      'frac': round(v[0] / v[1] / 1e12 / peak_tf, 3),
                      'frac_executed': round(v[3] / v[1] / 1e12 / peak_tf, 3),
                      'skipped_tap_share': round(1.0 - v[3] / v[0], 3) if v[0] else 0.0}
                     for t, v in agg.items() if v[1] > 0), key=lambda d: -d['us'])
    # the BatchNorm family (apply / backward / residual-gradient join): HBM-bound, algorithmic bytes per SURVEY §8d
    bn_bytes = bn_secs = 0.0
    bn_n = 0
    for tag, e0, e1, nbytes in op_rec or []:
        if nbytes > 0:
            bn_bytes += nbytes
            bn_secs += e0.elapsed_time(e1) * 1e-3
            bn_n += 1
    traffic = traffic_src = None
    tpath = os.path.join(ROOT, 'profiles', 'igemm_traffic.json')
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        traffic, traffic_src = tj.get('dram_bytes_per_launch'), tj.get('source')

    if rank == 0:
        ips = world * B * args.steps / secs
        ips_e2e = world * B * args.steps / secs_e2e
        line = {
            'metric': METRIC, 'value': ips, 'unit': 'images/sec', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': secs / args.steps * 1e3, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic',
            'config': {'workload': 'DeepLabV3+ 21-class 512x512 bf16 training, fwd+loss+bwd+SGD (BASELINE configs[3]); uint8 images '
                                   'and masks as the dataset yields them, post_fetch_fn normalisation fused into the stem pass',
                       'per_gpu_batch': B, 'global_batch': B * world, 'parallelism': 'dp%d' % world,
                       'cuda_graph': bool(use_graph), 'host_enqueue_ms_per_eager_step': round(host_enqueue_ms, 2),
                       'eager_images_per_sec': round(eager_ips, 1) if eager_ips else None,
                       'libb200seg_launches_per_step': eager_launches,
                       'l2': 'no explicit flush: one step streams several GB of activations (>> 126 MB L2)',
                       'backbone': 'ResNet-50 OS16 on libb200seg (tcgen05 convs incl. stride 2, fused BN/ReLU/residual, '
                                   'max pool); 7x7/2 stem = 4 row taps over the space-to-depth image (tcgen05); no cuDNN on the path'},
            'e2e': {'value': ips_e2e, 'unit': 'images/sec', 'h2d_bytes_per_step': x_host.numel() * x_host.element_size() + t_host.numel() * t_host.element_size(),
                    'd2h_bytes_per_step': 4, 'ms_per_step': secs_e2e / args.steps * 1e3},
            'gpu_launches': launches,
            'clocks': clocks,
            'roofline': {'bound': 'tensor', 'achieved': achieved, 'peak': peak_tf, 'unit': 'TFLOP/s',
                         'frac': achieved / peak_tf if peak_tf else None,
                         'flops_executed_per_step': ig_exec / max(prof_steps, 1),
                         'frac_executed': ig_exec / ig_secs / 1e12 / peak_tf if ig_secs > 0 and peak_tf else None,
                         'skipped_tap_share': 1.0 - ig_exec / ig_flops if ig_flops else None,
                         'traffic': traffic, 'traffic_source': traffic_src,
                         'algorithmic_bytes': ig_bytes / max(ig_n, 1),
                         'kernel': 'igemm_kernel (tcgen05 implicit-GEMM fprop+dgrad, %d launches/step)' % (ig_n // max(prof_steps, 1)),
                         'flops_per_step': ig_flops / max(prof_steps, 1), 'peak_source': pk['source'] + ', sustained',
                         'note': 'frac counts padding taps (SURVEY 8d nominal FLOPs); frac_executed leaves out the all-padding '
                                 '(tap, tile) pairs the kernel skips (ASPP d=12/18 skip 44 % / 61 %), so per-launch frac > 1 is '
                                 'nominal work not done, not time saved by magic; durations are eager per-launch CUDA events '
                                 'taken with the host running ahead of the GPU (a spin kernel precedes each profiled step), traffic / '
                                 'algorithmic_bytes are per launch averaged over the step'},
            'roofline_bn': {'bound': 'hbm', 'achieved': bn_bytes / bn_secs / 1e9 if bn_secs > 0 else None, 'peak': pk['hbm_gbs'],
                            'unit': 'GB/s', 'frac': bn_bytes / bn_secs / 1e9 / pk['hbm_gbs'] if bn_secs > 0 and pk['hbm_gbs'] else None,
                            'kernel': 'bn_apply / bn_bwd_reduce+apply / grad_add_masked (%d calls/step)' % (bn_n // max(prof_steps, 1)),
                            'algorithmic_bytes_per_step': bn_bytes / max(prof_steps, 1),
                            'ms_per_step': bn_secs / max(prof_steps, 1) * 1e3,
                            'bytes_rule': 'apply 2x tensor (+1 with a residual), backward 2x(dy+x)+dx, join dx r/w + operands'},
            'kernels': detail[:12],
        }
        if sync_diff is not None:
            line['param_sync_max_abs_diff'] = sync_diff
            # CUDA events around GradBucketer.finish()'s waits in the eager profiling steps: the part of the gradient
            # all-reduce that backward did not hide (rank 0; no nsys on this image)
            line['allreduce_exposed_ms_per_step'] = exposed_ms
        if world == 1 and not args.no_cpu_baseline:
            ips_cpu, dt, threads = cpu_reference_run(16, 1)      # ~10 s of host work
            ub, um, _ = cpu_unet_run(6, 1)                        # BASELINE.md §4 "config 1 exactly", ~5 s
            line['cpu_baseline'] = {'value': ips_cpu, 'unit': 'images/sec', 'cores': threads, 'kind': 'port', 'cpu': cpu_model(),
                                    'sample': 'oracle DeepLabV3+ 512x512 train step, batch %d, 1 warm-up + 16 timed steps '
                                              '(%.2f s/step)' % (CPU_SAMPLE_BATCH, dt),
                                    'config1_unet2_b4_320': {'best_s_per_step': ub, 'median_s_per_step': um,
                                                             'images_per_sec_best': 4 / ub, 'images_per_sec_median': 4 / um,
                                                             'what': 'UNet(2) batch 4 320x320 fwd+compute_loss+bwd, fp32, '
                                                                     '1 warm-up + 6 timed (BASELINE.md §4)'}}
        if args.detail:
            with open(args.detail, 'w') as f:
                ops, ops_by_shape = {}, {}
                for tag, e0, e1, _nb in op_rec or []:
                    us = e0.elapsed_time(e1) * 1e3 / prof_steps
                    for table, key in ((ops, tag.split(' ')[0]), (ops_by_shape, tag)):
                        o = table.setdefault(key, [0.0, 0])
                        o[0] += us
                        o[1] += 1
                counts = {t: v[2] // prof_steps for t, v in agg.items()}
                for d in detail:
                    d['per_step'] = counts.get(d['launch'])
                conv_us = {k2: sum(v[1] for t, v in agg.items() if t.startswith(k2)) * 1e6 / prof_steps
                           for k2 in ('conv_fprop', 'conv_dgrad', 'conv_wgrad')}
                json.dump({'ms_per_step': line['ms_per_step'], 'conv_us_per_step': conv_us,
                           'ops_us_per_step': {k2: [round(v[0], 1), v[1] // prof_steps]
                                               for k2, v in sorted(ops.items(), key=lambda kv: -kv[1][0])},
                           'ops_by_shape_us_per_step': {k2: [round(v[0], 1), v[1] // prof_steps]
                                                        for k2, v in sorted(ops_by_shape.items(), key=lambda kv: -kv[1][0])},
                           'kernels': detail}, f, indent=1)
        print(json.dumps(line), flush=True)
    if world > 1:
        _teardown(step, eager, bucketer)


def _teardown(*objs):
    """Leave a multi-rank run cleanly: the captured graphs (which hold NCCL kernels) go first, then the communicator.  A
    watchdog ends the process if the teardown has not finished in 30 s (round 1 saw communicator teardown block under a
    live graph): the JSON line is out by then, so the only effect is the exit path."""
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    sys.stderr.flush()
    done = threading.Event()

    def watchdog():
        if not done.wait(30.0):
            os._exit(0)
    threading.Thread(target=watchdog, daemon=True).start()
    for o in objs:
        if o is None:
            continue
        for name in ('_graph', '_loss'):
            if hasattr(o, name):
                setattr(o, name, None)
        if hasattr(o, 'remove'):
            o.remove()
    import gc
    gc.collect()
    torch.cuda.synchronize()
    dist.destroy_process_group()
    done.set()


# -----------------------------------------------------------------------------------------------------------
# the other BASELINE.json configurations (same contract, one JSON line)
# -----------------------------------------------------------------------------------------------------------
OTHER = {
    # name: (model, classes, H, W, per-GPU batch, train?, metric, workload)
    'infer': ('deeplab', 21, 512, 512, 16, False, 'inference images/sec @512x512 (DeepLabV3+ 21-class, bf16)',
              'DeepLabV3+ 21-class 512x512 bf16 inference batch 16 + mIoU confusion histogram (BASELINE configs[1])'),
    'hrnet': ('hrnet', 19, 512, 1024, 8, True, 'train images/sec @1024x512 (HRNet 19-class, bf16)',
              'HRNet 19-class 1024x512 bf16 training, fwd+loss+bwd+SGD (BASELINE configs[4])'),
    'unet': ('unet', 2, 320, 320, 4, True, 'train images/sec @320x320 (UNet 2-class, bf16)',
             'UNet 2-class 320x320 batch 4 fwd+loss+bwd+SGD on the GPU (BASELINE configs[0] shapes; MobileNetV2 encoder on '
             'torchvision/cuDNN — outside SURVEY §8 — head on libb200seg)'),
    'aspp': ('aspp', 21, 32, 32, 16, True, 'ASPP block fwd+bwd images/sec (2048-ch 32x32 input, rates 6/12/18, bf16)',
             'ASPP(2048, 256, [6, 12, 18]) isolated, forward + backward, batch 16 (BASELINE configs[2])'),
}


def run_other(args):
    import pytorch_segmentation_b200 as P
    from pytorch_segmentation_b200 import _lib, engine as E
    from pytorch_segmentation_b200.parallel import GradBucketer
    kind, nc, h, w, batch, train, metric, workload = OTHER[args.config]
    B = args.batch if args.batch_given else batch
    rank, local, world = (int(os.environ.get(k, d)) for k, d in (('RANK', '0'), ('LOCAL_RANK', '0'), ('WORLD_SIZE', '1')))
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py: no CUDA device — the B200 arm has no CPU fallback (use --impl reference)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)
    _lib.load()
    torch.manual_seed(1234)
    g = torch.Generator(device='cpu').manual_seed(100 + rank)
    bucketer = step = None
    if kind == 'aspp':
        model = P.ASPP(2048, 256, [6, 12, 18]).to(dev).train()
        x_host = torch.randn(B, h, w, 2048, generator=g).to(torch.bfloat16).pin_memory()
        gy = torch.randn(B, h, w, 256, generator=g).to(torch.bfloat16).to(dev).permute(0, 3, 1, 2)
        x_dev = x_host.to(dev).permute(0, 3, 1, 2).requires_grad_(True)
        t_host = None

        def fn(x=None):
            xin = x_dev if x is None else x
            xin.grad = None
            model.zero_grad(set_to_none=True)
            E.pack_all(E.packed_entries(model))
            y = model(xin)
            y.backward(gy)
            return y

        def fn_e2e():
            xx = x_host.to(dev, non_blocking=True).permute(0, 3, 1, 2).requires_grad_(True)
            y = fn(xx)
            return y.detach().float().abs().mean()
        h2d = x_host.numel() * 2
    else:
        cls = {'deeplab': P.DeepLabV3Plus, 'hrnet': P.HRNet, 'unet': P.UNet}[kind]
        model = cls(nc).to(dev)
        x_host = torch.randint(0, 256, (B, 3, h, w), generator=g, dtype=torch.uint8).pin_memory()
        t_host = torch.randint(0, nc, (B, h, w), generator=g, dtype=torch.uint8).pin_memory()
        x_dev, t_dev = x_host.to(dev), t_host.to(dev)
        h2d = x_host.numel() + t_host.numel()
        if train:
            model.train()
            opt = P.optim.SGD(model.parameters(), lr=1e-3, momentum=0.9, grad_scale=1.0 / world)
            if world > 1:
                bucketer = GradBucketer(model.parameters(), bucket_mb=args.bucket_mb, average=False, last_bucket_mb=args.last_bucket_mb).attach().sync_initial_state(model)
            step = P.TrainStep(model, opt, bucketer, graph=args.graph in ('on', 'auto'))
            eager = P.TrainStep(model, opt, bucketer, graph=False)

            def fn():
                return step(x_dev, t_dev)

            def fn_e2e():
                return step.step_prefetched(next_batch=(x_host, t_host))
        else:
            model.eval()
            cm = torch.zeros((nc, nc), dtype=torch.int64, device=dev)

            def infer(x, t):
                with torch.no_grad():
                    y = model(x)
                    tag = y._b200seg_class_map                # test.py:28-46 through evaluate.test's fused path
                    E.class_map_confusion(tag[0], nc, t, tag[2], cm)
                return y

            def fn():
                return infer(x_dev, t_dev)

            def fn_e2e():
                infer(x_host.to(dev, non_blocking=True), t_host.to(dev, non_blocking=True))
                return cm.sum()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(f, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            f()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / 1e3

    out_host = torch.zeros((), dtype=torch.float32).pin_memory()

    def e2e_step():
        out_host.copy_(fn_e2e().float(), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(args.warmup + (4 if step is not None and step.use_graph else 0)):
        fn()
    torch.cuda.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = _lib.launches()
    secs = timed(fn, args.steps)
    launches = step.launches_per_step * args.steps if step is not None and step.use_graph else _lib.launches() - l0
    clocks = sampler.stop() if sampler else None
    if step is not None:
        step.prefetch(x_host, t_host)
    for _ in range(2):
        e2e_step()
    secs_e2e = timed(e2e_step, args.steps)
    # per-launch CUDA-event timing of the tcgen05 convs (eager, weight-gradient lane off)
    lane_was, E._LANE_ON = E._LANE_ON, False
    E.PROFILE = []
    prof_fn = (lambda: eager(x_dev, t_dev)) if step is not None else fn
    for _ in range(2):
        torch.cuda._sleep(80_000_000)          # the host runs ahead: events bracket kernels, not launch latency
        prof_fn()
    torch.cuda.synchronize()
    rec, E.PROFILE = E.PROFILE, None
    E._LANE_ON = lane_was
    pk = peaks()
    peak_tf = pk['bf16_tflops_sustained'] or pk['bf16_tflops']
    agg = {}
    for tag, flops, e0, e1, info in rec:
        a = agg.setdefault(tag, [0.0, 0.0, 0, 0.0])
        a[0] += flops
        a[1] += e0.elapsed_time(e1) * 1e-3
        a[2] += 1
        a[3] += flops * info['run_share']
    ig = [v for t, v in agg.items() if t.startswith('conv_fprop') or t.startswith('conv_dgrad')]
    fl, sc, nn, ex = (sum(v[i] for v in ig) for i in range(4))
    detail = sorted(({'launch': t, 'tflops': round(v[0] / v[1] / 1e12, 1), 'us': round(v[1] / v[2] * 1e6, 1),
                      'frac': round(v[0] / v[1] / 1e12 / peak_tf, 3), 'frac_executed': round(v[3] / v[1] / 1e12 / peak_tf, 3)}
                     for t, v in agg.items() if v[1] > 0), key=lambda d: -d['us'])
    if rank == 0:
        line = {'metric': metric, 'value': world * B * args.steps / secs, 'unit': 'images/sec', 'n_gpus': world,
                'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': secs / args.steps * 1e3, 'higher_is_better': True,
                'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic',
                'config': {'workload': workload, 'per_gpu_batch': B, 'global_batch': B * world, 'parallelism': 'dp%d' % world,
                           'cuda_graph': bool(step is not None and step.use_graph),
                           'l2': 'no explicit flush: one step streams far more than the 126 MB L2' if kind != 'aspp' else
                                 'inputs + weights (67 + 47 MB) exceed what stays in the 126 MB L2 across a step'},
                'e2e': {'value': world * B * args.steps / secs_e2e, 'unit': 'images/sec', 'h2d_bytes_per_step': h2d,
                        'd2h_bytes_per_step': 4, 'ms_per_step': secs_e2e / args.steps * 1e3},
                'gpu_launches': launches, 'clocks': clocks,
                'roofline': {'bound': 'tensor', 'achieved': fl / sc / 1e12 if sc > 0 else None, 'peak': peak_tf,
                             'unit': 'TFLOP/s', 'frac': fl / sc / 1e12 / peak_tf if sc > 0 else None,
                             'frac_executed': ex / sc / 1e12 / peak_tf if sc > 0 else None, 'traffic': None,
                             'kernel': 'igemm_kernel (tcgen05 implicit-GEMM fprop+dgrad, %d launches/step)' % (nn // 2),
                             'peak_source': pk['source'] + ', sustained'},
                'kernels': detail[:12]}
        print(json.dumps(line), flush=True)
    if world > 1:
        _teardown(step, eager if step is not None else None, bucketer)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=40)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--batch', type=int, default=None, help='images per GPU per step (default: the configuration\'s)')
    ap.add_argument('--config', default='train', choices=['train', 'infer', 'aspp', 'hrnet', 'unet', 'unet_cpu'],
                    help='which BASELINE.json configuration (default: configs[3], the headline training step)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--detail', default=None, help='write the per-launch CUDA-event timings of every conv to this file')
    ap.add_argument('--bucket-mb', type=float, default=64.0, help='N > 1: gradient all-reduce bucket size')
    ap.add_argument('--last-bucket-mb', type=float, default=4.0, help='N > 1: cap of the bucket that completes last (exposed)')
    ap.add_argument('--sm-reserve', type=int, default=0, help='N > 1: SMs the conv kernels leave to NCCL')
    ap.add_argument('--nccl-ctas', type=int, default=0, help='N > 1: NCCL_MAX_CTAS (0 = NCCL default)')
    ap.add_argument('--graph', default='auto', choices=['auto', 'on', 'off'],
                    help='replay the step from a CUDA graph (auto = on; the bucketed NCCL all-reduces are captured with it)')
    args = ap.parse_args()
    args.batch_given = args.batch is not None
    if args.batch is None:
        args.batch = 16
    if args.impl == 'reference' or args.config == 'unet_cpu':
        run_reference(args)
    else:
        args.warmup = max(args.warmup, 3)
        (run_b200 if args.config == 'train' else run_other)(args)


if __name__ == '__main__':
    main()
