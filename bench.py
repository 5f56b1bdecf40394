#!/usr/bin/env python
"""bench.py — train images/sec @512x512 of DeepLabV3+ (21 classes, bf16) on N B200s (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch B]

One step = forward + compute_loss + backward (+ bucketed gradient all-reduce for N > 1) + SGD update on one
synthetic batch of B images per GPU (random-init weights, randn images, randint masks; per-rank seeds).
Launched by the driver as one process per GPU (torch.distributed.run) for N > 1.  Rank 0 prints ONE JSON line.

  value     whole-job images/sec with the step's inputs already resident in HBM
  e2e       the same through the public API with HOST buffers: pinned-host -> device copies of the images and
            masks and a device -> host read of the loss inside the timed region, every step
  roofline  the dominant kernel (tcgen05 implicit-GEMM conv, all fprop/dgrad launches of the step): nominal dense
            FLOPs (2*N*P*Q*K*C*R*S, padding taps counted — SURVEY.md §8d) / CUDA-event time, vs the measured bf16
            peak in MEASURED_PEAKS.json
  cpu_baseline / --impl reference
            the CPU oracle (oracle/restatement.py: the reference's own PyTorch calls, fp32) on the host cores on a
            bounded sample (batch 2 per step) of the same workload
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


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
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
            'cpu_baseline': {'value': ips, 'unit': 'images/sec', 'cores': threads, 'kind': 'port', 'sample': sample},
            'e2e': {'value': ips, 'unit': 'images/sec', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0}
    print(json.dumps(line), flush=True)


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
    bucketer = GradBucketer(model.parameters(), bucket_mb=args.bucket_mb, average=False).attach() if world > 1 else None

    g = torch.Generator(device='cpu').manual_seed(100 + rank)
    x_host = torch.randn(B, 3, H, W, generator=g).pin_memory()
    t_host = torch.randint(0, NC, (B, H, W), generator=g).pin_memory()
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

    # per-launch CUDA-event timing of the tcgen05 conv kernels over a few more steps (the events sit on the
    # launching stream; kept out of the throughput loop above so the headline has no instrumentation in it)
    # per-launch timing runs with the weight-gradient lane off: concurrent kernels would share SMs with the launch
    # being timed and understate it
    lane_was, E._LANE_ON = E._LANE_ON, False
    E.PROFILE = []
    E.OP_PROFILE = [] if args.detail else None
    prof_steps = min(args.steps, 3)
    for _ in range(prof_steps):
        eager(x_dev, t_dev)
    torch.cuda.synchronize()
    rec, E.PROFILE = E.PROFILE, None
    op_rec, E.OP_PROFILE = E.OP_PROFILE, None
    E._LANE_ON = lane_was
    agg = {}
    for tag, flops, e0, e1 in rec:
        a = agg.setdefault(tag, [0.0, 0.0, 0])
        a[0] += flops
        a[1] += e0.elapsed_time(e1) * 1e-3
        a[2] += 1
    pk = peaks()
    igemm = [(t, v) for t, v in agg.items() if t.startswith('conv_fprop') or t.startswith('conv_dgrad')]
    ig_flops, ig_secs, ig_n = (sum(v[i] for _, v in igemm) for i in range(3))
    peak_tf = pk['bf16_tflops_sustained'] or pk['bf16_tflops']
    achieved = ig_flops / ig_secs / 1e12 if ig_secs > 0 else 0.0
    detail = sorted(({'launch': t, 'tflops': round(v[0] / v[1] / 1e12, 1), 'us': round(v[1] / v[2] * 1e6, 1),
                      'frac': round(v[0] / v[1] / 1e12 / peak_tf, 3)} for t, v in agg.items() if v[1] > 0),
                    key=lambda d: -d['us'])
    traffic = None
    tpath = os.path.join(ROOT, 'profiles', 'igemm_traffic.json')
    if os.path.exists(tpath):
        traffic = json.load(open(tpath)).get('dram_bytes_per_launch')

    if rank == 0:
        ips = world * B * args.steps / secs
        ips_e2e = world * B * args.steps / secs_e2e
        line = {
            'metric': METRIC, 'value': ips, 'unit': 'images/sec', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': secs / args.steps * 1e3, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16', 'data': 'synthetic',
            'config': {'workload': 'DeepLabV3+ 21-class 512x512 bf16 training, fwd+loss+bwd+SGD (BASELINE configs[3])',
                       'per_gpu_batch': B, 'global_batch': B * world, 'parallelism': 'dp%d' % world,
                       'cuda_graph': bool(use_graph), 'host_enqueue_ms_per_eager_step': round(host_enqueue_ms, 2),
                       'libb200seg_launches_per_step': eager_launches,
                       'l2': 'no explicit flush: one step streams several GB of activations (>> 126 MB L2)',
                       'backbone': 'ResNet-50 OS16 on libb200seg (tcgen05 convs incl. stride 2, fused BN/ReLU/residual, '
                                   'max pool); 7x7/2 stem = 4 row taps over the space-to-depth image (tcgen05); no cuDNN on the path'},
            'e2e': {'value': ips_e2e, 'unit': 'images/sec', 'h2d_bytes_per_step': x_host.numel() * 4 + t_host.numel() * 8,
                    'd2h_bytes_per_step': 4, 'ms_per_step': secs_e2e / args.steps * 1e3},
            'gpu_launches': launches,
            'clocks': clocks,
            'roofline': {'bound': 'tensor', 'achieved': achieved, 'peak': peak_tf, 'unit': 'TFLOP/s',
                         'frac': achieved / peak_tf if peak_tf else None, 'traffic': traffic,
                         'kernel': 'igemm_kernel (tcgen05 implicit-GEMM fprop+dgrad, %d launches/step)' % (ig_n // max(prof_steps, 1)),
                         'flops_per_step': ig_flops / max(prof_steps, 1), 'peak_source': pk['source'] + ', sustained'},
            'kernels': detail[:12],
        }
        if world == 1 and not args.no_cpu_baseline:
            ips_cpu, dt, threads = cpu_reference_run(16, 1)      # ~10 s of host work
            line['cpu_baseline'] = {'value': ips_cpu, 'unit': 'images/sec', 'cores': threads, 'kind': 'port',
                                    'sample': 'oracle DeepLabV3+ 512x512 train step, batch %d, 1 warm-up + 16 timed steps '
                                              '(%.2f s/step)' % (CPU_SAMPLE_BATCH, dt)}
        if args.detail:
            with open(args.detail, 'w') as f:
                ops, ops_by_shape = {}, {}
                for tag, e0, e1 in op_rec or []:
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
        torch.cuda.synchronize()
        dist.barrier()
        if use_graph:
            # a captured step holds NCCL kernels: tearing the communicator down under a live graph (or the graph
            # after the communicator) can block at interpreter exit, so the ranks leave together right here
            sys.stdout.flush()
            sys.stderr.flush()
            os._exit(0)
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=40)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--batch', type=int, default=16, help='images per GPU per step')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--detail', default=None, help='write the per-launch CUDA-event timings of every conv to this file')
    ap.add_argument('--bucket-mb', type=float, default=64.0, help='N > 1: gradient all-reduce bucket size')
    ap.add_argument('--sm-reserve', type=int, default=0, help='N > 1: SMs the conv kernels leave to NCCL')
    ap.add_argument('--nccl-ctas', type=int, default=0, help='N > 1: NCCL_MAX_CTAS (0 = NCCL default)')
    ap.add_argument('--graph', default='auto', choices=['auto', 'on', 'off'],
                    help='replay the step from a CUDA graph (auto = on; the bucketed NCCL all-reduces are captured with it)')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        args.warmup = max(args.warmup, 3)
        run_b200(args)


if __name__ == '__main__':
    main()
