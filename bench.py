#!/usr/bin/env python
"""Benchmark of the serving hot path (BASELINE.json metric: ensemble queries/sec at 1/2/4/8 B200).

    python bench.py --gpus N --steps K --warmup W [--workload c3|c2] [--impl reference]

prints ONE JSON line (rank 0).  A step is one pass of the hot path over one batch of synthetic
queries: uint8 images in, ensemble-mean class probabilities out.

Workloads (BASELINE.json `configs`):
  c3 (default)  8-trial TfFeedForward ensemble (D 784, H 128, L 2), batch 65536 -- the
                configuration the north_star's targets are quoted on; it fits one GPU.  At N GPUs
                the 8 trials are placed 8/N per GPU (one per GPU at N = 8), every GPU sees the whole
                batch, and the partial means are summed onto rank 0 through NVLink peer memory (the trial kernel
                stores them into the owners' slots; rafiki_b200/csrc/peer.cu): strong scaling.
  c2            3-trial ensemble, batch 4096 per GPU; N > 1 runs independent replicas (weak scaling).

`value` is device-resident throughput (queries already in HBM); `e2e` goes through the pinned-host
pipeline (H2D of the uint8 batch and D2H of the probabilities inside the timed region).  The
`cpu_baseline` leg and `--impl reference` time the CPU oracle (`oracle/`, a NumPy restatement of
the reference's TensorFlow-1.12 path, which cannot be installed here) on the box's host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

WORKLOADS = {
    'c3': dict(trials=8, batch=65536, image_size=28, channels=1, hidden=128, layers=2, scaling='strong',
               name='8-trial TfFeedForward ensemble, 28x28, batch 65536 (BASELINE configs[2])'),
    'c2': dict(trials=3, batch=4096, image_size=28, channels=1, hidden=128, layers=2, scaling='weak',
               name='3-trial TfFeedForward ensemble, 28x28, batch 4096 per GPU (BASELINE configs[1])'),
    'c4': dict(trials=1, batch=65536, image_size=32, channels=3, hidden=2048, layers=3, scaling='strong', shard='rows',
               name='wide MLP trial (3x2048 hidden), 32x32x3 queries, batch 65536 sharded data-parallel '
                    '(BASELINE configs[3])'),
    'small': dict(trials=0, batch=262144, image_size=28, channels=1, hidden=0, layers=0, scaling='weak', small=True,
                  name='memory-bound kernels alone (decode / NEAREST resize / ensemble mean) at a working set larger than L2'),
    'c5': dict(trials=3, batch=65536, image_size=28, channels=1, hidden=128, layers=2, scaling='weak', sweep=True,
               name='serving latency sweep, batch 1..65536, 3-trial ensemble (BASELINE configs[4])'),
}
L2_BYTES = 126 * 1024 * 1024
FALLBACK_PEAKS = {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0}


def load_peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        p['_source'] = 'measured (MEASURED_PEAKS.json)'
        return p
    p = dict(FALLBACK_PEAKS)
    p['_source'] = 'fallback (B200_PROFILING.md)'
    return p


def make_workload(wl, rng_seed=0, n_batches=None):
    from rafiki_b200 import synth
    d = wl['image_size'] ** 2 * wl['channels']
    dead = synth.dead_mask(d)
    params = [synth.make_trial(t, wl['image_size'], wl['channels'], wl['layers'], wl['hidden'], dead=dead)
              for t in range(wl['trials'])]
    if n_batches is None:   # enough distinct batches that the rotation exceeds L2
        n_batches = max(2, -(-int(1.6 * L2_BYTES) // (wl['batch'] * d)))
    batches = [synth.make_queries(wl['batch'], wl['image_size'], wl['channels'], 'uniform', seed=rng_seed + i,
                                  dead=dead).reshape(wl['batch'], d) for i in range(n_batches)]
    return params, batches, d


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons sampled during the timed region: NVML every 50 ms when the bindings
    load, else nvidia-smi every 200 ms.  Built and started BEFORE the warm-up on every rank (NVML
    start-up costs milliseconds and must not sit between the barrier and the first timed step)."""
    QUERY = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')
    NAMES = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
    NVML_BITS = [0x8, 0x40, 0x20, 0x4]      # same order as NAMES (nvmlClocksThrottleReason*)

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []          # (sm_mhz, sm_max_mhz, [bool per NAMES])
        self._halt = threading.Event()
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                pr = torch.cuda.get_device_properties(index)
                bus = '%08x:%02x:%02x.0' % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
                self._h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
            except Exception:
                self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self._max = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
            self._nvml = pynvml
        except Exception:
            self._nvml = None

    def _sample_nvml(self):
        n = self._nvml
        sm = n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)
        bits = n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
        self.samples.append((float(sm), float(self._max), [bool(bits & b) for b in self.NVML_BITS]))

    def _sample_smi(self):
        out = subprocess.run(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.QUERY,
                              '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5)
        parts = [x.strip() for x in out.stdout.strip().split(',')]
        if len(parts) >= 6 and parts[0].replace('.', '', 1).isdigit():
            self.samples.append((float(parts[0]), float(parts[1]) if parts[1].replace('.', '', 1).isdigit() else None,
                                 [parts[2 + i].lower().startswith('active') for i in range(4)]))

    def run(self):
        while not self._halt.is_set():
            try:
                if self._nvml is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._halt.wait(0.05 if self._nvml is not None else 0.2)

    def mark(self):
        """Samples taken before this call (warm-up) are dropped from the summary."""
        self._from = len(self.samples)
        try:
            if self._nvml is not None:
                self._sample_nvml()
        except Exception:
            pass

    def stop(self):
        self._halt.set()
        self.join(3)
        try:
            if self._nvml is not None:
                self._sample_nvml()
        except Exception:
            pass
        self.samples = self.samples[getattr(self, '_from', 0):]
        sm = [s[0] for s in self.samples]
        mx = [s[1] for s in self.samples if s[1] is not None]
        reasons = [n for i, n in enumerate(self.NAMES) if any(s[2][i] for s in self.samples)]
        return {'sm_mhz': statistics.median(sm) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': reasons, 'samples': len(self.samples), 'source': 'nvml' if self._nvml is not None else 'nvidia-smi'}


def cpu_oracle_rate(params, batch, sample, repeats, threads):
    """queries/s of the CPU oracle (list in, list out, exactly the reference's composition) on the
    first `sample` queries of the batch."""
    from oracle import feedforward as ff     # the one place bench.py executes oracle/
    from threadpoolctl import threadpool_limits
    ch = int(params[0].get('channels', 1)) if isinstance(params[0], dict) else 1
    side = int(round((batch.shape[1] // ch) ** 0.5))
    shape = (sample, side, side) if ch == 1 else (sample, side, side, ch)
    queries = batch[:sample].reshape(shape).tolist()
    times = []
    with threadpool_limits(limits=threads):
        for _ in range(repeats):
            t0 = time.perf_counter()
            out = ff.serve_batch(params, queries)
            times.append(time.perf_counter() - t0)
    assert len(out) == sample
    return sample / min(times), times


def run_reference(args, wl):
    """The reference arm: the CPU port of the reference path (the oracle; TensorFlow 1.12 cannot be
    installed) on the box's host cores, list in / list out, a bounded sample of the batch per step."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return 0
    params, batches, d = make_workload(wl, n_batches=1)
    cores = os.cpu_count() or 1
    sample = min(wl['batch'], int(os.environ.get('RFK_REFERENCE_SAMPLE', '512')))   # ~0.2-0.4 s of CPU work per step
    for _ in range(max(1, min(args.warmup, 3))):
        cpu_oracle_rate(params, batches[0], min(sample, 256), 1, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_oracle_rate(params, batches[0], sample, 1, cores)
    dt = time.perf_counter() - t0
    value = sample * args.steps / dt
    one_thread, _ = cpu_oracle_rate(params, batches[0], min(sample, 256), 1, 1)      # BLAS pinned to one thread
    line = {
        'impl': 'reference', 'metric': 'ensemble_queries_per_sec', 'value': value, 'unit': 'queries/s',
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt / args.steps * 1e3,
        'higher_is_better': True, 'scaling': wl['scaling'], 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': wl['name'], 'trials': wl['trials'], 'batch': wl['batch'], 'sample_queries': sample,
                   'note': 'CPU oracle port of the reference path (TensorFlow 1.12 not installable); the rate is per '
                           'query, so each step runs the first {} queries of the {}-query batch through all {} trials '
                           '(list in / list out), not the whole batch'.format(sample, wl['batch'], wl['trials'])},
        'cpu_baseline': {'value': value, 'unit': 'queries/s', 'cores': cores, 'kind': 'port',
                         'sample': '{} queries x {} trials per step'.format(sample, wl['trials']),
                         'value_blas_1_thread': one_thread},
        'e2e': {'value': value, 'unit': 'queries/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def run_latency_sweep(args, wl, ens, xs, out, params, rank):
    """BASELINE configs[4]: p50 / p99 latency and queries/s per batch size, 3-trial ensemble."""
    import torch
    rows = []
    n_max = wl['batch']
    sizes = [1 << k for k in range(0, 17) if (1 << k) <= n_max]
    for bsz in sizes:
        x = [b[:bsz].contiguous() for b in xs]
        o = out[:bsz]
        for i in range(20):
            ens.forward(x[i % len(x)], out=o)
        torch.cuda.synchronize()
        lat, wall = [], []
        for i in range(min(max(args.steps, 200), 400)):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            a.record()
            ens.forward(x[i % len(x)], out=o)
            b.record()
            b.synchronize()
            wall.append((time.perf_counter() - t0) * 1e3)
            lat.append(a.elapsed_time(b))
        lat.sort()
        wall.sort()
        p50, p99 = lat[len(lat) // 2], lat[min(len(lat) - 1, int(len(lat) * 0.99))]
        rows.append({'batch': bsz, 'p50_ms': p50, 'p99_ms': p99, 'queries_per_s_at_p50': bsz / (p50 * 1e-3),
                     'host_wall_p50_ms': wall[len(wall) // 2], 'host_wall_p99_ms': wall[min(len(wall) - 1, int(len(wall) * 0.99))]})
    if rank == 0:
        cores = os.cpu_count() or 1
        cpu_rows = []
        if not args.no_cpu_baseline:
            host_batch = xs[0].cpu().numpy()
            for bsz in (1, 32, 1024, 4096):
                rate, times = cpu_oracle_rate(params, host_batch, bsz, 3 if bsz <= 1024 else 1, cores)
                cpu_rows.append({'batch': bsz, 'ms': min(times) * 1e3, 'queries_per_s': rate})
        print(json.dumps({'metric': 'ensemble_batch_latency_ms', 'unit': 'ms', 'higher_is_better': False,
                          'n_gpus': 1, 'dtype': args.compute, 'data': 'synthetic',
                          'config': {'workload': wl['name'], 'trials': wl['trials']},
                          'value': rows[-1]['p99_ms'], 'sweep': rows,
                          'cpu_baseline': {'kind': 'port', 'cores': cores, 'unit': 'ms', 'rows': cpu_rows,
                                           'sample': 'oracle serve_batch on the first b queries, best of 3'}}),
              flush=True)
    return 0


def run_small_kernels(args, wl):
    """Kernels 1 and 3 alone against the HBM roofline: achieved bytes/s of decode (-> bf16 / f32), the NEAREST
    resize and the ensemble mean, inputs rotated over more than L2, CUDA events; one JSON line."""
    import torch
    from rafiki_b200 import engine
    torch.cuda.set_device(0)
    peaks = load_peaks()
    peak = float(peaks.get('hbm_gbs', FALLBACK_PEAKS['hbm_gbs']))

    def timed(fn, iters):
        for i in range(max(args.warmup, 3)):
            fn(i)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(iters):
            fn(i)
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / iters * 1e-3

    iters = max(5, min(args.steps, 100))
    n = wl['batch']
    rows = []
    xs = [torch.randint(0, 256, (n, 784), dtype=torch.uint8, device='cuda') for _ in range(2)]      # 2 x 205 MB
    scale = torch.rand(784, device='cuda') + 0.5
    shift = torch.randn(784, device='cuda')
    for dt, sz in ((torch.bfloat16, 2), (torch.float32, 4)):
        t = timed(lambda i: engine.decode_u8(xs[i & 1], scale, shift, out_dtype=dt), iters)
        byts = n * 784 * (1 + sz)
        rows.append({'kernel': 'decode_u8 -> %s' % str(dt).split('.')[-1], 'n': n, 'us': t * 1e6, 'algorithmic_bytes': byts,
                     'gbs': byts / t / 1e9, 'frac': byts / t / 1e9 / peak})
    x28 = [x.view(n, 28, 28) for x in xs]
    t = timed(lambda i: engine.resize_decode_u8(x28[i & 1], 32), iters)
    byts = n * (784 + 1024)
    rows.append({'kernel': 'resize 28->32 (uint8)', 'n': n, 'us': t * 1e6, 'algorithmic_bytes': byts, 'gbs': byts / t / 1e9,
                 'frac': byts / t / 1e9 / peak})
    del xs, x28
    for w, r in ((3, 4 * 1048576), (8, 2 * 1048576)):
        sets = [[torch.rand(r, 10, device='cuda') for _ in range(w)] for _ in range(2)]
        dst = torch.empty(r, 10, device='cuda')
        t = timed(lambda i: engine.ensemble_mean(sets[i & 1], out=dst), iters)
        byts = r * 10 * 4 * (w + 1)
        rows.append({'kernel': 'ensemble_mean W=%d' % w, 'n': r, 'us': t * 1e6, 'algorithmic_bytes': byts, 'gbs': byts / t / 1e9,
                     'frac': byts / t / 1e9 / peak})
        del sets, dst
    worst = min(rows, key=lambda k: k['frac'])
    print(json.dumps({'metric': 'memory_bound_kernels_hbm_fraction', 'value': worst['frac'], 'unit': 'fraction of measured HBM copy bandwidth (worst kernel)',
                      'n_gpus': 1, 'steps': iters, 'warmup': max(args.warmup, 3), 'higher_is_better': True, 'dtype': 'u8/f32',
                      'data': 'synthetic', 'config': {'workload': wl['name']},
                      'roofline': {'bound': 'hbm', 'achieved': worst['gbs'], 'peak': peak, 'unit': 'GB/s', 'frac': worst['frac'],
                                   'traffic': None, 'kernel': worst['kernel'], 'peak_source': peaks['_source']},
                      'kernels': rows}), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=2000)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default='c3', choices=sorted(WORKLOADS))
    ap.add_argument('--compute', default='f16', choices=['f16', 'bf16', 'f32'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == 'reference':
        return run_reference(args, wl)
    if wl.get('small'):
        return run_small_kernels(args, wl)

    import torch
    import torch.distributed as dist
    from rafiki_b200 import engine, synth

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if args.gpus != world:
        if world == 1 and args.gpus > 1:
            raise SystemExit('launch with torchrun --nproc-per-node {} for --gpus {}'.format(args.gpus, args.gpus))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a GPU (no CPU fallback)')
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=dev)

    params, batches, d = make_workload(wl, n_batches=2 if wl.get('shard') == 'rows' else None)
    n, W, C = wl['batch'], wl['trials'], synth.NUM_CLASSES
    n_total = n
    if wl.get('shard') == 'rows':      # one trial, rows of the batch split over ranks, no collective
        mine = list(range(W))
        lo, hi = rank * n // world, (rank + 1) * n // world
        batches = [np.ascontiguousarray(b[lo:hi]) for b in batches]
        n = hi - lo
    elif wl['scaling'] == 'strong':    # trials dealt round-robin to ranks; every rank sees the whole batch
        mine = [t for t in range(W) if t % world == rank]
    else:                              # independent replicas
        mine = list(range(W))
    from rafiki_b200 import _native
    sampler = ClockSampler(local_rank)        # every rank, before the warm-up
    sampler.start()
    trials = [engine.Trial(params[t], compute=args.compute) for t in mine]
    ens = engine.Ensemble(trials, total_trials=W) if trials else None
    xs = [torch.from_numpy(b).to(dev) for b in batches]
    out = torch.zeros((n, C), dtype=torch.float32, device=dev)
    do_reduce = world > 1 and wl['scaling'] == 'strong' and wl.get('shard') != 'rows'
    peer = None
    if do_reduce:
        from rafiki_b200.distributed import PeerEnsemble
        peer = PeerEnsemble(ens, W, n, d, C, dev, root=0)       # one native call per step
        peer.set_inputs_resident(True)                          # the timed batches are resident before the first step

    last = [None]

    def step(i):
        if peer is not None:
            # one native call and (one trial per GPU) ONE kernel per rank and step: partial means go straight
            # into the owners' slots over NVLink, the cross-rank sum of batch i rides in the kernel of batch i + 1
            prev = peer.step_pipelined(xs[i % len(xs)])
            if prev is not None:
                last[0] = prev
        elif ens is not None:
            last[0] = ens.forward(xs[i % len(xs)], out=out)

    def drain():
        if peer is not None:
            res = peer.flush()          # completes the last batch (stand-alone sum of its slots)
            if res is not None:
                last[0] = res

    if wl.get('sweep'):
        sampler.stop()
        return run_latency_sweep(args, wl, ens, xs, out, params, rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warm = max(args.warmup, 3)
    for i in range(warm):
        step(i)
    drain()
    barrier()
    # rank skew from the barrier itself stays outside the timed region: a few synchronising steps
    # (at N > 1 every step ends in a cross-rank hand-shake), then the events
    if world > 1:
        for i in range(5):
            step(i)
        drain()
        torch.cuda.synchronize()
        dist.barrier()
    launches0 = _native.launch_count()
    sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):       # nothing but the steps between the two events: an event record between two
        step(i)                       # kernels would keep the second from starting while the first one drains
    drain()
    e1.record()
    torch.cuda.synchronize()
    launches = _native.launch_count() - launches0
    clocks = sampler.stop()
    barrier()
    # per-step times, measured apart (events between the steps): latency of one step, not the pipelined rate
    step_ev = [torch.cuda.Event(enable_timing=True) for _ in range(min(args.steps, 64) + 1)]
    step_ev[0].record()
    for i in range(len(step_ev) - 1):
        step(i)
        step_ev[i + 1].record()
    drain()
    torch.cuda.synchronize()
    barrier()
    ms = e0.elapsed_time(e1)
    per_step = sorted(step_ev[k].elapsed_time(step_ev[k + 1]) for k in range(len(step_ev) - 1))
    ms_median_step = per_step[len(per_step) // 2]
    if world > 1:
        t = torch.tensor([ms, ms_median_step], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_median_step = float(t[0].item()), float(t[1].item())
    units = (n_total if wl.get('shard') == 'rows' else n) * args.steps * (world if wl['scaling'] == 'weak' else 1)
    value = units / (ms * 1e-3)

    # ---- the N > 1 result against the single-GPU ensemble of the same batch (rank 0 holds all parameters)
    parity = None
    if do_reduce:
        step(0)
        drain()
        torch.cuda.synchronize()
        dist.barrier()
        if rank == 0:
            full = [engine.Trial(params[t], compute=args.compute) for t in range(W)]
            want = engine.Ensemble(full).forward(xs[0])
            parity = float((last[0].double() - want.double()).abs().max().item())
            for t in full:
                t.close()
        dist.barrier()

    # ---- per-batch latency (the metric's "p99 batch latency"): one step at a time, device timed
    lat = []
    for i in range(min(args.steps, 200)):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step(i)
        drain()
        b.record()
        b.synchronize()
        lat.append(a.elapsed_time(b))
    lat.sort()
    latency = {'p50_ms': lat[len(lat) // 2], 'p99_ms': lat[min(len(lat) - 1, int(len(lat) * 0.99))],
               'samples': len(lat), 'what': 'one batch, device resident, CUDA events'}

    # ---- end to end: pinned host -> device -> pinned host, three batches in flight
    e2e = None
    if peer is not None:
        # rank r: ITS rows of the batch, pinned host -> device over its own PCIe link (copy stream, overlapping
        # the previous step's kernels); all-gather of the slices over NVLink; fused kernel; peer reduce;
        # rank 0: probabilities -> pinned host
        h_in = torch.from_numpy(batches[0]).pin_memory()
        h_out = torch.empty((n, C), dtype=torch.float32).pin_memory()
        e2e_steps = max(4, min(args.steps, 40))
        peer.set_inputs_resident(False)          # from here on the queries arrive in-stream
        cur = torch.cuda.current_stream()
        copy = torch.cuda.Stream(device=dev)
        ev_up = [torch.cuda.Event(), torch.cuda.Event()]
        ev_g = torch.cuda.Event()
        up_bytes = 0

        def upload(k):
            nonlocal up_bytes
            copy.wait_event(ev_g)
            with torch.cuda.stream(copy):
                up_bytes = peer.upload(h_in)
                ev_up[k].record(copy)

        ev_g.record(cur)
        upload(0)
        t0 = None
        for it in range(3 + e2e_steps):
            if it == 3:
                barrier()
                t0 = time.perf_counter()
            cur.wait_event(ev_up[it & 1])
            xg = peer.gather(n)
            ev_g.record(cur)
            upload((it + 1) & 1)                 # next step's slice while this step computes
            res = peer.step(xg)
            if rank == 0:
                h_out.copy_(res, non_blocking=True)
        barrier()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt, float(up_bytes)], device=dev)
        dist.all_reduce(t[:1], op=dist.ReduceOp.MAX)
        dist.all_reduce(t[1:], op=dist.ReduceOp.SUM)
        dt = float(t[0].item())
        e2e_parity = None
        if rank == 0:
            full = [engine.Trial(params[t_], compute=args.compute) for t_ in range(W)]
            want = engine.Ensemble(full).forward(xs[0])
            e2e_parity = float((torch.from_numpy(h_out.numpy()).to(dev).double() - want.double()).abs().max().item())
            for t_ in full:
                t_.close()
        e2e = {'value': n * e2e_steps / dt, 'unit': 'queries/s', 'h2d_bytes_per_step': int(t[1].item()),
               'd2h_bytes_per_step': n * C * 4, 'steps': e2e_steps, 'parity_vs_single_gpu': e2e_parity,
               'note': 'each rank uploads only its 1/N of the rows from pinned host memory (own PCIe link, copy '
                       'stream), the slices are all-gathered over NVLink peer memory (rfk_peer_allgather_u8), '
                       'fused kernel, peer-memory reduce, rank 0 downloads the probabilities'}
    elif ens is not None:
        depth = 3
        pipe = engine.HostPipeline(ens, max_batch=n, depth=depth)
        for s in range(depth):
            pipe.stage(s, batches[s % len(batches)])
        for i in range(3):
            pipe.launch(i % depth)
            pipe.wait(i % depth)
        barrier()
        e2e_steps = max(4, min(args.steps, 40))
        t0 = time.perf_counter()
        inflight = []
        for i in range(e2e_steps):
            s = i % depth
            if len(inflight) == depth:
                pipe.wait(inflight.pop(0))
            pipe.launch(s)
            inflight.append(s)
        for s in inflight:
            pipe.wait(s)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e_units = (n_total if wl.get('shard') == 'rows' else n) * e2e_steps * (world if wl['scaling'] == 'weak' else 1)
        e2e = {'value': e_units / dt, 'unit': 'queries/s', 'h2d_bytes_per_step': pipe.h2d_bytes,
               'd2h_bytes_per_step': pipe.d2h_bytes, 'steps': e2e_steps, 'slots_in_flight': depth,
               'note': 'per rank: H2D of the whole uint8 batch + fused kernel + D2H of the probabilities'}
    # ---- the drop-in's own API, N = 1: Predictor.predict_batch with a uint8 ndarray (binary ingress) and with the
    #      reference's nested Python lists (what the CPU arm is fed), wall clock around the call
    api = {}
    if world == 1 and ens is not None and wl['channels'] == 1 and not wl.get('shard'):
        from rafiki_b200.predictor import Predictor
        pred = Predictor('bench-predictor', db=object())
        pred.attach_ensemble(ens, max_batch=n)
        S = wl['image_size']
        frames = batches[0].reshape(n, S, S)
        pred.predict_batch(frames)
        reps = 5
        t0 = time.perf_counter()
        for _ in range(reps):
            got = pred.predict_batch(frames)['predictions']
        dt = (time.perf_counter() - t0) / reps
        api['e2e_ndarray_api'] = {'value': n / dt, 'unit': 'queries/s', 'batch': n, 'ms': dt * 1e3,
                                  'what': 'Predictor.predict_batch(uint8 ndarray [n,28,28]) -> float32 ndarray, wall clock'}
        m = min(n, 8192)
        as_lists = frames[:m].tolist()
        t0 = time.perf_counter()
        got_l = pred.predict_batch(as_lists)['predictions']
        dt = time.perf_counter() - t0
        assert len(got_l) == m and abs(got_l[0][0] - float(got[0][0])) < 1e-6
        api['e2e_list_api'] = {'value': m / dt, 'unit': 'queries/s', 'batch': m, 'ms': dt * 1e3,
                               'what': 'Predictor.predict_batch(nested Python lists) -> nested lists (the reference\'s '
                                       'types; list -> uint8 conversion on the host dominates), wall clock'}

    if rank == 0:
        peaks = load_peaks()
        flops_per_query_local = sum(synth.trial_flops_per_query(params[t]) for t in mine)
        flops_per_launch = flops_per_query_local * n
        ms_launch = ms / args.steps
        achieved = flops_per_launch / (ms_launch * 1e-3) / 1e12
        peak_sus = float(peaks.get('bf16_tflops_sustained', FALLBACK_PEAKS['bf16_tflops_sustained']))
        peak_burst = float(peaks.get('bf16_tflops', FALLBACK_PEAKS['bf16_tflops']))
        # a timed region of a few ms on a cool chip is the burst regime; seconds at the power cap the sustained one
        capped = bool(clocks and 'sw_power_cap' in (clocks.get('reasons') or []))
        use_burst = ms < 1000.0 and not capped
        peak = peak_burst if use_burst else peak_sus
        traffic, traffic_source = None, None
        summ = os.path.join(ROOT, 'profiles', 'ncu_summary.json')
        if os.path.exists(summ) and world == 1:
            try:
                ent = json.load(open(summ)).get(args.workload, {})
                traffic = ent.get('fused_dram_bytes_per_launch')
                traffic_source = 'ncu capture {} (not measured in this run)'.format(ent.get('source', 'profiles/'))
            except Exception:
                traffic = None
        if do_reduce:
            par = ('trial-per-gpu x{}: partial means stored into peer memory over NVLink from the trial kernel, summed by '
                   'their owners; one native call per step (rfk_peer_ensemble_step_pipelined)'.format(world))
        elif wl.get('shard') == 'rows':
            par = 'rows sharded x{}'.format(world)
        else:
            par = 'replicas x{}'.format(world) if world > 1 else 'single gpu, fused ensemble'
        line = {
            'metric': 'ensemble_queries_per_sec', 'value': value, 'unit': 'queries/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': warm, 'ms_per_step': ms / args.steps, 'ms_median_step': ms_median_step,
            'higher_is_better': True, 'scaling': wl['scaling'], 'vs_baseline': None, 'dtype': args.compute,
            'data': 'synthetic',
            'config': {'workload': wl['name'], 'trials': W, 'batch': n, 'd': d, 'hidden_units': wl['hidden'],
                       'hidden_layers': wl['layers'], 'classes': C, 'parallelism': par,
                       'l2': 'rotating over {} distinct resident batches ({} MB > L2)'.format(
                           len(xs), len(xs) * n * d // (1 << 20))},
            'clocks': clocks,
            'latency': latency,
            'e2e': e2e,
            'gpu_launches': int(launches),
            'gpu_launches_source': 'rfk_launch_count() before/after the timed region on rank 0',
            'roofline': {'bound': 'tensor', 'achieved': achieved, 'peak': peak, 'unit': 'TFLOP/s',
                         'frac': achieved / peak, 'frac_burst': achieved / peak_burst,
                         'frac_sustained': achieved / peak_sus,
                         'peak_choice': 'burst' if use_burst else 'sustained',
                         'traffic': traffic, 'traffic_source': traffic_source,
                         'kernel': ('gemm_bias_act_kernel x4 + decode (whole step; duration = timed region / steps)'
                                    if wl.get('shard') == 'rows' else
                                    'fused_pair_kernel + ensemble_mean_vec4_kernel (whole step; duration = timed region / steps)'),
                         'peak_source': peaks['_source'] + ' (bf16 rates; same tensor rate for fp16)',
                         'flops_per_launch': flops_per_launch,
                         'hbm_gbs_algorithmic': n * (d + 4 * C) / (ms_launch * 1e-3) / 1e9},
        }
        line.update(api)
        if parity is not None:
            line['parity_vs_single_gpu'] = parity
            if not parity <= 1e-6:
                print(json.dumps(line), flush=True)
                raise SystemExit('N = {} result differs from the single-GPU ensemble by {}'.format(world, parity))
        if not args.no_cpu_baseline and world == 1:
            cores = os.cpu_count() or 1
            sample = min(n, 4096)
            rate, times = cpu_oracle_rate(params, batches[0], sample, 2, cores)
            line['cpu_baseline'] = {'value': rate, 'unit': 'queries/s', 'cores': cores, 'kind': 'port',
                                    'sample': 'first {} queries of batch 0 x {} trials, best of 2 ({:.1f} s each); '
                                              'NumPy restatement of the TF-1.12 path, list in / list out'.format(
                                                  sample, W, min(times))}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == '__main__':
    sys.exit(main())
