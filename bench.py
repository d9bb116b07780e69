#!/usr/bin/env python
"""Benchmark of the tally / decision hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--config c2|c3|c4|c5] [--metric keys|decisions]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One JSON line on stdout (rank 0).  BASELINE.json names two rates; either can be the headline:

  --metric keys       "keys matched/sec": a step is one pass of K1 (prefix match and count)
                      over the rank's resident key snapshot -- config 2: 1e8 synthetic keys x 64
                      queue prefixes, 32-byte heads, per GPU (weak scaling) -- and, with N > 1,
                      the NCCL all-reduce of the uint64[64] counts.
  --metric decisions  "scaling decisions/sec": a step is one replay of the rank's fleet under
                      the default policy over the resident trace -- config 3: 1e6 single-queue
                      deployments x 1e4 ticks per GPU -- and the all-reduce of the [S][3] sums.

The other leg rides in the same line ("replay" / "tally"), with the policy sweep, the uint16
trace, multi-queue fleets and the smaller kernels beside it.  --config picks the workload:
c2 / c3 as above (c3 defaults to --metric decisions), c4 = 2e6 queues x 1e4 ticks per GPU
(16M queues on 8 GPUs), c5 = 1.25e8 long-tailed keys per GPU cut to heads on the device by
kas_pack_heads, and a fleet of multi-queue resources (1-8 queues) of mixed job / deployment type.

Every leg is checked against the CPU oracle on sampled units (every rank, every N); a mismatch
aborts the run.  `--impl reference` times the CPU restatement of the reference's path (the
reference itself is Python + network I/O and cannot travel to the box; the oracle's C port with
every host thread is the strongest CPU form of it) without loading the product library.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, 'kiosk-autoscaler_b200')
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

N_PREFIX = 64
SEED = 2
DEFAULT_POLICY = (1, 0, 1)                      # scale.py:90-92 of the reference
SWEEP = [DEFAULT_POLICY] + [(k, a, b) for k in (1, 2, 5, 10, 50) for a in (0, 1) for b in (1, 4, 16, 64)]
FOUR = [(1, 0, 1), (2, 0, 4), (5, 1, 16), (10, 0, 64)]

CONFIGS = {
    'c2': dict(keys=100_000_000, long_tail=False, fleet_q=1_000_000, ticks=10_000, grouped=False, metric='keys',
               title='C2 keyspace tally: %(keys)d synthetic keys per GPU x 64 queue prefixes, 32-byte heads, '
                     'bit-exact vs tally_queues'),
    'c3': dict(keys=100_000_000, long_tail=False, fleet_q=1_000_000, ticks=10_000, grouped=False, metric='decisions',
               title='C3 policy replay: %(fleet_q)d single-queue deployments per GPU x %(ticks)d ticks, '
                     'KEYS_PER_POD/MIN_PODS/MAX_PODS sweep'),
    'c4': dict(keys=100_000_000, long_tail=False, fleet_q=2_000_000, ticks=10_000, grouped=False, metric='decisions',
               title='C4 sharded replay: %(fleet_q)d single-queue deployments per GPU x %(ticks)d ticks '
                     '(16M queues on 8 GPUs), all-reduce of pod-hours'),
    'c5': dict(keys=125_000_000, long_tail=True, fleet_q=2_000_000, ticks=10_000, grouped=True, metric='keys',
               title='C5: %(keys)d long-tailed keys per GPU (1e9 on 8 GPUs) cut to heads by kas_pack_heads, '
                     'fleet of %(fleet_q)d queues per GPU in resources of 1-8 queues, mixed job/deployment'),
}


def resolve_config(args):
    name = 'c3' if (args.config == 'c2' and args.metric == 'decisions') else args.config   # same sizes
    cfg = dict(CONFIGS[name])
    cfg['name'] = name
    for env, key in (('KAS_BENCH_KEYS', 'keys'), ('KAS_BENCH_REPLAY_Q', 'fleet_q'), ('KAS_BENCH_REPLAY_T', 'ticks')):
        if os.environ.get(env):
            cfg[key] = int(os.environ[env])
            cfg['overridden'] = True
    if args.metric:
        cfg['metric'] = args.metric
    cfg['title'] = cfg['title'] % cfg
    return cfg


def queue_names(n=N_PREFIX):
    base = ['predict', 'predict:zstack', 'track', 'tracker', 'train', 'segment', 'seg',
            'segment:nuclear', 'segment:cyto', 'spot', 'spots', 'caliban', 'polaris', 'mesmer',
            'mesmer:whole-cell', 'a:b', 'a:b:c', 'abc']
    i = 0
    while len(base) < n:
        base.append('queue-%02d-%s' % (i, 'x' * (i % 9)))
        i += 1
    return base[:n]


def mix64(z):
    """numpy form of the generator's mixer."""
    z = (np.asarray(z, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15))
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def fleet_groups(seed, q_first, n_queues):
    """CSR offsets of a fleet of multi-queue resources over queues [q_first, q_first + n_queues):
    resource sizes 1..8 from the mixer, the last one cut to fit.  Returns local offsets (uint64)."""
    with np.errstate(over='ignore'):
        sizes = (mix64(np.uint64(seed) ^ (np.arange(n_queues, dtype=np.uint64) + np.uint64(q_first))) % np.uint64(8)
                 ).astype(np.int64) + 1
    off = np.concatenate([[0], np.cumsum(sizes)])
    off = off[off <= n_queues]
    if off[-1] != n_queues:
        off = np.append(off, n_queues)
    return off.astype(np.uint64)


def resource_types(seed, r_first, n):
    """'deployment' / 'job' per resource (config 5: mixed types)."""
    with np.errstate(over='ignore'):
        bits = mix64(np.uint64(seed ^ 0x7A) ^ (np.arange(n, dtype=np.uint64) + np.uint64(r_first))) & np.uint64(1)
    return np.where(bits == 1, 'job', 'deployment')


def measured_peak():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        with open(p) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md)'


def ncu_traffic(kernel, alg_bytes):
    """DRAM bytes per launch from the committed ncu capture of this kernel, if the capture was made
    on a launch of this size (profiles/traffic.json); else None."""
    try:
        with open(os.path.join(ROOT, 'profiles', 'traffic.json')) as f:
            e = json.load(f).get(kernel)
        if isinstance(e, dict) and abs(e['algorithmic_bytes'] - alg_bytes) <= 0.001 * alg_bytes:
            return e['bytes']
    except Exception:
        pass
    return None


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    REASONS = {0x1: 'gpu_idle', 0x2: 'applications_clocks_setting', 0x4: 'sw_power_cap',
               0x8: 'hw_slowdown', 0x10: 'sync_boost', 0x20: 'sw_thermal_slowdown',
               0x40: 'hw_thermal_slowdown', 0x80: 'hw_power_brake_slowdown',
               0x100: 'display_clock_setting'}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.stop_flag = threading.Event()
        self.active = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.max_mhz = None

    def run(self):
        if not self.ok:
            return
        while not self.stop_flag.is_set():
            try:
                mhz = self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)
                try:
                    r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.samples.append((self.active.is_set(), mhz, r))
            except Exception:
                pass
            time.sleep(0.003)

    def summary(self):
        if not self.ok or not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': self.max_mhz, 'reasons': [], 'samples': 0}
        inside = [s for s in self.samples if s[0]] or self.samples[-3:]
        mhz = sorted(s[1] for s in inside)
        bits = 0
        for s in inside:
            bits |= s[2]
        reasons = [n for b, n in self.REASONS.items() if bits & b and n != 'gpu_idle']
        return {'sm_mhz': mhz[len(mhz) // 2], 'sm_max_mhz': self.max_mhz, 'reasons': reasons,
                'samples': len(inside)}


# --------------------------------------------------------------------------
# CPU arm (oracle port; no product code)
# --------------------------------------------------------------------------
def cpu_sample_keys(steps, warmup):
    """Keys per CPU step: about 30 s of work for the whole run at ~1.3e7 keys/s."""
    total = 4.0e8
    return int(max(2_000_000, min(16_000_000, total / max(1, steps + warmup))))


def cpu_sample_queues(steps, warmup, ticks):
    total = 2.0e10                              # (queue, tick) decisions for the whole run
    return int(max(2_000, min(100_000, total / max(1, steps + warmup) / max(1, ticks))))


def run_reference(args):
    """The CPU restatement of the reference's path on the host cores (rank 0 only).  Inputs come
    from the oracle's own generators; libkas.so is never loaded here."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    from oracle import coracle
    cfg = resolve_config(args)
    threads = coracle.max_threads()
    names = queue_names()
    if cfg['metric'] == 'keys':
        n_sample = int(os.environ.get('KAS_BENCH_CPU_KEYS', cpu_sample_keys(args.steps, args.warmup)))
        W = coracle.rec_bytes_for(names)
        heads = coracle.synth_heads(SEED, 0, n_sample, names, W, long_tail=cfg['long_tail'], threads=threads)

        def step():
            return coracle.match_counts_heads(names, heads, threads=threads)
        units, unit, metric = n_sample, 'keys/s', 'keys matched/sec'
        sample = ('%d keys x %d queues of the %s keyspace per step (bounded sample of %d), C restatement, '
                  'one full pass per queue like the reference' % (n_sample, N_PREFIX, cfg['name'], cfg['keys']))
        extra_cfg = {'cpu_sample_keys': n_sample}
    else:
        T = cfg['ticks']
        n_q = int(os.environ.get('KAS_BENCH_CPU_QUEUES', cpu_sample_queues(args.steps, args.warmup, T)))
        trace = coracle.trace_fill(SEED + 1, T, 0, n_q, threads=threads)
        go = fleet_groups(SEED + 3, 0, n_q) if cfg['grouped'] else None

        def step():
            return coracle.replay(trace, None, go, None, [DEFAULT_POLICY], threads=threads)
        units, unit, metric = n_q * T, '(queue,tick) decisions/s', 'scaling decisions/sec'
        sample = ('%d queues x %d ticks of the %s trace per step (bounded sample of %d queues), C restatement '
                  'of the scale() loop, default policy' % (n_q, T, cfg['name'], cfg['fleet_q']))
        extra_cfg = {'cpu_sample_queues': n_q}
    for _ in range(args.warmup):
        step()
    secs = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        step()
        secs.append(time.perf_counter() - t0)
    value = len(secs) * units / sum(secs)
    config = bench_config(cfg, args.gpus)
    config.update(extra_cfg)
    config['cpu_sample'] = sample
    line = {
        'impl': 'reference', 'metric': metric, 'value': value, 'unit': unit,
        'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': 1000.0 * sum(secs) / len(secs), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'u8' if cfg['metric'] == 'keys' else 'int32',
        'data': 'synthetic', 'config': config,
        'cpu_baseline': {'value': value, 'unit': unit, 'cores': threads, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': unit, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    emit(line)


def bench_config(cfg, n_gpus):
    out = {'workload': cfg['title'], 'config': cfg['name'], 'keys_per_gpu': cfg['keys'], 'n_prefix': N_PREFIX,
           'rec_bytes': 32, 'long_tail_keys': cfg['long_tail'], 'fleet_queues_per_gpu': cfg['fleet_q'],
           'ticks': cfg['ticks'], 'multi_queue_resources': cfg['grouped'],
           'sharding': ('keys and resources split by rank (contiguous ranges), integer totals all-reduced '
                        'once (NCCL)') if n_gpus > 1 else 'none',
           'l2': 'inputs (%.1f GB of heads, %.0f GB of trace per GPU) larger than the 126 MB L2; no flush needed'
                 % (cfg['keys'] * 32 / 1e9, cfg['fleet_q'] * cfg['ticks'] * 4 / 1e9)}
    if cfg.get('overridden'):
        out['sizes_overridden_by_env'] = True
    return out


# --------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------
class Ctx(object):
    """What every leg needs: torch, the device, the ranks, the communicator."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from autoscaler import engine
        self.torch, self.dist, self.engine = torch, dist, engine
        self.world = int(os.environ.get('WORLD_SIZE', '1'))
        self.rank = int(os.environ.get('RANK', '0'))
        self.local = int(os.environ.get('LOCAL_RANK', '0'))
        if not torch.cuda.is_available():
            raise SystemExit('bench.py needs a CUDA device (there is no CPU fallback); '
                             'use --impl reference for the CPU arm')
        torch.cuda.set_device(self.local)
        self.dev = torch.device('cuda', self.local)
        self.numa = bind_to_gpu_numa_node(self.local) if self.world > 1 else 'unchanged'
        self.comm = None
        if self.world > 1:
            dist.init_process_group('nccl', device_id=self.dev)
            self.comm = engine.Comm.from_torch_group()       # libkas's own NCCL communicator
            # one SM for the all-reduce that runs beside the next step's match kernel
            engine.set_reserved_sms(int(os.environ.get('KAS_BENCH_RESERVED_SMS', 1)))
        self.peak, self.peak_src = measured_peak()
        self.args = args

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        if self.world == 1:
            return float(x)
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def all_true(self, flag):
        if self.world == 1:
            return bool(flag)
        t = self.torch.tensor([1 if flag else 0], dtype=self.torch.int64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return bool(t.item())

    def timed(self, step, steps, warmup, sampler=None):
        """warmup untimed calls of step(i), then exactly `steps` of them between CUDA events on the
        launching stream, bracketed by barrier + synchronize; returns ms per step, max over ranks."""
        torch = self.torch
        for i in range(warmup):
            step(i)
        self.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if sampler is not None:
            sampler.active.set()
        a.record()
        for i in range(steps):
            step(warmup + i)
        b.record()
        torch.cuda.synchronize()
        if sampler is not None:
            sampler.active.clear()
        self.barrier()
        return self.max_over_ranks(a.elapsed_time(b) / steps)

    def roofline(self, kernel, alg_bytes, ms, traffic_key=None):
        gbs = alg_bytes / (ms * 1e-3) / 1e9
        return {'bound': 'hbm', 'kernel': kernel, 'achieved': gbs, 'peak': self.peak, 'unit': 'GB/s',
                'frac': gbs / self.peak, 'peak_source': self.peak_src,
                'algorithmic_bytes_per_launch': int(alg_bytes), 'kernel_ms': ms,
                'traffic': ncu_traffic(traffic_key, alg_bytes) if traffic_key else None}


def bind_to_gpu_numa_node(local):
    """Keep this rank's threads (and the pinned memory they first touch) on the NUMA node that owns
    the GPU's PCIe root: the end-to-end legs are bound by the host-to-device copy."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bdf = pynvml.nvmlDeviceGetPciInfo(h).busId
        bdf = (bdf.decode() if isinstance(bdf, bytes) else bdf).lower()
        if len(bdf.split(':')[0]) == 8:
            bdf = bdf[4:]
        with open('/sys/bus/pci/devices/%s/numa_node' % bdf) as f:
            node = int(f.read().strip())
        nodes = [d for d in os.listdir('/sys/devices/system/node') if d.startswith('node')]
        if node < 0 or len(nodes) < 2:
            return 'one NUMA node (gpu on node %d): unchanged' % node
        with open('/sys/devices/system/node/node%d/cpulist' % node) as f:
            cpus = set()
            for part in f.read().strip().split(','):
                lo, _, hi = part.partition('-')
                cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return 'bound to NUMA node %d of the GPU (%d cpus)' % (node, len(cpus))
        return 'unchanged (no allowed cpu on node %d)' % node
    except Exception as err:   # not fatal: only the e2e legs are affected
        return 'unchanged (%s)' % type(err).__name__


def tally_leg(cx, cfg, headline, sampler):
    """K1 over the rank's resident snapshot, the oracle check of a slice, and the end-to-end forms."""
    torch, engine, dev, world, rank, args = cx.torch, cx.engine, cx.dev, cx.world, cx.rank, cx.args
    from autoscaler import snapshot as ksnap
    from oracle import coracle
    names = queue_names()
    W = engine.rec_bytes_for(names)
    assert W == 32
    n = cfg['keys']
    first = rank * n                       # each rank owns its own contiguous key range
    table = torch.from_numpy(engine.build_prefix_table(names, W)).to(dev)
    out = {}

    t0 = time.perf_counter()
    host_blob = host_off = None
    if cfg['long_tail']:
        # config 5: the snapshot arrives as variable-length keys; the heads are cut on the device
        _, blob, off = engine.synth_keys_host(SEED, first, n, names, long_tail=True, want_heads=False, want_blob=True)
        host_blob = torch.from_numpy(blob).pin_memory()
        host_off = torch.from_numpy(off.view(np.int64)).pin_memory()
        del blob, off
        gen_s = time.perf_counter() - t0
        d_blob = host_blob.to(dev, non_blocking=True)
        d_off = host_off.to(dev, non_blocking=True)
        d_heads = engine.pack_heads_device(d_blob, d_off, n, W).reshape(-1)
        torch.cuda.synchronize()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        for _ in range(5):
            engine.check(engine.lib().kas_pack_heads(engine.ptr(d_blob), engine.ptr(d_off), n, 0, W,
                                                     engine.ptr(d_heads), engine.stream_ptr()), 'kas_pack_heads')
        p1.record()
        torch.cuda.synchronize()
        pack_ms = p0.elapsed_time(p1) / 5
        lens = np.diff(host_off.numpy())
        ing_bytes = int(np.minimum(lens, W).sum()) + 8 * (n + 1) + n * W
        out['ingest'] = {'keys': n, 'mean_key_bytes': float(host_blob.numel()) / n, 'max_key_bytes': int(lens.max()),
                         'device_pack': {'ms': pack_ms, 'keys_per_sec': n / (pack_ms * 1e-3),
                                         'roofline': cx.roofline('pack_heads_kernel', ing_bytes, pack_ms)}}
        del d_blob, d_off
        host_heads = None
    else:
        host_heads = torch.empty(n * W, dtype=torch.uint8).pin_memory()
        engine.synth_keys_host(SEED, first, n, names, rec_bytes=W, heads_out=host_heads)
        gen_s = time.perf_counter() - t0
        d_heads = torch.empty(n * W, dtype=torch.uint8, device=dev)
        d_heads.copy_(host_heads, non_blocking=True)
    torch.cuda.synchronize()

    # ---- the step: K1 (+ the all-reduce on a second stream, double-buffered counts) ----------
    NB = 3                                    # count buffers in rotation
    counts = [torch.empty(N_PREFIX, dtype=torch.int64, device=dev) for _ in range(NB)]
    comm_stream = torch.cuda.Stream(device=dev) if world > 1 else None
    filled = [torch.cuda.Event() for _ in range(NB)]
    reduced = [torch.cuda.Event() for _ in range(NB)]
    main = torch.cuda.current_stream(dev)

    def step(i):
        b = i % NB
        if world > 1:
            main.wait_event(reduced[b])          # the all-reduce of step i-NB is done with this buffer
        engine.match_count(d_heads, n, W, table, N_PREFIX, d_counts=counts[b], accumulate=False)
        if world > 1:
            filled[b].record(main)
            with torch.cuda.stream(comm_stream):
                comm_stream.wait_event(filled[b])
                cx.comm.allreduce_sum_(counts[b])
                reduced[b].record(comm_stream)

    def drain():
        if world > 1:
            main.wait_stream(comm_stream)

    steps, warmup = (args.steps, max(3, args.warmup)) if headline else (max(10, min(args.steps, 50)), 3)

    # the all-reduces still in flight are part of the timed region: drain before the stop event
    def timed_step(i):
        step(i)
        if i == warmup + steps - 1 or i == warmup - 1:
            drain()
    ms = cx.timed(timed_step, steps, warmup, sampler if headline else None)
    last = (warmup + steps - 1) % NB
    counts_box = engine.counts_to_host(counts[last]).astype(np.int64)
    out.update({'ms_per_step': ms, 'value': world * n / (ms * 1e-3), 'steps': steps, 'warmup': warmup,
                'launches_per_step': 1})

    # kernel alone (no collective) for the roofline
    if world == 1:
        k_ms = ms
    else:
        k_ms = cx.timed(lambda i: engine.match_count(d_heads, n, W, table, N_PREFIX, d_counts=counts[0]),
                        max(10, min(args.steps, 100)), 3)
    engine.match_count(d_heads, n, W, table, N_PREFIX, d_counts=counts[0])
    local_counts = engine.counts_to_host(counts[0]).astype(np.int64)
    out['roofline'] = cx.roofline('match_count_kernel<2,true,*> (32-byte records, 256-bit loads)',
                                  n * W + 8 * N_PREFIX, k_ms, 'match_count_kernel')

    # ---- every rank: a slice of its own range against the oracle ---------------------------
    n_chk = min(n, int(os.environ.get('KAS_BENCH_CHECK_KEYS', 2_000_000)))
    thr = max(1, coracle.max_threads() // max(1, world))
    lo = (n - n_chk) // 2
    d_slice = d_heads[lo * W:(lo + n_chk) * W]
    got = engine.counts_to_host(engine.match_count(d_slice, n_chk, W, table, N_PREFIX)).astype(np.int64)
    want_heads = coracle.synth_heads(SEED, first + lo, n_chk, names, W, long_tail=cfg['long_tail'], threads=thr)
    want = coracle.match_counts_heads(names, want_heads, threads=thr).astype(np.int64)
    slice_ok = bool((got == want).all())
    box_ok = True
    if world > 1:       # the all-reduced counts are the sum of the ranks' own counts
        t = torch.from_numpy(local_counts.copy()).to(dev)
        cx.dist.all_reduce(t)
        box_ok = bool((t.cpu().numpy() == counts_box).all())
    ok = cx.all_true(slice_ok and box_ok)
    if not ok:
        raise SystemExit('bench.py: GPU counts differ from the CPU oracle on a sampled slice (rank %d: slice %s, '
                         'box sum %s) -- a throughput number for wrong results is not a result'
                         % (rank, slice_ok, box_ok))
    out['checks'] = {'oracle_slice_equal_all_ranks': ok, 'slice_keys_per_rank': n_chk,
                     'keys_matched_fraction': float(counts_box.sum()) / (world * n),
                     'local_counts_sum': int(local_counts.sum()), 'host_generation_s': gen_s}

    # ---- end to end: host snapshot -> H2D -> K1 -> counts on the host ----------------------
    e2e_steps = max(2, min(args.steps, 4))
    chunk_keys = int(os.environ.get('KAS_BENCH_CHUNK_KEYS', 8_388_608))
    streamer = ksnap.StreamingTally(names, chunk_keys=chunk_keys, device=dev)
    reduce = (lambda t: cx.comm.allreduce_sum_(t)) if world > 1 else None

    def wall_timed(fn, reps):
        fn()
        cx.barrier()
        w0 = time.perf_counter()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(reps):
            fn()
        g1.record()
        torch.cuda.synchronize()
        return cx.max_over_ranks(max(g0.elapsed_time(g1) / reps, (time.perf_counter() - w0) * 1e3 / reps))

    if host_heads is not None:
        e_ms = wall_timed(lambda: streamer.enqueue(host_heads, n, reduce=reduce), e2e_steps)
        if not cx.all_true(bool((streamer.host_counts.numpy() == counts_box).all())):
            raise SystemExit('bench.py: end-to-end counts differ from the resident-data counts')
        h2d = n * W
        # the bare copy, same chunks, same buffers, no kernel: the ceiling the pipeline can reach
        def bare():
            for lo_k in range(0, n, chunk_keys):
                hi_k = min(n, lo_k + chunk_keys)
                streamer.stage[(lo_k // chunk_keys) & 1][:(hi_k - lo_k) * W].copy_(
                    host_heads[lo_k * W:hi_k * W], non_blocking=True)
        c_ms = wall_timed(bare, e2e_steps)
        out['e2e'] = {'value': world * n / (e_ms * 1e-3), 'unit': 'keys/s', 'h2d_bytes_per_step': h2d,
                      'd2h_bytes_per_step': 8 * N_PREFIX, 'ms_per_step': e_ms, 'steps': e2e_steps,
                      'matches_resident_result': True,
                      'h2d_gbs_per_rank': h2d / (e_ms * 1e-3) / 1e9,
                      'h2d_ceiling_gbs_per_rank': h2d / (c_ms * 1e-3) / 1e9,
                      'h2d_frac': c_ms / e_ms,
                      'bound': 'PCIe host-to-device copy (h2d_frac = bare pinned copy time / pipeline time, '
                               'all ranks copying at once)',
                      'path': 'autoscaler.snapshot.StreamingTally.enqueue: pinned host heads -> chunked '
                              'double-buffered H2D -> kas_match_count_accumulate -> counts D2H'}
    # the raw form: variable-length keys (blob + offsets) in pinned host memory, heads cut on the device
    raw_n = n if host_blob is not None else min(n, int(os.environ.get('KAS_BENCH_RAW_KEYS', 32_000_000)))
    if host_blob is None:
        _, blob, off = engine.synth_keys_host(SEED, first, raw_n, names, want_heads=False, want_blob=True)
        host_blob = torch.from_numpy(blob).pin_memory()
        host_off = torch.from_numpy(off.view(np.int64)).pin_memory()
        del blob, off
    blob_chunk = 512 << 20
    chunks = streamer.plan_blob_chunks(host_off.numpy().view(np.uint64), blob_chunk)
    r_ms = wall_timed(lambda: streamer.enqueue_blob(host_blob, host_off, chunks, blob_chunk, reduce=None), e2e_steps)
    raw_counts = streamer.host_counts.numpy().astype(np.int64).copy()
    want_raw = engine.counts_to_host(engine.match_count(d_heads, raw_n, W, table, N_PREFIX)).astype(np.int64)
    if not cx.all_true(bool((raw_counts == want_raw).all())):
        raise SystemExit('bench.py: raw-snapshot end-to-end counts differ from the resident-data counts')
    raw_bytes = int(host_blob.numel()) + 8 * (raw_n + 1)
    e2e_raw = {'value': world * raw_n / (r_ms * 1e-3), 'unit': 'keys/s', 'keys_per_rank': raw_n,
               'h2d_bytes_per_step': raw_bytes, 'd2h_bytes_per_step': 8 * N_PREFIX, 'ms_per_step': r_ms,
               'h2d_gbs_per_rank': raw_bytes / (r_ms * 1e-3) / 1e9, 'matches_resident_result': True,
               'path': 'StreamingTally.enqueue_blob: pinned host blob + uint64 offsets -> chunked H2D -> '
                       'kas_pack_heads -> kas_match_count_accumulate -> counts D2H (nothing packed on the host)'}
    if 'e2e' in out:
        out['e2e_raw'] = e2e_raw
    else:
        e2e_raw['launches'] = 2 * len(chunks)
        out['e2e'] = e2e_raw
    out['n'] = n
    out['names'] = names
    out['table'] = table
    out['d_heads'] = d_heads
    del streamer, host_blob, host_off
    return out


def replay_check_samples(cx, cfg, seed, q_first, Q, T, go, runs, n_samples=16):
    """Rebuild sampled resources' whole trajectories with the oracle's generator twin and replay
    them on the CPU; compare with the rows of every run (name, policies, totals on the host)."""
    from oracle import coracle
    rng = np.random.default_rng(1234 + cx.rank)
    R = Q if go is None else len(go) - 1
    picks = sorted(set(rng.integers(0, R, size=n_samples).tolist() + [0, R - 1]))
    thr = max(1, coracle.max_threads() // max(1, cx.world))
    ok = True
    for r in picks:
        q0, q1 = (r, r + 1) if go is None else (int(go[r]), int(go[r + 1]))
        cols = coracle.trace_fill(seed, T, q_first + q0, q1 - q0, threads=thr)
        local_go = None if go is None else np.array([0, q1 - q0], dtype=np.uint64)
        for name, pols, tot in runs:
            want = coracle.replay(cols, None, local_go, None, pols, threads=thr)
            ok = ok and tot['pod_ticks'][r].tolist() == want['pod_ticks'][0].tolist() \
                and tot['n_patches'][r].tolist() == want['n_patches'][0].tolist() \
                and tot['final_current'][r].tolist() == want['final'][0].tolist()
            if not ok:
                raise SystemExit('bench.py: replay `%s` differs from the CPU oracle on resource %d (rank %d)'
                                 % (name, r, cx.rank))
    return cx.all_true(ok), len(picks)


def replay_leg(cx, cfg, headline, sampler):
    """K2 over the rank's resident trace: default policy, sweep, uint16 trace; sampled-resource
    oracle check; CPU baseline; end to end from a host trace."""
    torch, engine, dev, world, rank, args = cx.torch, cx.engine, cx.dev, cx.world, cx.rank, cx.args
    from autoscaler import replay as kreplay
    from oracle import coracle
    Q, T = cfg['fleet_q'], cfg['ticks']
    seed = SEED + 1
    q_first = rank * Q
    trace = engine.synth_trace(seed, T, q_first, Q, device=dev)
    go = d_go = None
    R = Q
    if cfg['grouped']:
        go = fleet_groups(SEED + 3, q_first, Q)
        R = len(go) - 1
        d_go = torch.from_numpy(go.view(np.int64)).to(dev)
    torch.cuda.synchronize()
    reduce = (lambda t: kreplay.allreduce_sum_(t, comm=cx.comm))

    def run(tr, policies, want_totals):
        out = engine.replay(tr, T, Q, policies, d_group_off=d_go, n_resources=R, want_totals=want_totals)
        reduce(out['sums'])
        return out

    def timed(tr, policies, reps, warm, smp=None):
        ms = cx.timed(lambda i: run(tr, policies, False), reps, warm, smp)
        return ms, world * Q * T * len(policies) / (ms * 1e-3)

    if headline:
        steps, warmup = args.steps, max(3, args.warmup)
    else:
        steps, warmup = 10, 3
    ms1, dps1 = timed(trace, [DEFAULT_POLICY], steps, warmup, sampler if headline else None)
    msS, dpsS = timed(trace, SWEEP, 5, 2)
    runs = []
    o1 = run(trace, [DEFAULT_POLICY], True)
    runs.append(('default policy, int32', [DEFAULT_POLICY], engine.totals_to_host(o1['totals'])))
    sums1 = o1['sums'].cpu().numpy()
    oS = run(trace, SWEEP, True)
    runs.append(('sweep, int32', SWEEP, engine.totals_to_host(oS['totals'])))
    sumsS = oS['sums'].cpu().numpy()
    del o1, oS
    kernel1 = 'replay_grouped_sorted_kernel' if cfg['grouped'] else 'replay_identity_kernel'
    rep = {'metric': 'scaling decisions/sec', 'unit': '(queue,tick,policy) decisions/s',
           'workload': '%s: %d queues in %d %s per GPU x %d ticks, int32 tick-major trace resident in HBM '
                       '(%.0f GB per GPU)' % (cfg['name'], Q, R, 'multi-queue resources (1-8 queues)' if cfg['grouped']
                                              else 'single-queue deployments', T, Q * T * 4 / 1e9),
           'ms_per_step': ms1, 'value': dps1, 'steps': steps, 'warmup': warmup,
           's1': {'policies': 1, 'ms': ms1, 'decisions_per_sec': dps1,
                  'roofline': cx.roofline(kernel1, Q * T * 4, ms1, kernel1)},
           'sweep': {'policies': len(SWEEP), 'ms': msS, 'decisions_per_sec': dpsS, 'reps': 5, 'warmup': 2,
                     'bound': 'integer issue (41 decisions per 4 loaded bytes)'},
           'pod_ticks_default_policy': int(sums1[0, 0]),
           'pod_hours_default_policy_interval_5s': int(sums1[0, 0]) * 5 / 3600.0}
    if cfg['grouped']:
        ms4, dps4 = timed(trace, FOUR, 5, 2)
        rep['four_policies'] = {'policies': 4, 'ms': ms4, 'decisions_per_sec': dps4}

    # patch log with the body field of each resource's type (config 5: mixed job / deployment),
    # over the first ticks of the trace (the log of all 1e4 ticks is tens of GB)
    if cfg['grouped']:
        Tp = min(T, int(os.environ.get('KAS_BENCH_PATCHLOG_TICKS', 1000)))
        types = resource_types(SEED, 0, R)
        head = trace[:Tp]
        t0 = time.perf_counter()
        off, ev = engine.patch_events(head, Tp, Q, DEFAULT_POLICY, d_group_off=d_go, n_resources=R)
        log = kreplay.PatchLog(off, ev, types)
        ev_s = time.perf_counter() - t0
        n_pat = int(engine.replay(head, Tp, Q, [DEFAULT_POLICY], d_group_off=d_go, n_resources=R,
                                  want_totals=False)['sums'][0, 1].item())
        rng = np.random.default_rng(99 + rank)
        fields = set()
        for r in rng.integers(0, R, size=8).tolist() + [int(np.argmax(types == 'job')), int(np.argmax(types == 'deployment'))]:
            q0, q1 = int(go[r]), int(go[r + 1])
            cols = coracle.trace_fill(seed, Tp, q_first + q0, q1 - q0, threads=2)
            dense = coracle.replay(cols, None, np.array([0, q1 - q0], dtype=np.uint64), None, [DEFAULT_POLICY],
                                   want_dense=True, threads=2)['dense'][0, :, 0]
            prev, want = 0, []
            field = kreplay.RESOURCE_FIELD[types[r]]           # autoscaler.py:232 (job), :236 (deployment)
            for t, v in enumerate(dense.tolist()):
                if v != prev:
                    want.append((t, {'spec': {field: v}}))
                prev = v
            if log.bodies(r) != want:
                raise SystemExit('bench.py: patch bodies of resource %d differ from the oracle' % r)
            fields.add(field)
        if len(log) != n_pat:
            raise SystemExit('bench.py: patch log length differs from the replay patch count')
        rep['patch_log'] = {'ticks': Tp, 'events': len(log), 'seconds_incl_alloc_and_d2h': ev_s,
                            'bodies_equal_oracle_on_samples': True, 'body_fields_seen': sorted(fields),
                            'job_fraction': float((types == 'job').mean())}
        del off, ev, log, head

    # the same trace stored as uint16 (half the bytes); the synthetic depths are below 2^16
    if int(trace.max().item()) < 65536:
        trace16 = trace.to(torch.int16)      # same bits as uint16 below 2^15; the library reads uint16
        del trace
        trace = trace16
        ms16, dps16 = timed(trace, [DEFAULT_POLICY], 10, 3)
        msS16, dpsS16 = timed(trace, SWEEP, 5, 2)
        o16 = run(trace, [DEFAULT_POLICY], True)
        runs.append(('default policy, uint16', [DEFAULT_POLICY], engine.totals_to_host(o16['totals'])))
        oS16 = run(trace, SWEEP, True)
        runs.append(('sweep, uint16', SWEEP, engine.totals_to_host(oS16['totals'])))
        if not bool((o16['sums'].cpu().numpy() == sums1).all()) or not bool((oS16['sums'].cpu().numpy() == sumsS).all()):
            raise SystemExit('bench.py: uint16 and int32 traces replay to different integers')
        del o16, oS16
        k16 = 'replay_grouped_sorted_kernel' if cfg['grouped'] else 'replay_packed16_kernel'
        rep['s1_uint16'] = {'policies': 1, 'ms': ms16, 'decisions_per_sec': dps16,
                            'roofline': cx.roofline(k16, Q * T * 2, ms16, k16)}
        rep['sweep_uint16'] = {'policies': len(SWEEP), 'ms': msS16, 'decisions_per_sec': dpsS16, 'reps': 5, 'warmup': 2}
    del trace

    ok, n_picks = replay_check_samples(cx, cfg, seed, q_first, Q, T, go, runs)
    rep['oracle_sample_equal'] = ok
    rep['oracle_samples_per_rank'] = n_picks
    rep['oracle_checked_runs'] = [r[0] for r in runs]
    # box-wide sums are the sums of the ranks' rows (the all-reduce added the right things)
    own = np.stack([runs[0][2]['pod_ticks'][:, 0].sum(), runs[0][2]['n_patches'][:, 0].astype(np.int64).sum(),
                    runs[0][2]['final_current'][:, 0].astype(np.int64).sum()])
    t = torch.from_numpy(own.astype(np.int64)).to(dev)
    if world > 1:
        cx.dist.all_reduce(t)
    rep['allreduced_sums_equal_sum_of_rows'] = cx.all_true(bool((t.cpu().numpy() == sums1[0]).all()))
    if not rep['allreduced_sums_equal_sum_of_rows']:
        raise SystemExit('bench.py: all-reduced sums differ from the sum of per-resource totals')
    del runs

    # ---- end to end: pinned host trace -> H2D -> replay -> sums on the host ------------------
    T_e = max(1, min(T, int(os.environ.get('KAS_BENCH_E2E_TRACE_BYTES', 4 << 30)) // (Q * 4)))
    dtrace = engine.synth_trace(seed, T_e, q_first, Q, device=dev)
    htrace = torch.empty((T_e, Q), dtype=torch.int32).pin_memory()
    htrace.copy_(dtrace)
    want_sums = engine.replay(dtrace, T_e, Q, [DEFAULT_POLICY], d_group_off=d_go, n_resources=R,
                              want_totals=False)['sums'].cpu().numpy()
    del dtrace
    got = {}

    def e2e_once():
        res = kreplay.replay_fleet(htrace, [DEFAULT_POLICY], group_off=go, want_totals=False)
        got['sums'] = res.sums
    e2e_once()
    cx.barrier()
    w0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        e2e_once()
    e_ms = cx.max_over_ranks((time.perf_counter() - w0) * 1e3 / reps)
    if not cx.all_true(bool((got['sums'] == want_sums).all())):
        raise SystemExit('bench.py: end-to-end replay sums differ from the resident-trace sums')
    rep['e2e'] = {'value': world * Q * T_e / (e_ms * 1e-3), 'unit': '(queue,tick) decisions/s',
                  'h2d_bytes_per_step': Q * T_e * 4, 'd2h_bytes_per_step': 24, 'ms_per_step': e_ms, 'steps': reps,
                  'h2d_gbs_per_rank': Q * T_e * 4 / (e_ms * 1e-3) / 1e9, 'matches_resident_result': True,
                  'sample': 'the first %d ticks of the fleet (a %d-byte pinned host trace per rank): the step is '
                            'the H2D copy, so the rate does not depend on the tick count' % (T_e, Q * T_e * 4),
                  'path': 'autoscaler.replay.replay_fleet(pinned int32 host trace) -> H2D -> kas_replay -> sums D2H'}
    # the same from a uint16 host trace (half the bytes over PCIe)
    if int(htrace.max().item()) < 32768:
        h16 = htrace.to(torch.int16).pin_memory()

        def e2e16_once():
            got['sums16'] = kreplay.replay_fleet(h16, [DEFAULT_POLICY], group_off=go, want_totals=False).sums
        e2e16_once()
        cx.barrier()
        w0 = time.perf_counter()
        for _ in range(reps):
            e2e16_once()
        e16_ms = cx.max_over_ranks((time.perf_counter() - w0) * 1e3 / reps)
        if not cx.all_true(bool((got['sums16'] == want_sums).all())):
            raise SystemExit('bench.py: end-to-end uint16 replay sums differ from the resident-trace sums')
        rep['e2e_uint16'] = {'value': world * Q * T_e / (e16_ms * 1e-3), 'unit': '(queue,tick) decisions/s',
                             'h2d_bytes_per_step': Q * T_e * 2, 'd2h_bytes_per_step': 24, 'ms_per_step': e16_ms,
                             'h2d_gbs_per_rank': Q * T_e * 2 / (e16_ms * 1e-3) / 1e9, 'matches_resident_result': True}
        del h16
    del htrace

    # ---- CPU baseline: the oracle's scale() loop on a slice of the same trace ----------------
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = coracle.max_threads()
        n_q = int(os.environ.get('KAS_BENCH_CPU_QUEUES', min(Q, 50_000)))
        cols = coracle.trace_fill(seed, T, 0, n_q, threads=threads)
        cgo = fleet_groups(SEED + 3, 0, n_q) if cfg['grouped'] else None
        t0 = time.perf_counter()
        coracle.replay(cols, None, cgo, None, [DEFAULT_POLICY], threads=threads)
        c_s = time.perf_counter() - t0
        n_qs = max(1, n_q // 10)
        t0 = time.perf_counter()
        coracle.replay(cols[:, :n_qs].copy(), None, fleet_groups(SEED + 3, 0, n_qs) if cfg['grouped'] else None, None,
                       SWEEP, threads=threads)
        cs_s = time.perf_counter() - t0
        from oracle import oracle
        n_py = 4
        t0 = time.perf_counter()
        for j in range(n_py):
            oracle.replay_resource([[int(v)] for v in cols[:, j]], None, DEFAULT_POLICY[0], DEFAULT_POLICY[1],
                                   DEFAULT_POLICY[2], 0)
        p_s = time.perf_counter() - t0
        rep['cpu_baseline'] = {
            'value': n_q * T / c_s, 'unit': '(queue,tick) decisions/s', 'cores': threads, 'kind': 'port',
            'sample': '%d queues x %d ticks of the same trace, C restatement of the scale() loop '
                      '(autoscaler.py:244-273), default policy; %.1f s' % (n_q, T, c_s),
            'sweep': {'value': n_qs * T * len(SWEEP) / cs_s, 'unit': '(queue,tick,policy) decisions/s',
                      'sample': '%d queues x %d ticks x %d policies; %.1f s' % (n_qs, T, len(SWEEP), cs_s)},
            'python_port': {'value': n_py * T / p_s, 'unit': '(queue,tick) decisions/s', 'cores': 1,
                            'sample': '%d queues x %d ticks, line-by-line Python port of scale(); %.1f s'
                                      % (n_py, T, p_s)}}
    return rep


def extras_leg(cx, cfg, tally):
    """Numbers for the kernels the headline does not touch (N = 1): other record widths and table
    forms of K1, the glob kernel, the patch-event kernels, multi-queue and small fleets."""
    torch, engine, dev = cx.torch, cx.engine, cx.dev
    from autoscaler import replay as kreplay
    from oracle import coracle
    out = {}

    def time_ms(fn, reps=10, warm=3):
        return cx.timed(lambda i: fn(), reps, warm)

    # K1 at other record widths / table forms
    n = int(os.environ.get('KAS_BENCH_EXTRA_KEYS', 50_000_000))
    k1 = {}
    cases = {
        'w16_8names': ['q%d' % i for i in range(8)],
        'w48_64names': ['queue-name-number-%02d-%s' % (i, 'y' * (i % 12)) for i in range(64)],
        'w64_64names': ['a-rather-long-queue-name-for-wide-records-%02d-%s' % (i, 'z' * (i % 8)) for i in range(64)],
        'w32_1024names_two_slot_table': ['q%04d%s' % (i, 'abcdefghij'[:i % 11]) for i in range(1024)],
        'w32_2names_scale_py_default': ['predict', 'track'],
    }
    for label, names in cases.items():
        W = engine.rec_bytes_for(names)
        heads = torch.empty(n * W, dtype=torch.uint8).pin_memory()
        engine.synth_keys_host(SEED + 11, 0, n, names, rec_bytes=W, heads_out=heads)
        d = heads.to(dev)
        tbl = torch.from_numpy(engine.build_prefix_table(names, W)).to(dev)
        cnt = torch.empty(len(names), dtype=torch.int64, device=dev)
        ms = time_ms(lambda: engine.match_count(d, n, W, tbl, len(names), d_counts=cnt))
        n_chk = 200_000
        got = engine.counts_to_host(engine.match_count(d[:n_chk * W], n_chk, W, tbl, len(names))).astype(np.int64)
        want = coracle.match_counts_heads(names, heads[:n_chk * W].numpy().reshape(n_chk, W)).astype(np.int64)
        if not (got == want).all():
            raise SystemExit('bench.py: K1 case %s differs from the oracle' % label)
        k1[label] = {'keys': n, 'rec_bytes': W, 'names': len(names), 'ms': ms, 'keys_per_sec': n / (ms * 1e-3),
                     'oracle_equal_on_%d_keys' % n_chk: True,
                     'roofline': cx.roofline('match_count_kernel', n * W + 8 * len(names), ms)}
        del heads, d, tbl, cnt
    out['k1_variants'] = k1

    # general glob kernel: whole keys, live-glob queue names
    n_g = int(os.environ.get('KAS_BENCH_GLOB_KEYS', 16_000_000))
    names = queue_names()
    _, blob, off = engine.synth_keys_host(SEED + 5, 0, n_g, names, want_heads=False, want_blob=True)
    d_blob = torch.from_numpy(blob).to(dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    pats = [b'processing-pre*:*', b'processing-[st]*:*', b'processing-queue-0?-*:*', b'processing-track:*']
    ms = time_ms(lambda: engine.glob_count(d_blob, d_off, n_g, pats), reps=5, warm=2)
    n_chk = 20_000
    got = engine.counts_to_host(engine.glob_count(d_blob, d_off, n_chk, pats)).astype(np.int64)
    want = [sum(coracle.glob_match(p, bytes(blob[off[i]:off[i + 1]])) for i in range(n_chk)) for p in pats]
    if got.tolist() != want:
        raise SystemExit('bench.py: glob kernel differs from the oracle')
    g_bytes = int(blob.size) + 8 * (n_g + 1)
    out['glob'] = {'keys': n_g, 'patterns': len(pats), 'ms': ms, 'key_pattern_tests_per_sec': n_g * len(pats) / (ms * 1e-3),
                   'oracle_equal_on_%d_keys' % n_chk: True,
                   'roofline': cx.roofline('glob_count_kernel (keys read once for all patterns)', g_bytes, ms)}
    del d_blob, d_off, blob, off

    # patch-event kernels (count + scan + emit) and multi-queue / small fleets
    Tg = int(os.environ.get('KAS_BENCH_EXTRA_TICKS', 2000))
    Qg = int(os.environ.get('KAS_BENCH_EXTRA_QUEUES', 1_000_000))
    trace = engine.synth_trace(3, Tg, 0, Qg, device=dev)
    pol = (5, 0, 16)

    def events_once():
        return engine.patch_events(trace, Tg, Qg, pol)
    t0 = time.perf_counter()
    off_e, ev = events_once()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    off_e, ev = events_once()
    ev_s = time.perf_counter() - t0
    col = coracle.trace_fill(3, Tg, 777, 1)
    dense = coracle.replay(col, None, None, None, [pol], want_dense=True)['dense'][0, :, 0]
    mine = ev[int(off_e[777]):int(off_e[778])]
    prev, want = 0, []
    for t, v in enumerate(dense.tolist()):
        if v != prev:
            want.append((t, prev, v))
        prev = v
    if [(int(e['tick']), int(e['from_pods']), int(e['to_pods'])) for e in mine] != want:
        raise SystemExit('bench.py: patch events differ from the oracle')
    out['patch_events'] = {'queues': Qg, 'ticks': Tg, 'events': int(len(ev)), 'seconds_incl_alloc_and_d2h': ev_s,
                           'oracle_equal_on_sample': True,
                           'note': 'two passes over the trace (count, emit) + scan; the wall time includes the '
                                   'event buffer allocation and its copy to the host'}
    del off_e, ev

    go = fleet_groups(5, 0, Qg)
    R = len(go) - 1
    d_go = torch.from_numpy(go.view(np.int64)).to(dev)
    grouped = {}
    for dt in ('int32', 'uint16'):
        tr = trace if dt == 'int32' else trace.to(torch.int16)
        esz = 4 if dt == 'int32' else 2
        ms1 = time_ms(lambda: engine.replay(tr, Tg, Qg, [DEFAULT_POLICY], d_group_off=d_go, n_resources=R, want_totals=False))
        ms4 = time_ms(lambda: engine.replay(tr, Tg, Qg, FOUR, d_group_off=d_go, n_resources=R, want_totals=False), reps=5)
        o = engine.replay(tr, Tg, Qg, FOUR, d_group_off=d_go, n_resources=R)
        tot = engine.totals_to_host(o['totals'])
        for r in (0, R // 3, R - 1):
            q0, q1 = int(go[r]), int(go[r + 1])
            w = coracle.replay(coracle.trace_fill(3, Tg, q0, q1 - q0), None, np.array([0, q1 - q0], dtype=np.uint64),
                               None, FOUR)
            if tot['pod_ticks'][r].tolist() != w['pod_ticks'][0].tolist() or \
                    tot['n_patches'][r].tolist() != w['n_patches'][0].tolist():
                raise SystemExit('bench.py: multi-queue replay differs from the oracle')
        grouped[dt] = {'s1_ms': ms1, 's1_queue_ticks_per_sec': Qg * Tg / (ms1 * 1e-3),
                       'four_policies_ms': ms4, 'four_policies_decisions_per_sec': 4 * Qg * Tg / (ms4 * 1e-3),
                       'oracle_equal_on_samples': True,
                       'roofline': cx.roofline('replay_grouped_sorted_kernel', Qg * Tg * esz, ms1, 'replay_grouped_sorted_kernel')}
        del o, tot
    out['multi_queue_fleet'] = {'queues': Qg, 'resources': R, 'ticks': Tg, 'group_sizes': '1-8', **grouped}
    # the `only_running` reading under a pod start-up delay of three ticks (autoscaler.py:176-177): the
    # plain kernel with a window of the last specs per thread, element-wise loads
    delay = {}
    for label, dgo, dR in (('single_queue', None, Qg), ('multi_queue', d_go, R)):
        msd = time_ms(lambda: engine.replay(trace, Tg, Qg, [DEFAULT_POLICY], d_group_off=dgo, n_resources=dR,
                                            want_totals=False, startup_delay=3), reps=5)
        od = engine.replay(trace, Tg, Qg, [DEFAULT_POLICY], d_group_off=dgo, n_resources=dR, startup_delay=3)
        td = engine.totals_to_host(od['totals'])
        for r in (1, dR // 2):
            q0, q1 = (r, r + 1) if dgo is None else (int(go[r]), int(go[r + 1]))
            w = coracle.replay(coracle.trace_fill(3, Tg, q0, q1 - q0), None,
                               None if dgo is None else np.array([0, q1 - q0], dtype=np.uint64), None,
                               [DEFAULT_POLICY], startup_delay=3)
            if td['pod_ticks'][r].tolist() != w['pod_ticks'][0].tolist() or \
                    td['n_patches'][r].tolist() != w['n_patches'][0].tolist():
                raise SystemExit('bench.py: start-up-delay replay differs from the oracle')
        delay[label] = {'ms': msd, 'queue_ticks_per_sec': Qg * Tg / (msd * 1e-3), 'oracle_equal_on_samples': True,
                        'roofline': cx.roofline('replay_grouped_kernel<DELAY>', Qg * Tg * 4, msd)}
        del od, td
    out['startup_delay_3_ticks'] = {'queues': Qg, 'ticks': Tg, **delay}
    del trace, d_go

    # the reference's own case: ONE resource, QUEUES='predict,track', 10k ticks, the 41-policy sweep
    rng = np.random.default_rng(0)
    c1 = (rng.integers(0, 40, size=(10_000, 2)) * (rng.random((10_000, 2)) < 0.6)).astype(np.int32)
    d_c1 = torch.from_numpy(c1).to(dev)
    d_go1 = torch.tensor([0, 2], dtype=torch.int64, device=dev)
    ms = time_ms(lambda: engine.replay(d_c1, 10_000, 2, SWEEP, d_group_off=d_go1, n_resources=1, want_totals=True), reps=5)
    o = engine.replay(d_c1, 10_000, 2, SWEEP, d_group_off=d_go1, n_resources=1)
    w = coracle.replay(c1, None, np.array([0, 2], dtype=np.uint64), None, SWEEP)
    if engine.totals_to_host(o['totals'])['pod_ticks'][0].tolist() != w['pod_ticks'][0].tolist():
        raise SystemExit('bench.py: small-fleet replay differs from the oracle')
    out['one_resource_c1'] = {'queues': 2, 'ticks': 10_000, 'policies': len(SWEEP), 'ms': ms,
                              'decisions_per_sec': 2 * 10_000 * len(SWEEP) / (ms * 1e-3), 'oracle_equal': True}
    return out


def run_ours(args):
    cfg = resolve_config(args)
    cx = Ctx(args)
    torch, engine = cx.torch, cx.engine
    sampler = ClockSampler(cx.local)
    sampler.start()
    keys_headline = cfg['metric'] == 'keys'

    tally = tally_leg(cx, cfg, keys_headline, sampler)
    cpu = None
    if cx.rank == 0 and cx.world == 1 and not args.no_cpu:
        cpu = tally_cpu_baseline(cx, cfg, tally)
    extras = None
    if cx.world == 1 and not args.no_extras:
        d_heads = tally.pop('d_heads')
        del d_heads
        torch.cuda.empty_cache()
        extras = extras_leg(cx, cfg, tally)
    tally.pop('d_heads', None)
    tally.pop('table', None)
    tally.pop('names', None)
    torch.cuda.empty_cache()
    rep = None
    if not args.no_replay or not keys_headline:
        rep = replay_leg(cx, cfg, not keys_headline, sampler)

    sampler.stop_flag.set()
    sampler.join(timeout=2)
    clocks = sampler.summary()

    if cx.rank == 0:
        head = tally if keys_headline else rep
        line = {
            'metric': 'keys matched/sec' if keys_headline else 'scaling decisions/sec',
            'value': head['value'], 'unit': 'keys/s' if keys_headline else '(queue,tick) decisions/s',
            'n_gpus': cx.world, 'steps': head['steps'], 'warmup': head['warmup'],
            'ms_per_step': head['ms_per_step'], 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'u8' if keys_headline else 'int32', 'data': 'synthetic',
            'config': bench_config(cfg, cx.world),
            'roofline': tally['roofline'] if keys_headline else rep['s1']['roofline'],
            'cpu_baseline': cpu if keys_headline else (rep or {}).get('cpu_baseline'),
            'e2e': tally['e2e'] if keys_headline else rep['e2e'],
            'gpu_launches': head['steps'], 'clocks': clocks,
            'collective': ('one ncclAllReduce(int64 sum) per step, issued by libkas (kas_allreduce_i64) on a second '
                           'stream, overlapped with the next step (the match kernel leaves %s SM free for it); '
                           'drained inside the timed region' % os.environ.get('KAS_BENCH_RESERVED_SMS', 1))
                          if cx.world > 1 else 'none',
            'checks': {'host_affinity': cx.numa},
        }
        line['checks'].update(tally['checks'])
        if keys_headline:
            line['replay'] = rep
            if 'e2e_raw' in tally:
                line['e2e_raw'] = tally['e2e_raw']
        else:
            line['tally'] = {k: tally[k] for k in ('value', 'ms_per_step', 'roofline', 'e2e') if k in tally}
            line['tally']['cpu_baseline'] = cpu
            if rep is not None:
                for k in ('oracle_sample_equal', 'allreduced_sums_equal_sum_of_rows'):
                    line['checks']['replay_' + k] = rep[k]
        if rep is not None and keys_headline:
            line['checks']['replay_oracle_sample_equal'] = rep['oracle_sample_equal']
        if 'ingest' in tally:
            line['ingest'] = tally['ingest']
        if extras is not None:
            line['extras'] = extras
        emit(line)
    if cx.world > 1:
        cx.barrier()
        cx.comm.destroy()
        cx.dist.destroy_process_group()


def tally_cpu_baseline(cx, cfg, tally):
    """The oracle's C port on a bounded sample of the same keyspace (all host threads), the same
    sample through the GPU path, and the line-by-line Python port on one core."""
    from oracle import coracle, oracle
    engine = cx.engine
    names, W = tally['names'], 32
    threads = coracle.max_threads()
    n_sample = min(cfg['keys'], int(os.environ.get('KAS_BENCH_CPU_KEYS', 16_000_000)))
    heads = coracle.synth_heads(SEED, 0, n_sample, names, W, long_tail=cfg['long_tail'], threads=threads)
    coracle.match_counts_heads(names, heads[:1000], threads=threads)
    t0 = time.perf_counter()
    ccounts = coracle.match_counts_heads(names, heads, threads=threads)
    dt = time.perf_counter() - t0
    c = engine.match_count(tally['d_heads'], n_sample, W, tally['table'], N_PREFIX)
    agree = bool((engine.counts_to_host(c).astype(np.int64) == ccounts.astype(np.int64)).all())
    if not agree:
        raise SystemExit('bench.py: GPU counts differ from the CPU oracle on the sample -- '
                         'a throughput number for wrong results is not a result')
    n_py = 20_000
    blob, off = coracle.synth_blob(SEED, 0, n_py, names, long_tail=cfg['long_tail'])
    keys = [bytes(blob[off[i]:off[i + 1]]) for i in range(n_py)]
    t0 = time.perf_counter()
    oracle.match_counts(names, keys)
    pdt = time.perf_counter() - t0
    return {'value': n_sample / dt, 'unit': 'keys/s', 'cores': threads, 'kind': 'port',
            'sample': '%d keys x %d queues of the same keyspace, C restatement, one full pass '
                      'per queue like the reference; %.1f s' % (n_sample, N_PREFIX, dt),
            'gpu_counts_equal_on_sample': agree,
            'python_port': {'value': n_py / pdt, 'unit': 'keys/s', 'cores': 1,
                            'sample': '%d keys x %d queues, line-by-line Python port; %.1f s'
                                      % (n_py, N_PREFIX, pdt)}}


_REAL_STDOUT = None


def quiet_stdout():
    """Everything but the result line goes to stderr (NCCL prints its version on
    stdout at the first collective); the JSON line is written to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + '\n').encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--config', default='c2', choices=sorted(CONFIGS))
    ap.add_argument('--metric', default=None, choices=['keys', 'decisions'])
    ap.add_argument('--no-replay', action='store_true')
    ap.add_argument('--no-cpu', action='store_true')
    ap.add_argument('--no-extras', action='store_true')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
