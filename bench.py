#!/usr/bin/env python
"""Benchmark of the tally/decision hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One JSON line on stdout (rank 0).  A "step" is one pass of K1 (prefix match and
count) over the rank's resident key snapshot: BASELINE.json config 2, 1e8
synthetic keys x 64 queue prefixes, 32-byte heads, per GPU (weak scaling); with
N > 1 the step ends with the NCCL all-reduce of the uint64[64] counts.  The
same line carries the replay kernel's numbers (BASELINE config 3, 1M queues x
10k ticks per GPU) under "replay".

`--impl reference` times the CPU restatement of the reference's path (the
reference itself is Python + network I/O and cannot travel to the box; the
oracle's C port with every host thread is the strongest CPU form of it).
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

N_KEYS_PER_GPU = int(os.environ.get('KAS_BENCH_KEYS', 100_000_000))
N_PREFIX = 64
REPLAY_Q = int(os.environ.get('KAS_BENCH_REPLAY_Q', 1_000_000))
REPLAY_T = int(os.environ.get('KAS_BENCH_REPLAY_T', 10_000))
LONG_TAIL = os.environ.get('KAS_BENCH_LONG_TAIL', '0') == '1'   # long-tailed key lengths (BASELINE config 5)
SEED = 2
SWEEP = [(1, 0, 1)] + [(k, a, b) for k in (1, 2, 5, 10, 50) for a in (0, 1) for b in (1, 4, 16, 64)]


def queue_names(n=N_PREFIX):
    base = ['predict', 'predict:zstack', 'track', 'tracker', 'train', 'segment', 'seg',
            'segment:nuclear', 'segment:cyto', 'spot', 'spots', 'caliban', 'polaris', 'mesmer',
            'mesmer:whole-cell', 'a:b', 'a:b:c', 'abc']
    i = 0
    while len(base) < n:
        base.append('queue-%02d-%s' % (i, 'x' * (i % 9)))
        i += 1
    return base[:n]


def measured_peak():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        with open(p) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json hbm_gbs)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md)'


def ncu_traffic(kernel):
    """DRAM bytes per launch from the committed ncu capture summary, or None."""
    try:
        with open(os.path.join(ROOT, 'profiles', 'traffic.json')) as f:
            return json.load(f).get(kernel)
    except Exception:
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
            time.sleep(0.005)

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
# CPU arm (oracle port)
# --------------------------------------------------------------------------
def cpu_tally_sample(n_sample, threads):
    """Time the oracle's C restatement (one full pass per queue, like the reference)
    on n_sample keys of the same synthetic keyspace.  Returns (keys/s, seconds)."""
    from autoscaler import engine
    from oracle import coracle
    names = queue_names()
    W = engine.rec_bytes_for(names)
    heads, _, _ = engine.synth_keys_host(SEED, 0, n_sample, names, long_tail=LONG_TAIL, rec_bytes=W)
    coracle.match_counts_heads(names, heads[:1000], threads=threads)
    t0 = time.perf_counter()
    counts = coracle.match_counts_heads(names, heads, threads=threads)
    dt = time.perf_counter() - t0
    return n_sample / dt, dt, counts


def cpu_python_port_sample(n_sample):
    """The line-by-line Python port (1 core, like the single-threaded reference)."""
    from autoscaler import engine
    from oracle import oracle
    names = queue_names()
    _, blob, off = engine.synth_keys_host(SEED, 0, n_sample, names, want_heads=False, want_blob=True)
    keys = [bytes(blob[off[i]:off[i + 1]]) for i in range(n_sample)]
    t0 = time.perf_counter()
    oracle.match_counts(names, keys)
    dt = time.perf_counter() - t0
    return n_sample / dt, dt


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    from oracle import coracle
    threads = coracle.max_threads()
    n_sample = int(os.environ.get('KAS_BENCH_CPU_KEYS', 16_000_000))
    for _ in range(min(args.warmup, 1)):
        cpu_tally_sample(min(n_sample, 200_000), threads)
    rates, secs = [], []
    steps = max(1, min(args.steps, 5))
    for _ in range(steps):
        r, dt, _ = cpu_tally_sample(n_sample, threads)
        rates.append(r)
        secs.append(dt)
    value = len(rates) * n_sample / sum(secs)
    sample = '%d keys x %d queues of the C2 keyspace per step (bounded sample of 1e8)' % (
        n_sample, N_PREFIX)
    line = {
        'impl': 'reference', 'metric': 'keys matched/sec', 'value': value, 'unit': 'keys/s',
        'n_gpus': args.gpus, 'steps': steps, 'warmup': min(args.warmup, 1),
        'ms_per_step': 1000.0 * sum(secs) / len(secs), 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'u8', 'data': 'synthetic',
        'config': bench_config(args.gpus),
        'cpu_baseline': {'value': value, 'unit': 'keys/s', 'cores': threads, 'kind': 'port',
                         'sample': sample},
        'e2e': {'value': value, 'unit': 'keys/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    emit(line)


def bench_config(n_gpus):
    return {'workload': 'C2 keyspace tally: %d synthetic keys per GPU x %d queue prefixes, '
                        '32-byte heads, bit-exact vs tally_queues' % (N_KEYS_PER_GPU, N_PREFIX),
            'keys_per_gpu': N_KEYS_PER_GPU, 'n_prefix': N_PREFIX, 'rec_bytes': 32, 'long_tail_keys': LONG_TAIL,
            'sharding': 'keys split by rank, counts all-reduced (NCCL)' if n_gpus > 1 else 'none',
            'l2': 'inputs (%.1f GB per GPU) larger than the 126 MB L2; no flush needed'
                  % (N_KEYS_PER_GPU * 32 / 1e9)}


# --------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from autoscaler import engine, replay as kreplay

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device (there is no CPU fallback); '
                         'use --impl reference for the CPU arm')
    torch.cuda.set_device(local)
    numa = 'unchanged'
    if world > 1:
        # keep this rank's host threads (and the pinned memory they first touch) on the CPUs next
        # to its GPU: the end-to-end leg is bound by the host-to-device copy
        try:
            import pynvml
            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
            numa = 'cpu affinity of the GPU (NVML)'
        except Exception as err:   # not fatal: only the e2e leg is affected
            numa = 'unchanged (%s)' % type(err).__name__
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    dev = torch.device('cuda', local)

    names = queue_names()
    W = engine.rec_bytes_for(names)
    assert W == 32
    n = N_KEYS_PER_GPU
    first = rank * n                       # each rank owns its own contiguous key range
    table = torch.from_numpy(engine.build_prefix_table(names, W)).to(dev)

    host_heads = torch.empty(n * W, dtype=torch.uint8).pin_memory()
    t0 = time.perf_counter()
    engine.synth_keys_host(SEED, first, n, names, long_tail=LONG_TAIL, rec_bytes=W, heads_out=host_heads)
    gen_s = time.perf_counter() - t0
    d_heads = torch.empty(n * W, dtype=torch.uint8, device=dev)
    d_heads.copy_(host_heads, non_blocking=True)
    d_counts = torch.empty(N_PREFIX, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()

    def step():
        engine.match_count(d_heads, n, W, table, N_PREFIX, d_counts=d_counts, accumulate=False)
        if world > 1:
            dist.all_reduce(d_counts)

    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(max(3, args.warmup)):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.active.set()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    sampler.active.clear()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms_total = e0.elapsed_time(e1)
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    ms_per_step = ms_total / args.steps
    value = world * n / (ms_per_step * 1e-3)
    counts_box = engine.counts_to_host(d_counts).astype(np.int64)

    # Kernel time for the roofline.  On one GPU a step IS one launch of the match kernel, so
    # the timed region above already measured it (CUDA events around exactly these launches);
    # with several ranks a step also holds the all-reduce, so the kernel is timed on its own.
    if world == 1:
        k_ms = ms_per_step
    else:
        for _ in range(3):
            engine.match_count(d_heads, n, W, table, N_PREFIX, d_counts=d_counts)
        torch.cuda.synchronize()
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        kn = max(10, min(args.steps, 200))
        k0.record()
        for _ in range(kn):
            engine.match_count(d_heads, n, W, table, N_PREFIX, d_counts=d_counts)
        k1.record()
        torch.cuda.synchronize()
        k_ms = k0.elapsed_time(k1) / kn
    engine.match_count(d_heads, n, W, table, N_PREFIX, d_counts=d_counts)
    local_counts = engine.counts_to_host(d_counts).astype(np.int64)
    peak, peak_src = measured_peak()
    alg_bytes = n * W + 8 * N_PREFIX
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    roofline = {'bound': 'hbm', 'kernel': 'match_count_kernel<2,true,*> (32-byte records, 256-bit loads)', 'achieved': achieved,
                'peak': peak, 'unit': 'GB/s', 'frac': achieved / peak, 'peak_source': peak_src,
                'algorithmic_bytes_per_launch': alg_bytes, 'kernel_ms': k_ms,
                'traffic': ncu_traffic('match_count_kernel')}

    # ---- end to end: host heads -> H2D -> K1 -> counts on the host -----------------
    e2e_steps = max(2, min(args.steps, 4))
    from autoscaler import snapshot as ksnap
    streamer = ksnap.StreamingTally(names, chunk_keys=8_388_608, device=dev)
    out_host = streamer.host_counts

    def e2e_step():
        # the public end-to-end call: pinned host heads -> H2D -> K1 -> counts on the host
        streamer.enqueue(host_heads, n, reduce=(dist.all_reduce if world > 1 else None))

    e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    w0 = time.perf_counter()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    for _ in range(e2e_steps):
        e2e_step()
    g1.record()
    torch.cuda.synchronize()
    e2e_ms = g0.elapsed_time(g1) / e2e_steps
    e2e_wall_ms = (time.perf_counter() - w0) * 1e3 / e2e_steps
    t = torch.tensor([max(e2e_ms, e2e_wall_ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item())
    e2e_ok = bool((out_host.numpy() == counts_box).all())
    if not e2e_ok:
        raise SystemExit('bench.py: end-to-end counts differ from the resident-data counts')
    e2e = {'value': world * n / (e2e_ms * 1e-3), 'unit': 'keys/s', 'h2d_bytes_per_step': n * W,
           'd2h_bytes_per_step': 8 * N_PREFIX, 'ms_per_step': e2e_ms, 'steps': e2e_steps,
           'matches_resident_result': e2e_ok,
           'path': 'autoscaler.snapshot.StreamingTally: pinned host heads -> chunked double-buffered '
                   'H2D -> kas_match_count_accumulate -> counts D2H'}
    del streamer

    # ---- ingest: variable-length keys -> packed heads (long-tailed lengths, C5-like) ------
    ingest = None
    if rank == 0 and not args.no_cpu:
        n_ing = int(os.environ.get('KAS_BENCH_INGEST_KEYS', 16_000_000))
        _, iblob, ioff = engine.synth_keys_host(SEED + 7, 0, n_ing, names, long_tail=True,
                                                want_heads=False, want_blob=True)
        t0 = time.perf_counter()
        hh = engine.pack_heads_host(iblob, ioff, W)
        host_s = time.perf_counter() - t0
        d_blob = torch.from_numpy(iblob).to(dev)
        d_off = torch.from_numpy(ioff.view(np.int64)).to(dev)
        for _ in range(3):
            dh = engine.pack_heads_device(d_blob, d_off, n_ing, W)
        torch.cuda.synchronize()
        i0, i1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        i0.record()
        for _ in range(10):
            dh = engine.pack_heads_device(d_blob, d_off, n_ing, W)
        i1.record()
        torch.cuda.synchronize()
        pack_ms = i0.elapsed_time(i1) / 10
        same = bool((dh.cpu().numpy() == hh).all())
        if not same:
            raise SystemExit('bench.py: device and host head packers disagree')
        # bytes the packer has to move: the first min(len, W) bytes of every key, one offset
        # per key, and the W-byte head written out (the tails of long keys are never read)
        lens = np.diff(ioff.astype(np.int64))
        ing_bytes = int(np.minimum(lens, W).sum()) + 8 * (n_ing + 1) + n_ing * W
        ingest = {'keys': n_ing, 'mean_key_bytes': float(iblob.size) / n_ing,
                  'device_pack': {'ms': pack_ms, 'keys_per_sec': n_ing / (pack_ms * 1e-3),
                                  'roofline': {'bound': 'hbm', 'achieved': ing_bytes / (pack_ms * 1e-3) / 1e9,
                                               'peak': peak, 'unit': 'GB/s',
                                               'frac': ing_bytes / (pack_ms * 1e-3) / 1e9 / peak,
                                               'algorithmic_bytes_per_launch': ing_bytes}},
                  'host_pack': {'keys_per_sec': n_ing / host_s, 'threads': engine.host_threads()},
                  'device_equals_host': same}
        del d_blob, d_off, dh, hh, iblob, ioff

    # ---- replay (K2), BASELINE config 3 per GPU -------------------------------------
    rep = None
    if not args.no_replay:
        Q, T = REPLAY_Q, REPLAY_T
        trace = engine.synth_trace(SEED + 1, T, rank * Q, Q, device=dev)
        torch.cuda.synchronize()

        def timed(policies, reps, warm):
            S = len(policies)
            for _ in range(warm):
                out = engine.replay(trace, T, Q, policies, want_totals=True)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()          # ranks enter the timed region together
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                out = engine.replay(trace, T, Q, policies, want_totals=True)
                kreplay.allreduce_sum_(out['sums'])
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b) / reps
            tt = torch.tensor([ms], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms = float(tt.item())
            return ms, world * Q * T * S / (ms * 1e-3), out

        ms1, dps1, out1 = timed([(1, 0, 1)], 10, 3)
        msS, dpsS, outS = timed(SWEEP, 2, 1)
        gbs1 = Q * T * 4 / (ms1 * 1e-3) / 1e9
        rep = {'metric': 'scaling decisions/sec', 'unit': '(queue,tick,policy) decisions/s',
               'workload': 'C3 policy replay: %d single-queue deployment resources per GPU x %d '
                           'ticks, int32 tick-major trace resident in HBM (%.0f GB per GPU)'
                           % (Q, T, Q * T * 4 / 1e9),
               's1': {'policies': 1, 'ms': ms1, 'decisions_per_sec': dps1,
                      'roofline': {'bound': 'hbm', 'achieved': gbs1, 'peak': peak, 'unit': 'GB/s',
                                   'frac': gbs1 / peak, 'algorithmic_bytes_per_launch': Q * T * 4,
                                   'traffic': ncu_traffic('replay_identity_kernel')}},
               'sweep': {'policies': len(SWEEP), 'ms': msS, 'decisions_per_sec': dpsS,
                         'bound': 'integer ALU (41 decisions per 4 loaded bytes)'},
               'pod_ticks_default_policy': int(out1['sums'][0, 0].item())}
        # the same trace stored as uint16 (BASELINE config 3's "uint16 variant": half the bytes);
        # the synthetic depths are below 2^16, so the integers are the same
        if int(trace.max().item()) < 65536:
            trace16 = trace.to(torch.int16)      # same bits as uint16 below 2^15; the library reads uint16
            del trace
            trace = trace16
            ms16, dps16, out16 = timed([(1, 0, 1)], 10, 3)
            msS16, dpsS16, outS16 = timed(SWEEP, 2, 1)
            if int(out16['sums'][0, 0].item()) != rep['pod_ticks_default_policy'] or \
                    not bool((outS16['sums'] == outS['sums']).all()):
                raise SystemExit('bench.py: uint16 and int32 traces replay to different integers')
            gbs16 = Q * T * 2 / (ms16 * 1e-3) / 1e9
            rep['s1_uint16'] = {'policies': 1, 'ms': ms16, 'decisions_per_sec': dps16,
                                'roofline': {'bound': 'hbm', 'achieved': gbs16, 'peak': peak, 'unit': 'GB/s',
                                             'frac': gbs16 / peak, 'algorithmic_bytes_per_launch': Q * T * 2}}
            rep['sweep_uint16'] = {'policies': len(SWEEP), 'ms': msS16, 'decisions_per_sec': dpsS16}
        del trace

    clocks = None
    sampler.stop_flag.set()
    sampler.join(timeout=2)
    clocks = sampler.summary()

    # ---- CPU baseline (rank 0, N = 1 only) ---------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        from oracle import coracle
        threads = coracle.max_threads()
        n_sample = int(os.environ.get('KAS_BENCH_CPU_KEYS', 16_000_000))
        rate, dt, ccounts = cpu_tally_sample(n_sample, threads)
        # the same sample through the GPU path must give the same counts
        c = engine.match_count(d_heads, n_sample, W, table, N_PREFIX)
        agree = bool((engine.counts_to_host(c).astype(np.int64) == ccounts.astype(np.int64)).all())
        if not agree:
            raise SystemExit('bench.py: GPU counts differ from the CPU oracle on the sample -- '
                             'a throughput number for wrong results is not a result')
        prate, pdt = cpu_python_port_sample(20_000)
        cpu = {'value': rate, 'unit': 'keys/s', 'cores': threads, 'kind': 'port',
               'sample': '%d keys x %d queues of the same keyspace, C restatement, one full pass '
                         'per queue like the reference; %.1f s' % (n_sample, N_PREFIX, dt),
               'gpu_counts_equal_on_sample': agree,
               'python_port': {'value': prate, 'unit': 'keys/s', 'cores': 1,
                               'sample': '20000 keys x %d queues, line-by-line Python port; %.1f s'
                                         % (N_PREFIX, pdt)}}

    if rank == 0:
        line = {
            'metric': 'keys matched/sec', 'value': value, 'unit': 'keys/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': max(3, args.warmup), 'ms_per_step': ms_per_step,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'u8',
            'data': 'synthetic', 'config': bench_config(world), 'roofline': roofline,
            'cpu_baseline': cpu, 'e2e': e2e, 'gpu_launches': args.steps, 'clocks': clocks,
            'replay': rep, 'ingest': ingest,
            'checks': {'keys_matched_fraction': float(counts_box.sum()) / (world * n),
                       'host_generation_s': gen_s, 'host_affinity': numa,
                       'local_counts_sum': int(local_counts.sum())},
        }
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


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
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--no-replay', action='store_true')
    ap.add_argument('--no-cpu', action='store_true')
    args = ap.parse_args()
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_ours(args)


if __name__ == '__main__':
    main()
