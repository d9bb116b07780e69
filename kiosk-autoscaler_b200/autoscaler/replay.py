"""Offline replay of the scale loop over a queue-depth trace (K2) and the
sharding of keys / resources over the GPUs of one box.

One tick of the reference is ``Autoscaler.scale`` (autoscaler.py:244-273 of the
reference): tally, read current pods, sum of per-queue decisions, clip, patch.
With a cluster that applies every patch, the replicas after tick t are the
current pods of tick t+1, so a resource is a sequential recurrence over ticks
while resources (and policies) are independent -- that is what is sharded.
"""
from __future__ import annotations

import numpy as np

from autoscaler import engine

RESOURCE_FIELD = {'deployment': 'replicas', 'job': 'parallelism'}   # autoscaler.py:232,236


class ReplayResult(object):
    """Per (resource, policy) reductions of a replay, on the host.

    pod_ticks[r][s]  sum over ticks of the replicas after the tick
    n_patches[r][s]  ticks on which scale_resource would have patched
    final[r][s]      replicas after the last tick
    sums[s]          (sum_r pod_ticks, sum_r n_patches, sum_r final)
    dense[s][t][r]   replicas after each tick (only when requested)
    """

    def __init__(self, policies, totals=None, sums=None, dense=None, n_ticks=0):
        self.policies = [tuple(int(v) for v in p) for p in policies]
        self.n_ticks = n_ticks
        self.pod_ticks = self.n_patches = self.final = None
        if totals is not None:
            self.pod_ticks = totals['pod_ticks'].astype(np.int64)
            self.n_patches = totals['n_patches'].astype(np.int64)
            self.final = totals['final_current'].astype(np.int64)
        self.sums = sums
        self.dense = dense
        self._cur0 = None
        self._bodies_from_dense = True

    def pod_hours(self, interval_seconds=5):
        """Box-wide pod-hours per policy; INTERVAL as in scale.py:84."""
        return [int(s[0]) * interval_seconds / 3600.0 for s in self.sums]

    def patch_bodies(self, resource_type='deployment', resource=0, policy=0):
        """The bodies scale_resource would have sent (autoscaler.py:232,236),
        reconstructed from the dense trajectory of one resource.  A replay with
        failing patch calls or a start-up delay makes calls that leave no trace in the
        trajectory (a failed call; a repeated patch of the value already in the spec):
        take those from ``patch_events`` (PatchLog.bodies)."""
        if not self._bodies_from_dense:
            raise ValueError('this replay ran with patch_fail / startup_delay: its patch calls '
                             'cannot be read off the trajectory; use autoscaler.replay.patch_events')
        if self.dense is None:
            raise ValueError('patch_bodies needs the dense trajectory (want_dense=True)')
        field = RESOURCE_FIELD[resource_type]
        out = []
        prev = int(self._cur0[resource])
        for t, v in enumerate(self.dense[policy, :, resource]):
            v = int(v)
            if v != prev:
                out.append((t, {'spec': {field: v}}))
            prev = v
        return out


def sweep_report(result, interval_seconds=5, cost_per_pod_hour=1.0, first_pod=None):
    """Per-policy summary of a replay: box-wide pod-hours (pod_ticks x INTERVAL / 3600,
    INTERVAL as in scale.py:84), cost, number of patch calls (churn) and whether the
    policy is on the cost / churn Pareto front.  All derived on the host from the
    integer sums, so the report is identical at every world size.  ``first_pod``: the
    list ``first_pod_latency`` returned for the same policies adds the responsiveness
    columns (bursts served / unserved, mean and longest wait for the first pod)."""
    rows = []
    for i, ((kpp, mn, mx), s) in enumerate(zip(result.policies, result.sums)):
        hours = int(s[0]) * interval_seconds / 3600.0
        rows.append(dict(keys_per_pod=kpp, min_pods=mn, max_pods=mx, pod_ticks=int(s[0]),
                         pod_hours=hours, cost=hours * cost_per_pod_hour, n_patches=int(s[1]),
                         final_pods=int(s[2])))
        if first_pod is not None:
            f = first_pod[i]
            rows[-1].update(bursts_served=f['n_served'], bursts_unserved=f['n_unserved'],
                            max_wait_ticks=f['wait_max'], mean_wait_ticks=f['mean_wait'],
                            mean_wait_seconds=f['mean_wait'] * interval_seconds,
                            bursts_pending=f['n_pending'])
    for a in rows:
        a['pareto'] = not any(
            (b['pod_ticks'] <= a['pod_ticks'] and b['n_patches'] <= a['n_patches'] and
             (b['pod_ticks'] < a['pod_ticks'] or b['n_patches'] < a['n_patches'])) for b in rows)
    return rows


def _check_type(resource_type, managed=('deployment', 'job')):
    if resource_type not in managed:
        raise ValueError('`resource_type` must be one of {}. Got {}.'.format(set(managed), resource_type))


def _to_device_i32(x, torch, what):
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        if x.dtype != torch.int32:
            raise TypeError('%s must be int32' % what)
        return x.contiguous().to('cuda')
    a = np.asarray(x)
    if a.size and (a.min() < -(2 ** 31) or a.max() >= 2 ** 31):
        raise OverflowError('%s must fit in int32' % what)
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.int32)).to('cuda')


def _to_device_trace(x, torch, what, like=None):
    """A trace on the device: uint16 stays uint16 (half the bytes), everything else int32."""
    if x is None:
        return None
    if isinstance(x, torch.Tensor):
        if x.dtype not in (torch.int32, torch.uint16, torch.int16):
            raise TypeError('%s must be int32 or uint16' % what)
        x = x.contiguous()
        # a pinned host tensor is copied asynchronously on the current stream
        return x.to('cuda', non_blocking=x.device.type == 'cpu' and x.is_pinned())
    a = np.asarray(x)
    want_u16 = a.dtype == np.uint16 if like is None else like.dtype in (torch.uint16, torch.int16)
    if want_u16:
        if a.size and (a.min() < 0 or a.max() > 65535):
            raise OverflowError('%s does not fit the uint16 trace format' % what)
        # int16 storage: same bits, universally supported by torch; the library reads uint16
        return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint16).view(np.int16)).to('cuda')
    return _to_device_i32(a, torch, what)


def _to_device_fail(patch_fail, torch, T, R):
    if patch_fail is None:
        return None
    a = np.ascontiguousarray(np.asarray(patch_fail) != 0, dtype=np.uint8)
    if a.shape != (T, R):
        raise ValueError('patch_fail must be [T][R] = [%d][%d]' % (T, R))
    return torch.from_numpy(a).to('cuda')


def _delay_for(resource_type, startup_delay):
    """Only a deployment has an `only_running` reading (autoscaler.py:176-177 of the
    reference); for a job get_current_pods returns spec.parallelism whatever the flag
    (autoscaler.py:185-190), so the start-up delay does not apply."""
    d = engine._check_delay(startup_delay)
    return d if resource_type == 'deployment' else 0


def replay_fleet(llen, policies, inprog=None, group_off=None, cur0=None, want_totals=True,
                 want_dense=False, resource_type='deployment', patch_fail=None, startup_delay=0):
    """Replay many resources at once.

    llen       int32 or uint16 [T][Q] (numpy or CUDA tensor), tick-major
    inprog     None | [Q] | [T][Q]
    group_off  None (every queue is its own resource) | uint64 [R+1]
    cur0       None | int32 [R]
    policies   sequence of (keys_per_pod, min_pods, max_pods)
    patch_fail None | bool [T][R]: the patch call of that tick and resource raises
               ApiException, which ``scale`` swallows (autoscaler.py:267-273 of the
               reference): the call counts, the replicas stay
    startup_delay  0: the controller reads spec.replicas, as ``scale`` does.  D > 0: it
               reads status.available_replicas (``get_current_pods(only_running=True)``,
               autoscaler.py:176-177) and a replica takes D ticks in the spec to become
               available (model in include/kas.h); results stay spec.replicas per tick
    """
    _check_type(resource_type)
    torch = engine.require_cuda()
    d_llen = _to_device_trace(llen, torch, 'llen')
    T, Q = int(d_llen.shape[0]), int(d_llen.shape[1])
    if inprog is not None and np.ndim(inprog) == 2:
        d_ip = _to_device_trace(inprog, torch, 'inprog', like=d_llen)
    else:
        d_ip = _to_device_i32(inprog, torch, 'inprog')
    mode = 0 if d_ip is None else (1 if d_ip.dim() == 1 else 2)
    d_go = None
    R = Q
    if group_off is not None:
        go = np.ascontiguousarray(group_off, dtype=np.uint64)
        if go[0] != 0 or go[-1] != Q or (np.diff(go.astype(np.int64)) < 0).any():
            raise ValueError('group_off must rise from 0 to Q')
        R = len(go) - 1
        d_go = torch.from_numpy(go.view(np.int64)).to('cuda')
    d_c0 = _to_device_i32(cur0, torch, 'cur0')
    out = engine.replay(d_llen, T, Q, policies, d_inprog=d_ip, inprog_mode=mode, d_group_off=d_go,
                        n_resources=R, d_cur0=d_c0, want_totals=want_totals, want_sums=True,
                        want_dense=want_dense, d_patch_fail=_to_device_fail(patch_fail, torch, T, R),
                        startup_delay=_delay_for(resource_type, startup_delay))
    res = ReplayResult(policies,
                       totals=engine.totals_to_host(out['totals']) if want_totals else None,
                       sums=out['sums'].cpu().numpy(),
                       dense=out['dense'].cpu().numpy() if want_dense else None, n_ticks=T)
    res._cur0 = np.zeros(R, dtype=np.int64) if cur0 is None else np.asarray(cur0, dtype=np.int64)
    res._bodies_from_dense = patch_fail is None and _delay_for(resource_type, startup_delay) == 0
    return res


class PatchLog(object):
    """The patch calls of a replay, per resource (CSR) and in tick order:
    ``events[off[r]:off[r+1]]`` with fields tick / from_pods / to_pods."""

    def __init__(self, off, events, resource_types):
        self.off, self.events, self.resource_types = off, events, resource_types

    def __len__(self):
        return int(self.off[-1])

    def of(self, r):
        return self.events[int(self.off[r]):int(self.off[r + 1])]

    def bodies(self, r):
        """(tick, body) pairs as scale_resource would have sent them
        (autoscaler.py:232 for a job, :236 for a deployment)."""
        rt = self.resource_types if isinstance(self.resource_types, str) else self.resource_types[r]
        field = RESOURCE_FIELD[rt]
        return [(int(e['tick']), {'spec': {field: int(e['to_pods'])}}) for e in self.of(r)]


def patch_events(llen, policy, inprog=None, group_off=None, cur0=None, resource_types='deployment',
                 patch_fail=None, startup_delay=0):
    """Patch-event stream of ONE policy over a fleet (same inputs as replay_fleet).
    ``resource_types`` is one type for all resources or a sequence with one
    entry per resource ('deployment' / 'job'); it only selects the body field --
    except under ``startup_delay``, which only deployments have: a fleet replayed
    with a delay must be of one type."""
    types = [resource_types] if isinstance(resource_types, str) else list(resource_types)
    for rt in types:
        _check_type(rt)
    if engine._check_delay(startup_delay) and len(set(types)) > 1:
        raise ValueError('startup_delay needs a fleet of one resource type (jobs have no '
                         'available_replicas); replay deployments and jobs separately')
    delay = _delay_for(types[0], startup_delay) if types else 0
    torch = engine.require_cuda()
    d_llen = _to_device_trace(llen, torch, 'llen')
    T, Q = int(d_llen.shape[0]), int(d_llen.shape[1])
    if inprog is not None and np.ndim(inprog) == 2:
        d_ip = _to_device_trace(inprog, torch, 'inprog', like=d_llen)
    else:
        d_ip = _to_device_i32(inprog, torch, 'inprog')
    mode = 0 if d_ip is None else (1 if d_ip.dim() == 1 else 2)
    d_go, R = None, Q
    if group_off is not None:
        go = np.ascontiguousarray(group_off, dtype=np.uint64)
        if go[0] != 0 or go[-1] != Q or (np.diff(go.astype(np.int64)) < 0).any():
            raise ValueError('group_off must rise from 0 to Q')
        R = len(go) - 1
        d_go = torch.from_numpy(go.view(np.int64)).to('cuda')
    if not isinstance(resource_types, str) and len(types) != R:
        raise ValueError('resource_types must have one entry per resource')
    d_c0 = _to_device_i32(cur0, torch, 'cur0')
    off, ev = engine.patch_events(d_llen, T, Q, policy, d_inprog=d_ip, inprog_mode=mode,
                                  d_group_off=d_go, n_resources=R, d_cur0=d_c0,
                                  d_patch_fail=_to_device_fail(patch_fail, torch, T, R),
                                  startup_delay=delay)
    return PatchLog(off, ev, resource_types)


def first_pod_latency(llen, policies, inprog=None, group_off=None, cur0=None,
                      resource_type='deployment', patch_fail=None, startup_delay=0, want_rows=False,
                      comm=None):
    """Time to the first pod after a burst, per policy, over a fleet (same inputs as
    ``replay_fleet``; definition in include/kas.h).  Returns one dict per policy with the
    box-wide integers n_served, n_unserved, wait_sum, wait_max, pending_ticks, n_pending (when
    ``want_rows``), mean_wait (ticks) and, with ``want_rows``, ``rows`` (numpy structured [R]).
    With a communicator / process group the sums are all-reduced (the maximum as a maximum)."""
    _check_type(resource_type)
    torch = engine.require_cuda()
    d_llen = _to_device_trace(llen, torch, 'llen')
    T, Q = int(d_llen.shape[0]), int(d_llen.shape[1])
    if inprog is not None and np.ndim(inprog) == 2:
        d_ip = _to_device_trace(inprog, torch, 'inprog', like=d_llen)
    else:
        d_ip = _to_device_i32(inprog, torch, 'inprog')
    mode = 0 if d_ip is None else (1 if d_ip.dim() == 1 else 2)
    d_go, R = None, Q
    if group_off is not None:
        go = np.ascontiguousarray(group_off, dtype=np.uint64)
        if go[0] != 0 or go[-1] != Q or (np.diff(go.astype(np.int64)) < 0).any():
            raise ValueError('group_off must rise from 0 to Q')
        R = len(go) - 1
        d_go = torch.from_numpy(go.view(np.int64)).to('cuda')
    d_c0 = _to_device_i32(cur0, torch, 'cur0')
    d_pf = _to_device_fail(patch_fail, torch, T, R)
    delay = _delay_for(resource_type, startup_delay)
    out = []
    for pol in policies:
        rows, sums = engine.first_pod_latency(d_llen, T, Q, pol, d_inprog=d_ip, inprog_mode=mode,
                                              d_group_off=d_go, n_resources=R, d_cur0=d_c0,
                                              d_patch_fail=d_pf, startup_delay=delay, want_rows=want_rows)
        sums = _allreduce_first_pod(sums, comm)
        served = int(sums[0])
        d = dict(n_served=served, n_unserved=int(sums[1]), wait_sum=int(sums[2]), wait_max=int(sums[3]),
                 pending_ticks=int(sums[4]), mean_wait=(int(sums[2]) / served) if served else 0.0,
                 n_pending=int((rows['pending'] > 0).sum()) if rows is not None else None)
        if want_rows:
            d['rows'] = rows
        out.append(d)
    return out


def _allreduce_first_pod(sums, comm):
    """[served, unserved, wait_sum, max wait, pending]: four sums and one maximum over the ranks,
    carried by ONE integer sum all-reduce (every rank puts its maximum in its own slot)."""
    import torch
    import torch.distributed as dist
    if comm is not None:
        world, rank, dev = comm.world_size, comm.rank, 'cuda'
    elif dist.is_available() and dist.is_initialized():
        world, rank = dist.get_world_size(), dist.get_rank()
        dev = 'cuda' if dist.get_backend() == 'nccl' else 'cpu'
    else:
        world = 1
    if world == 1:
        return sums
    v = np.zeros(4 + world, dtype=np.int64)
    v[:4] = [sums[0], sums[1], sums[2], sums[4]]
    v[4 + rank] = sums[3]
    t = torch.from_numpy(v).to(dev)
    allreduce_sum_(t, comm=comm)
    v = t.cpu().numpy()
    return np.array([v[0], v[1], v[2], v[4:].max(), v[3]], dtype=np.int64)


class StreamingReplay(object):
    """Replay a trace that sits in host memory -- or is longer than the GPU's memory -- in chunks
    of ticks.  The scale loop keeps no state between ticks but the pods Kubernetes holds
    (autoscaler.py:252 of the reference), so a chunk only needs the final pods of the chunk before
    it: chunk k+1 resumes from chunk k's totals (``kas_replay_args.d_resume``) and
    ``kas_totals_chain`` adds the totals up.  Chunks are copied into two device staging buffers
    on a copy stream while the previous chunk is replayed; the result is, bit for bit, that of
    one replay of the whole trace.  (A start-up delay cannot be chunked: its availability window
    is not part of the totals.)"""

    def __init__(self, policies, n_queues, group_off=None, chunk_ticks=512, trace_dtype='int32',
                 per_tick_inprog=False, device='cuda'):
        torch = engine.require_cuda()
        self.policies = [tuple(int(v) for v in p) for p in policies]
        self.Q = int(n_queues)
        self.device = torch.device(device)
        self.chunk_ticks = int(chunk_ticks)
        self.d_go, self.R = None, self.Q
        if group_off is not None:
            go = np.ascontiguousarray(group_off, dtype=np.uint64)
            if go[0] != 0 or go[-1] != self.Q or (np.diff(go.astype(np.int64)) < 0).any():
                raise ValueError('group_off must rise from 0 to Q')
            self.R = len(go) - 1
            self.d_go = torch.from_numpy(go.view(np.int64)).to(self.device)
        self.tdtype = torch.int16 if trace_dtype in ('uint16', np.uint16) else torch.int32
        S = len(self.policies)
        n_stage = 2 if not per_tick_inprog else 4
        self.stage = [torch.empty((self.chunk_ticks, self.Q), dtype=self.tdtype, device=self.device)
                      for _ in range(n_stage)]
        self.free = [torch.cuda.Event() for _ in range(2)]
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self.acc = torch.empty((self.R, S, 16), dtype=torch.uint8, device=self.device)
        self.tmp = torch.empty((self.R, S, 16), dtype=torch.uint8, device=self.device)
        self.sums_acc = torch.empty((S, 3), dtype=torch.int64, device=self.device)
        self.sums_tmp = torch.empty((S, 3), dtype=torch.int64, device=self.device)
        self.per_tick_inprog = per_tick_inprog
        self.chunks = 0

    def _host(self, x, what):
        torch = engine.require_cuda()
        if x is None:
            return None
        if not isinstance(x, torch.Tensor):
            a = np.ascontiguousarray(x)
            if self.tdtype == torch.int16:
                if a.size and (a.min() < 0 or a.max() > 65535):
                    raise OverflowError('%s does not fit the uint16 trace format' % what)
                a = a.astype(np.uint16).view(np.int16)
            else:
                a = a.astype(np.int32, copy=False)
            x = torch.from_numpy(a)
        if x.dtype == torch.uint16:
            x = x.view(torch.int16)
        if x.dtype != self.tdtype or x.dim() != 2 or x.shape[1] != self.Q:
            raise TypeError('%s must be [T][%d] of the streaming replay\'s trace type' % (what, self.Q))
        return x.contiguous()

    def run(self, host_trace, inprog=None, cur0=None):
        """host_trace: [T][Q] numpy array or CPU tensor (pinned for asynchronous copies).
        inprog: None, per-queue int32 [Q], or -- with per_tick_inprog -- a second [T][Q] host trace.
        cur0: pods before the first tick, int32 [R].  Returns a ``ReplayResult``."""
        torch = engine.require_cuda()
        llen = self._host(host_trace, 'host_trace')
        T = int(llen.shape[0])
        tick_ip = self._host(inprog, 'inprog') if self.per_tick_inprog else None
        d_ipq = None if self.per_tick_inprog else _to_device_i32(inprog, torch, 'inprog')
        d_c0 = _to_device_i32(cur0, torch, 'cur0')
        main = torch.cuda.current_stream(self.device)
        for b in range(2):
            self.free[b].record(main)
        for k, t0 in enumerate(range(0, T, self.chunk_ticks)):
            nt = min(self.chunk_ticks, T - t0)
            b = k & 1
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(self.free[b])
                self.stage[b][:nt].copy_(llen[t0:t0 + nt], non_blocking=True)
                if tick_ip is not None:
                    self.stage[2 + b][:nt].copy_(tick_ip[t0:t0 + nt], non_blocking=True)
                ready = torch.cuda.Event()
                ready.record(self.copy_stream)
            main.wait_event(ready)
            first = k == 0
            engine.replay(self.stage[b][:nt], nt, self.Q, self.policies,
                          d_inprog=self.stage[2 + b][:nt] if tick_ip is not None else d_ipq,
                          inprog_mode=2 if tick_ip is not None else (0 if d_ipq is None else 1),
                          d_group_off=self.d_go, n_resources=self.R,
                          d_cur0=d_c0 if first else None, d_resume=None if first else self.acc,
                          out_totals=self.acc if first else self.tmp,
                          out_sums=self.sums_acc if first else self.sums_tmp)
            if not first:
                engine.totals_chain(self.acc, self.tmp, self.sums_acc, self.sums_tmp)
            self.free[b].record(main)
            self.chunks += 1
        res = ReplayResult(self.policies, totals=engine.totals_to_host(self.acc),
                           sums=self.sums_acc.cpu().numpy(), n_ticks=T)
        res._bodies_from_dense = False
        return res


def replay_one(queue_names, llen_trace, policies, inprog=None, current_pods=0,
               resource_type='deployment', managed=('deployment', 'job'), want_dense=False,
               patch_fail=None, startup_delay=0):
    """One resource fed by ``len(queue_names)`` queues (what one reference pod
    manages): llen_trace [T][Q] in the order of ``queue_names``."""
    _check_type(resource_type, managed)
    a = np.asarray(llen_trace)
    if a.ndim != 2 or a.shape[1] != len(queue_names):
        raise ValueError('llen_trace must be [T][%d]' % len(queue_names))
    # (a scaler with ONE queue is a single-queue resource: no groups, so that a long trace can be
    # replayed tick-parallel)
    groups = None if len(queue_names) == 1 else np.array([0, len(queue_names)], dtype=np.uint64)
    return replay_fleet(a, policies, inprog=inprog, group_off=groups,
                        cur0=np.array([current_pods], dtype=np.int64), want_dense=want_dense,
                        resource_type=resource_type, startup_delay=startup_delay,
                        patch_fail=None if patch_fail is None else np.asarray(patch_fail).reshape(-1, 1))


# --------------------------------------------------------------------------
# sharding over the GPUs of one box
# --------------------------------------------------------------------------
def shard_range(n_units, rank, world_size, align=1):
    """Contiguous share [lo, hi) of ``n_units`` for ``rank``; boundaries are
    multiples of ``align`` (4 keeps the vectorised replay path)."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError('bad rank/world_size')
    blocks = (n_units + align - 1) // align
    lo = (blocks * rank) // world_size * align
    hi = (blocks * (rank + 1)) // world_size * align
    return min(lo, n_units), min(hi, n_units)


def shard_groups(group_off, rank, world_size):
    """Share of RESOURCES for ``rank`` (all queues of a resource stay together):
    returns (r_lo, r_hi, q_lo, q_hi, local_group_off)."""
    go = np.asarray(group_off, dtype=np.uint64)
    r_lo, r_hi = shard_range(len(go) - 1, rank, world_size)
    q_lo, q_hi = int(go[r_lo]), int(go[r_hi])
    return r_lo, r_hi, q_lo, q_hi, (go[r_lo:r_hi + 1] - go[r_lo]).astype(np.uint64)


def allreduce_sum_(t, group=None, comm=None):
    """Box-wide integer totals: ONE all-reduce.  With ``comm`` (an ``engine.Comm``) libkas
    issues the NCCL all-reduce itself on the current stream; otherwise torch.distributed
    carries it (NCCL for CUDA tensors, gloo in the CPU tests).  Integer sums are order
    independent, so the result is bit-exact at every world size.  No-op without a
    communicator or an initialised process group."""
    if comm is not None:
        return comm.allreduce_sum_(t) if comm.world_size > 1 else t
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t
