"""Asynchronous offload of saved activations to pinned host memory.

Reference: ``runtime/activation_checkpointing/offload_activations.py`` (``_ActivationOffloadEngine``,
``CheckpointHiddenStatesOffload``, ``CheckpointActivationOffloadStats``) and the DeepCompile ``offload_activation`` pass.

Design: one :class:`ActivationOffloadEngine` per context. ``offload(t)`` starts a device-to-host copy on a side stream
into a pinned buffer taken from a pool keyed by (shape, dtype) - pinned allocations are expensive, reuse is what makes
this cheap - and returns a ticket; the device tensor stays referenced until its copy has finished (a bounded "in
flight" window, so at most ``max_in_flight`` copies keep device memory alive). The most recent ``keep_last`` tensors
are never copied: backward consumes saved tensors in reverse order, so they would come straight back. ``restore(ticket)``
brings a tensor back on the side stream and makes the compute stream wait for it; the host buffer returns to the pool
when the copy back has completed. On a CPU-only run tensors pass through untouched (and the statistics say so)."""
import contextlib
from dataclasses import dataclass

import torch
from torch.autograd.graph import saved_tensors_hooks


@dataclass
class CheckpointActivationOffloadStats:
    saved_tensors_seen: int = 0
    marked_tensors: int = 0
    offloaded_tensors: int = 0
    restored_tensors: int = 0
    skipped_marked_tensors: int = 0
    kept_last_tensors: int = 0
    offloaded_bytes: int = 0
    restored_bytes: int = 0


class _Ticket:
    __slots__ = ("tid", )

    def __init__(self, tid):
        self.tid = tid


class ActivationOffloadEngine:

    def __init__(self, use_pin_memory=True, use_streams=True, min_offload_bytes=1024, max_in_flight=2,
                 max_cpu_buffer_pool_count=64, keep_last_count=1):
        for name, v in (("keep_last_count", keep_last_count), ("max_in_flight", max_in_flight),
                        ("max_cpu_buffer_pool_count", max_cpu_buffer_pool_count),
                        ("min_offload_bytes", min_offload_bytes)):
            if v < 0:
                raise ValueError(f"{name} must be >= 0, got {v}")
        self.cuda = torch.cuda.is_available()
        self.use_pin_memory = use_pin_memory and self.cuda
        self.use_streams = use_streams and self.cuda
        self.min_offload_bytes = min_offload_bytes
        self.max_in_flight = max_in_flight
        self.max_pool = max_cpu_buffer_pool_count
        self.keep_last_count = keep_last_count
        self.stats = CheckpointActivationOffloadStats()
        self.side = torch.cuda.Stream() if self.use_streams else None
        self._next = 0
        self._host = {}  # tid -> (host tensor, device, pool key, copy-done event)
        self._in_flight = {}  # tid -> device tensor kept alive until its copy is done
        self._kept = {}  # tid -> device tensor never offloaded (most recent ones)
        self._pool = {}  # key -> [host tensors]
        self._pool_count = 0
        self._returning = []  # (key, host tensor, event): buffers whose copy back is still running

    # ---- pinned buffer pool -------------------------------------------------------------------------------------
    def _key(self, t):
        return (tuple(t.shape), t.dtype)

    def _take(self, t):
        self._reap_returning()
        key = self._key(t)
        bufs = self._pool.get(key)
        if bufs:
            self._pool_count -= 1
            return bufs.pop(), key
        return torch.empty(t.shape, dtype=t.dtype, device="cpu", pin_memory=self.use_pin_memory), key

    def _give(self, key, host):
        if self._pool_count < self.max_pool:
            self._pool.setdefault(key, []).append(host)
            self._pool_count += 1

    def _reap_returning(self, force=False):
        still = []
        for key, host, ev in self._returning:
            if force and ev is not None:
                ev.synchronize()
            if ev is None or force or ev.query():
                self._give(key, host)
            else:
                still.append((key, host, ev))
        self._returning = still

    # ---- offload / restore ---------------------------------------------------------------------------------------
    def should_skip(self, t):
        return (not t.is_cuda or t.numel() * t.element_size() < self.min_offload_bytes
                or isinstance(t, torch.nn.Parameter) or (t.numel() > 1 and 0 in t.stride()))

    def _start_copy(self, tid, t):
        host, key = self._take(t)
        ev = None
        if self.use_streams:
            self.side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.side):
                host.copy_(t.detach(), non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.side)
            self._in_flight[tid] = t
            # bound the device memory held by copies still running
            for old in [i for i in self._in_flight if i <= tid - self.max_in_flight]:
                self._host[old][3].synchronize()
                del self._in_flight[old]
        else:
            host.copy_(t.detach())
        self._host[tid] = (host, t.device, key, ev)
        self.stats.offloaded_tensors += 1
        self.stats.offloaded_bytes += t.numel() * t.element_size()

    def offload(self, t):
        """Returns a ticket for :meth:`restore`, or ``t`` itself when it is not worth / possible to offload."""
        self.stats.saved_tensors_seen += 1
        if not torch.is_tensor(t) or self.should_skip(t):
            if torch.is_tensor(t) and t.is_cuda:
                self.stats.skipped_marked_tensors += 1
            return t
        self.stats.marked_tensors += 1
        self._next += 1
        tid = self._next
        if self.keep_last_count > 0:
            self._kept[tid] = t
            self.stats.kept_last_tensors += 1
            while len(self._kept) > self.keep_last_count:
                old = min(self._kept)
                self._start_copy(old, self._kept.pop(old))
                self.stats.kept_last_tensors -= 1
        else:
            self._start_copy(tid, t)
        return _Ticket(tid)

    def restore(self, ticket):
        if not isinstance(ticket, _Ticket):
            return ticket
        tid = ticket.tid
        if tid in self._kept:
            return self._kept.pop(tid)
        if tid in self._in_flight:  # never left the device's reach: reuse it, retire the host copy
            t = self._in_flight.pop(tid)
            host, _, key, ev = self._host.pop(tid)
            self._returning.append((key, host, ev))
            return t
        host, device, key, ev = self._host.pop(tid)
        if self.use_streams:
            with torch.cuda.stream(self.side):
                if ev is not None:
                    self.side.wait_event(ev)
                out = host.to(device, non_blocking=True)
                back = torch.cuda.Event()
                back.record(self.side)
            torch.cuda.current_stream().wait_event(back)
            out.record_stream(torch.cuda.current_stream())
            self._returning.append((key, host, back))
        else:
            out = host.to(device)
            self._give(key, host)
        self.stats.restored_tensors += 1
        self.stats.restored_bytes += out.numel() * out.element_size()
        return out

    def reset(self):
        """Drop everything still tracked (end of a step); pooled host buffers are kept for the next step."""
        self._reap_returning(force=True)
        for tid, (host, _, key, ev) in list(self._host.items()):
            if ev is not None:
                ev.synchronize()
            self._give(key, host)
        self._host.clear()
        self._in_flight.clear()
        self._kept.clear()


_MARK = "_ds_offload_activation"


def mark_for_offload(tensor):
    """Tag a tensor (e.g. the hidden-state input of a checkpointed layer) so that
    :class:`CheckpointHiddenStatesOffload` with ``only_marked=True`` moves exactly these to the host."""
    setattr(tensor, _MARK, True)
    return tensor


class CheckpointHiddenStatesOffload(contextlib.AbstractContextManager):
    """``with CheckpointHiddenStatesOffload(): loss = model(x); loss.backward()`` - tensors autograd saves inside the
    context live in pinned host memory between forward and backward. With non-reentrant activation checkpointing the
    saved tensors are the layer inputs, so this is "CPU checkpointing" with asynchronous copies. ``only_marked``:
    offload only tensors tagged by :func:`mark_for_offload`."""

    def __init__(self, use_pin_memory=True, use_streams=True, min_offload_bytes=1024, max_fwd_stash_count=2,
                 max_cpu_buffer_pool_count=64, keep_last_count=1, only_marked=False, engine=None):
        self.engine = engine or ActivationOffloadEngine(use_pin_memory, use_streams, min_offload_bytes,
                                                        max_fwd_stash_count, max_cpu_buffer_pool_count,
                                                        keep_last_count)
        self.only_marked = only_marked
        self._hooks = None

    @property
    def stats(self):
        return self.engine.stats

    def _pack(self, t):
        if self.only_marked and not getattr(t, _MARK, False):
            self.engine.stats.saved_tensors_seen += 1
            return t
        return self.engine.offload(t)

    def __enter__(self):
        self._hooks = saved_tensors_hooks(self._pack, self.engine.restore)
        self._hooks.__enter__()
        return self

    def __exit__(self, *exc):
        self._hooks.__exit__(*exc)
        self._hooks = None
        return False
