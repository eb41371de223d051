"""ZeRO-3 unit coordinator: gather / release / prefetch of flat units around module execution.

The reference walks *every* ``nn.Module`` with four hooks and a recorded trace
(``partitioned_param_coordinator.py:73-690``, ``parameter_offload.py:331-603``). Here the granule is the
:class:`FlatUnit` (a transformer block): one forward-pre / forward-post hook per unit module, a tensor hook on
the unit outputs for pre-backward, and the per-parameter grad hooks of the optimizer for post-backward. The
execution order recorded in the first iteration drives a one-unit-ahead prefetch on a side stream.

A module may use parameters owned by another unit (tied embeddings, "external parameters" in the reference,
``partition_parameters.py register_external_parameter``): ``_needed_units`` resolves that statically from
``module.parameters()`` plus anything registered through :func:`register_external_parameter`.
"""
import torch

from ... import comm as dist
from ...utils.nvtx import instrument_w_nvtx
from .units import FlatUnit


class UnitCoordinator:

    def __init__(self, zopt, prefetch=True, prefetch_bucket_size=None, max_live_parameters=None,
                 max_reuse_distance=None):
        self.z = zopt
        self.prefetch = prefetch
        # reference knobs (partitioned_param_coordinator.py:73-690): elements that may be gathered ahead of use, cap on
        # gathered elements alive at once, and "do not release what is needed again within this many elements"
        self.prefetch_bucket_size = int(prefetch_bucket_size) if prefetch_bucket_size else 0
        self.max_live_parameters = int(max_live_parameters) if max_live_parameters else 0
        self.max_reuse_distance = int(max_reuse_distance) if max_reuse_distance else 0
        self.gathers_launched = 0  # statistics (tests, tuning)
        self._plan = None  # profile-guided schedule (compile/planner.py); None = the heuristics below
        # reference aux subsystems: event counters (zero_optimization.log_trace_cache_warnings / wall_clock_breakdown)
        # and the cross-rank sanity checks of "enable_sanity_checks" (every rank must gather the same units in the same
        # order, or the collectives deadlock / corrupt)
        self.param_profiler = None
        self.sanity_checks = False
        self._gather_log = []
        self._prof = None  # active UnitProfiler (compile/profiler.py)
        self.handles = []
        self.fwd_order = []  # unit ids in first-forward order
        self._trace_done = False
        self._fwd_pos = {}
        self._in_backward_pass = False
        self._module_units = {}
        self._external = {}  # module -> [params]
        self._slot_buffers = {}
        self._install()

    # ---- hook installation ---------------------------------------------------------------------
    def _install(self):
        mods = {}
        for u in self.z.units:
            if u.module is not None and not u.persistent:
                mods.setdefault(u.module, []).append(u)
        self._own_units = mods
        for m in mods:
            self.handles.append(m.register_forward_pre_hook(self._pre_forward, with_kwargs=True))
            self.handles.append(m.register_forward_hook(self._post_forward))

    def remove_hooks(self):
        for h in self.handles:
            h.remove()
        self.handles = []

    def register_external_parameter(self, module, param):
        self._external.setdefault(module, []).append(param)
        self._module_units.pop(module, None)
        if module not in self._own_units:
            self._own_units[module] = []
            self.handles.append(module.register_forward_pre_hook(self._pre_forward, with_kwargs=True))
            self.handles.append(module.register_forward_hook(self._post_forward))

    def _needed_units(self, module):
        """Non-persistent units owning any parameter reachable from ``module`` that no child unit-module covers."""
        if module in self._module_units:
            return self._module_units[module]
        covered = set()
        for child in module.modules():
            if child is not module and child in self._own_units:
                for p in child.parameters():
                    covered.add(id(p))
        need, seen = [], set()
        plist = list(module.parameters()) + list(self._external.get(module, []))
        for p in plist:
            if id(p) in covered and p not in self._external.get(module, []):
                continue
            ent = self.z._unit_of_param.get(id(p))
            if ent is None:
                continue
            u = ent[0]
            if not u.persistent and u.uid not in seen:
                seen.add(u.uid)
                need.append(u)
        self._module_units[module] = need
        return need

    # ---- gather / release ----------------------------------------------------------------------
    @instrument_w_nvtx
    def gather(self, u: FlatUnit, wait=True):
        """Launch (if needed) and optionally complete the all-gather of a unit's parameters."""
        z = self.z
        if not u.gathered and u.gather_work is None:
            self.gathers_launched += 1
            if self.sanity_checks:
                self._gather_log.append(u.uid)
            if self.param_profiler is not None:
                from . import partitioned_param_profiler as pp
                bwd = self._in_backward_pass
                name = (pp.BACKWARD_FETCH_SUBMIT if bwd else pp.FORWARD_FETCH_SUBMIT) if wait else \
                    (pp.BACKWARD_PREFETCH_SUBMIT if bwd else pp.FORWARD_PREFETCH_SUBMIT)
                self.param_profiler.count(name, u.numel_padded)
                self.param_profiler.count(pp.BACKWARD_ALL_GATHER if bwd else pp.FORWARD_ALL_GATHER, u.numel_padded)
            use_side = z._use_streams and z.overlap_comm
            src = u.param_shard
            group = u.dp_group
            if z._hpz_group is not None and self._in_backward_pass and u.sec_shard is not None:
                src, group = u.sec_shard, z._hpz_group
            world = dist.get_world_size(group) if group is not None and dist.is_initialized() else 1
            ev = u.opt_event
            if ev is not None:  # this unit's slice of a pipelined optimizer step must have landed
                (z.ag_stream if (use_side and world > 1) else torch.cuda.current_stream()).wait_event(ev)
                u.opt_event = None
            if world <= 1 and src.numel() == u.numel_padded and src.dtype == u.dtype:
                # a single shard IS the full parameter: alias it instead of "gathering" through a copy
                u.full_param = src
                u.gather_work = True
                use_side = None
            prof = self._prof if use_side is not None else None
            if use_side is None:
                pass
            elif use_side:
                with torch.cuda.stream(z.ag_stream):
                    full = torch.empty(u.numel_padded, dtype=u.dtype, device=u.device)
                    tok = prof.gather_begin(u, z.ag_stream) if prof is not None else None
                    self._launch_gather(u, full, src, group)
                    if prof is not None:
                        prof.gather_end(tok, z.ag_stream)
                    u.gather_event = torch.cuda.Event()
                    u.gather_event.record(z.ag_stream)
            else:
                full = torch.empty(u.numel_padded, dtype=u.dtype, device=u.device)
                tok = prof.gather_begin(u, None) if prof is not None else None
                self._launch_gather(u, full, src, group)
                if prof is not None:
                    prof.gather_end(tok, None)
            if use_side is not None:
                u.full_param = full
                u.gather_work = True
        if wait and not u.gathered:
            if self.param_profiler is not None:
                from . import partitioned_param_profiler as pp
                self.param_profiler.count(pp.BACKWARD_FETCH_WAIT if self._in_backward_pass else pp.FORWARD_FETCH_WAIT,
                                          u.numel_padded)
            if u.gather_event is not None:
                torch.cuda.current_stream().wait_event(u.gather_event)
                u.full_param.record_stream(torch.cuda.current_stream())
                u.gather_event = None
            for p, view in zip(u.params, u.view_params(u.full_param)):
                p.data = view
                p.ds_status = "AVAILABLE"
            u.gathered = True
            u.gather_work = None
            if z.hpz_size > 1 and u.sec_shard is None and not self._in_backward_pass:
                self._make_secondary(u)

    def _launch_gather(self, u, full, src, group):
        world = dist.get_world_size(group) if group is not None and dist.is_initialized() else 1
        if world <= 1:
            full[:src.numel()].copy_(src)
            return
        st = self.z.storages.get((u.group_key, u.dtype))
        if self.z.symm is not None and st is not None and getattr(st, "lp_symm", None) is not None \
                and src.data_ptr() == u.param_shard.data_ptr() and not self.z.quantized_weights:
            # fused path 2 (v1): pull every peer's shard straight out of its symmetric lp storage over NVLink
            esz = src.element_size()
            if self.z._ag_on_copy_engines:
                self.z.symm.all_gather_ce(st.lp_symm, u.shard_offset * esz, u.shard_numel * esz, full,
                                          need_epoch=getattr(u, "_ready_epoch", 0))
            else:
                self.z.symm.all_gather(st.lp_symm, u.shard_offset * esz, u.shard_numel * esz, full)
            return
        if self.z.quantized_weights and self.z.symm is not None and u.dtype in (torch.bfloat16, torch.float16) \
                and src.data_ptr() == u.param_shard.data_ptr() and u.shard_numel % 256 == 0 \
                and group is u.dp_group:
            # fused path 3 (qwZ): quantize -> NVLink pull of FP8 + scales -> dequantize, one kernel
            self.z.symm.all_gather_fp8(src, self._fp8_stage(u), full)
            return
        if self.z.quantized_weights and u.dtype in (torch.bfloat16, torch.float16) and full.is_cuda:
            from ...ops.quantizer import quantized_all_gather
            quantized_all_gather(full, src, group)
            return
        if dist.get_backend(group) == "gloo":
            src = src.clone()
        if world * src.numel() == full.numel():
            dist.all_gather_into_tensor(full, src, group=group)
        else:  # hpZ secondary gather: group is a subgroup holding replicated sub-shards
            dist.all_gather_into_tensor(full, src, group=group)

    def _fp8_stage(self, u):
        """Round-robin symmetric staging buffers for the FP8 payload + scales (the kernel itself waits for the peers
        to be done with the buffer's previous use)."""
        pool = getattr(self, "_fp8_pool", None)
        if pool is None:
            nbytes = self.z.symm.fp8_stage_bytes(max(x.shard_numel for x in self.z.units), 256)
            pool = self._fp8_pool = [self.z.symm.alloc(nbytes, want_multicast=False) for _ in range(3)]
            self._fp8_next = 0
        stage = pool[self._fp8_next]
        self._fp8_next = (self._fp8_next + 1) % len(pool)
        return stage

    def _make_secondary(self, u):
        """hpZ: keep an intra-group replica so the backward re-gather stays inside the (smaller) group."""
        z = self.z
        hw = z.hpz_size
        n = u.numel_padded // hw
        r = dist.get_rank(z._hpz_group)
        u.sec_shard = u.full_param[r * n:(r + 1) * n].clone()

    def release(self, u: FlatUnit, force=False):
        if u.persistent or not (u.gathered or u.gather_work):
            return
        if getattr(u, "resident", False) and not force:
            return  # planned to stay gathered until the optimizer step rewrites the shards (on_step_done)
        if u.gather_work and not u.gathered:
            self.gather(u, wait=True)
        self.z._release_params(u)

    # ---- module hooks --------------------------------------------------------------------------
    def _pre_forward(self, module, args, kwargs):
        need = self._needed_units(module)
        for u in need:
            if not self._trace_done and u.uid not in self._fwd_pos:
                self._fwd_pos[u.uid] = len(self.fwd_order)
                self.fwd_order.append(u.uid)
            self.gather(u, wait=False)
        if self._prof is not None and need and not self._in_backward_pass:
            self._prof.step_begin("fwd", need)
        if self.prefetch and self._trace_done and need and not self._in_backward_pass:
            if self._plan is not None:
                self._launch_planned([self._fwd_pos.get(u.uid, -1) for u in need])
            else:
                pos = max(self._fwd_pos.get(u.uid, -1) for u in need) + 1
                self._prefetch_from(pos, +1)
        for u in need:
            self.gather(u, wait=True)
        return None

    # ---- profile-guided plan (DeepCompile counterpart, compile/planner.py) -----------------------
    def set_plan(self, plan):
        """``plan``: :class:`compile.planner.UnitPlan` over the timeline [forward positions, backward positions] of
        ``fwd_order``, or None to go back to the heuristics. Resident units stay gathered across micro-steps."""
        for u in self.z.units:
            u.resident = False
        self._plan = plan
        if plan is not None:
            for uid in plan.resident:
                self.z.units[uid].resident = True

    def _launch_planned(self, steps):
        n = len(self.fwd_order)
        for s in steps:
            for t in self._plan.launch_at.get(s, ()):
                pos = t if t < n else 2 * n - 1 - t
                u = self.z.units[self.fwd_order[pos]]
                if t >= n and u.reduced_this_pass:
                    continue
                self.gather(u, wait=False)

    def _live_numel(self):
        return sum(u.numel_padded for u in self.z.units if not u.persistent and (u.gathered or u.gather_work))

    def _prefetch_from(self, pos, step):
        """Launch gathers of the units that follow in execution order (``step`` +1 forward, -1 backward): always the
        next one, further ones while the elements in flight stay within ``stage3_prefetch_bucket_size`` and the live
        elements within ``stage3_max_live_parameters``."""
        budget, first = self.prefetch_bucket_size, True
        while 0 <= pos < len(self.fwd_order):
            nxt = self.z.units[self.fwd_order[pos]]
            if step < 0 and nxt.reduced_this_pass:
                break
            if not (nxt.gathered or nxt.gather_work):
                if not first:
                    if nxt.numel_padded > budget:
                        break
                    if self.max_live_parameters and self._live_numel() + nxt.numel_padded > self.max_live_parameters:
                        break
                self.gather(nxt, wait=False)
                budget -= nxt.numel_padded
            first = False
            pos += step

    def _keep_for_backward(self, u):
        """Forward is done with ``u``: skip the release when its backward use is closer than
        ``stage3_max_reuse_distance`` elements away (the units after it run forward and backward in between) and the
        live set stays under ``stage3_max_live_parameters`` (reference ``__params_to_release``)."""
        if self._plan is not None and self._trace_done and u.uid in self._fwd_pos:
            return torch.is_grad_enabled() and self._fwd_pos[u.uid] in self._plan.keep
        if not (self.max_reuse_distance and self._trace_done and torch.is_grad_enabled()) or u.uid not in self._fwd_pos:
            return False
        pos = self._fwd_pos[u.uid]
        dist = 2 * sum(self.z.units[uid].numel_padded for uid in self.fwd_order[pos + 1:])
        if dist > self.max_reuse_distance:
            return False
        return not self.max_live_parameters or self._live_numel() <= self.max_live_parameters

    def _post_forward(self, module, args, output):
        need = self._needed_units(module)
        grad_mode = torch.is_grad_enabled()
        if grad_mode:
            self._attach_pre_backward(module, output)
        if not self._in_backward_pass:
            for u in need:
                if self._still_needed_by_parent(module, u):
                    continue
                if grad_mode and self._keep_for_backward(u):
                    continue
                self.release(u)
        return None

    def _still_needed_by_parent(self, module, u):
        # units attached to a module that is an ancestor of the running one stay until their own post-forward
        return u.module is not None and u.module is not module and module in set(u.module.modules())

    def _attach_pre_backward(self, module, output):
        tensors = []
        _collect_tensors(output, tensors)
        tensors = [t for t in tensors if t.requires_grad]
        if not tensors:
            return
        fired = {"done": False}

        def hook(grad):
            if not fired["done"]:
                fired["done"] = True
                self._pre_backward(module)
            return grad

        for t in tensors:
            t.register_hook(hook)

    def _pre_backward(self, module):
        self._in_backward_pass = True
        self.z._ensure_epilogue()
        need = self._needed_units(module)
        for u in need:
            self.gather(u, wait=False)
        if self._prof is not None and need:
            self._prof.step_begin("bwd", need)
        if self.prefetch and need and self._plan is not None and self._trace_done:
            n = len(self.fwd_order)
            self._launch_planned([2 * n - 1 - self._fwd_pos[u.uid] for u in need if u.uid in self._fwd_pos])
        elif self.prefetch and need:
            pos = min(self._fwd_pos.get(u.uid, 1 << 30) for u in need) - 1
            self._prefetch_from(pos, -1)
        for u in need:
            self.gather(u, wait=True)

    # ---- optimizer callbacks -------------------------------------------------------------------
    def on_unit_backward_done(self, u):
        if not u.persistent:
            self.release(u)

    def on_backward_end(self):
        self._in_backward_pass = False
        if self._prof is not None:
            self._prof.pass_end()
        if self.fwd_order:
            self._trace_done = True
        for u in self.z.units:
            if not u.persistent and (u.gathered or u.gather_work):
                self.release(u)

    def check_gather_order(self):
        """``enable_sanity_checks``: the unit ids gathered since the last check must be identical on every rank."""
        from .utils import assert_ints_same_as_other_ranks
        log, self._gather_log = self._gather_log, []
        if dist.is_initialized() and dist.get_world_size() > 1:
            assert_ints_same_as_other_ranks(log)

    def on_step_done(self):
        if self.sanity_checks:
            self.check_gather_order()
        z = self.z
        if z._use_streams and z.overlap_comm:
            z.ag_stream.wait_stream(torch.cuda.current_stream())  # shards were rewritten by the step
        for u in self.z.units:
            u.sec_shard = None  # hpZ replicas are stale after the update (reference stage3.py:706)
        if self._plan is not None:
            # resident units: the gathered copy is stale now; refresh it once per optimizer step (in first-use
            # order, every rank the same), overlapping whatever runs before the next forward needs them
            for uid in self.fwd_order:
                u = self.z.units[uid]
                if getattr(u, "resident", False) and not u.persistent:
                    self.release(u, force=True)
                    self.gather(u, wait=False)

    def release_all(self):
        for u in self.z.units:
            if not u.persistent:
                self.release(u, force=True)

    def gather_all(self):
        for u in self.z.units:
            self.gather(u, wait=True)


def _collect_tensors(obj, out):
    if torch.is_tensor(obj):
        out.append(obj)
    elif isinstance(obj, (list, tuple)):
        for o in obj:
            _collect_tensors(o, out)
    elif isinstance(obj, dict):
        for o in obj.values():
            _collect_tensors(o, out)
    elif hasattr(obj, "__dict__") and not isinstance(obj, torch.nn.Module):
        for o in vars(obj).values():
            if torch.is_tensor(o) or isinstance(o, (list, tuple, dict)):
                _collect_tensors(o, out)
