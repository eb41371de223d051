"""ZenFlow: stall-free optimizer offload (reference ``runtime/zenflow/zenflow_stage_1_and_2.py:47``,
``engine_stage3.py``, ``zenflow_utils.py``).

Idea: most of the gradient energy sits in a small share of the coordinates. Those *important* coordinates are updated
on the GPU every step by a small resident optimizer; the rest of the gradient is accumulated and handed to the CPU
optimizer only every ``update_interval`` steps, asynchronously, so the GPU never waits for the host.

Adaptation to the flat-unit design: the reference selects whole *columns* of each 2-D weight (all ranks agree through
an all-reduce of column norms, and the selected columns are updated redundantly). Our optimizer state is a flat shard
per rank, so importance is measured per 256-element *block of the rank's shard* and the selected blocks are updated
by the owning rank only - no extra collective and no redundancy."""
import threading

import torch

BLOCK = 256


class ZenFlowController:

    def __init__(self, zopt, cfg, total_steps_hint=None):
        self.z = zopt
        self.topk_ratio = cfg.topk_ratio
        self.select_interval = 1 if cfg.select_interval == "auto" else int(cfg.select_interval)
        self.update_interval = 4 if cfg.update_interval == "auto" else int(cfg.update_interval)
        self.overlap_step = bool(cfg.overlap_step)
        self.full_warm_up_rounds = int(cfg.full_warm_up_rounds)
        self.step_no = 0
        self.per_storage = {}
        self._thread = None
        for key, st in zopt.storages.items():
            nblk = (st.numel + BLOCK - 1) // BLOCK
            k = max(1, int(nblk * self.topk_ratio))
            dev = st.device
            self.per_storage[key] = dict(
                nblk=nblk, k=k, idx=None,
                p=torch.zeros(k * BLOCK, dtype=torch.float32, device=dev),      # GPU copies of the selected blocks
                m=torch.zeros(k * BLOCK, dtype=torch.float32, device=dev),
                v=torch.zeros(k * BLOCK, dtype=torch.float32, device=dev),
                steps=0,
                acc=torch.zeros(st.numel, dtype=torch.float32, device="cpu",
                                pin_memory=torch.cuda.is_available()),            # unimportant gradient, accumulated
                acc_steps=0)

    # ---- selection ------------------------------------------------------------------------------
    def _element_index(self, blocks, numel):
        idx = (blocks.view(-1, 1) * BLOCK + torch.arange(BLOCK, device=blocks.device)).view(-1)
        return idx.clamp_(max=numel - 1)

    def _select(self, key, st, grad):
        ps = self.per_storage[key]
        pad = ps["nblk"] * BLOCK - st.numel
        g = grad.float()
        if pad:
            g = torch.cat([g, g.new_zeros(pad)])
        energy = g.view(ps["nblk"], BLOCK).pow(2).sum(1)
        blocks = energy.topk(ps["k"]).indices.sort().values
        new_idx = self._element_index(blocks, st.numel)
        if ps["idx"] is not None:
            # write the GPU copies of the previously selected blocks back before re-selecting
            self._flush_selected(key, st)
        ps["idx"] = new_idx
        master = st.master if st.master.device == new_idx.device else st.master.to(new_idx.device)
        ps["p"].copy_(master[new_idx])
        ps["m"].zero_()
        ps["v"].zero_()
        ps["steps"] = 0

    def _flush_selected(self, key, st):
        ps = self.per_storage[key]
        idx = ps["idx"]
        st.master[idx.to(st.master.device)] = ps["p"].to(st.master.device)

    # ---- the step -------------------------------------------------------------------------------
    @torch.no_grad()
    def step(self, grad_scale):
        """Called instead of the plain offloaded optimizer step. Returns True when the (slow) CPU update ran/was queued."""
        z, opt = self.z, self.z.optimizer
        self.step_no += 1
        warm = self.step_no <= self.full_warm_up_rounds
        group0 = opt.param_groups[0]
        lr, (b1, b2), eps, wd = group0["lr"], group0["betas"], group0["eps"], group0["weight_decay"]
        self._join()
        for key, st in z.storages.items():
            ps = self.per_storage[key]
            grad = st.grad.float() * grad_scale if grad_scale != 1.0 else st.grad.float()
            if warm:
                ps["acc"].add_(grad.cpu())
                ps["acc_steps"] += 1
                continue
            if ps["idx"] is None or (self.step_no - 1) % self.select_interval == 0:
                self._select(key, st, grad)
            idx = ps["idx"]
            # important coordinates: AdamW on the GPU, every step
            g_sel = grad[idx]
            ps["steps"] += 1
            ps["m"].mul_(b1).add_(g_sel, alpha=1 - b1)
            ps["v"].mul_(b2).addcmul_(g_sel, g_sel, value=1 - b2)
            bc1, bc2 = 1 - b1**ps["steps"], 1 - b2**ps["steps"]
            upd = (ps["m"] / bc1) / ((ps["v"] / bc2).sqrt() + eps)
            ps["p"].mul_(1 - lr * wd).add_(upd, alpha=-lr)
            self._write_lp(st, idx, ps["p"])
            # everything else: accumulate for the host
            rest = grad.clone()
            rest[idx] = 0
            ps["acc"].add_(rest.cpu())
            ps["acc_steps"] += 1
        if warm or self.step_no % self.update_interval == 0:
            self._cpu_update()
            return True
        return False

    def _write_lp(self, st, idx, values):
        z = self.z
        if z.partition_params:
            st.lp[idx] = values.to(st.lp.dtype)
        else:
            for u in st.units:
                a, b = u.shard_offset, u.shard_offset + u.shard_numel
                m = (idx >= a) & (idx < b)
                if m.any():
                    u.param_shard[(idx[m] - a)] = values[m].to(u.param_shard.dtype)

    def _cpu_update(self):
        """CPU AdamW over the accumulated (mean) gradient of the unimportant coordinates; the selected coordinates are
        masked (their gradient there is zero and weight decay must not be applied twice)."""
        z, opt = self.z, self.z.optimizer

        def work():
            grads, lp = {}, {}
            for (key, a, b, gi, u, v) in z._segments:
                ps = self.per_storage[key]
                st = z.storages[key]
                g = ps["acc"][a:b]
                if ps["acc_steps"] > 1:
                    g.div_(ps["acc_steps"])
                grads[v] = g
                lp[v] = st.lp_host[a:b]
            opt.step_flat(grads=grads, lp_params=lp, grad_scale=1.0)
            for key, st in z.storages.items():
                ps = self.per_storage[key]
                ps["acc"].zero_()
                ps["acc_steps"] = 0
                if ps["idx"] is not None:  # selected coordinates are owned by the GPU optimizer
                    hidx = ps["idx"].cpu()
                    st.master[hidx] = ps["p"].cpu()
                    st.lp_host[hidx] = ps["p"].cpu().to(st.lp_host.dtype)
                if z.partition_params:
                    st.lp.copy_(st.lp_host, non_blocking=True)
                else:
                    for u in st.units:
                        u.param_shard.copy_(st.lp_host[u.shard_offset:u.shard_offset + u.shard_numel],
                                            non_blocking=True)

        if self.overlap_step:
            self._thread = threading.Thread(target=work, name="zenflow-cpu-adam", daemon=True)
            self._thread.start()
        else:
            work()

    def _join(self):
        if self._thread is not None:
            self._thread.join()
            self._thread = None

    def finalize(self):
        self._join()
        for key, st in self.z.storages.items():
            if self.per_storage[key]["idx"] is not None:
                self._flush_selected(key, st)
