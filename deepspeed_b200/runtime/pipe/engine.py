"""``PipelineEngine``: executes the 1F1B instruction stream of :mod:`.schedule` over a :class:`PipelineModule`.

Reference: ``runtime/pipe/engine.py`` (``train_batch`` :341, ``eval_batch`` :431, ``_exec_schedule`` :1396,
``_exec_forward_pass`` :722, ``_exec_backward_pass`` :815, send/recv activations+grads :1047-1139,
``_aggregate_total_loss`` :596, tied-weight + DP gradient reduction :278-296). ZeRO-0/1 only under PP (same rule as
the reference ``engine.py:2390``).
"""
import torch

from ... import comm as dist
from ...utils.logging import log_dist
from ..engine import DeepSpeedEngine
from . import p2p, schedule
from .module import PipelineModule


class PipelineEngine(DeepSpeedEngine):
    ID_TO_DTYPE = p2p._DTYPES

    def __init__(self, has_bool_tensors=False, *args, **kwargs):
        model = kwargs.get("model")
        assert isinstance(model, PipelineModule), "model must be a PipelineModule"
        self.grid = model.mpu()
        kwargs["mpu"] = self.grid
        super().__init__(*args, **kwargs)
        assert self.zero_optimization_stage() < 2, "ZeRO-2 and ZeRO-3 are incompatible with pipeline parallelism"
        self.enable_backward_allreduce = False
        self.has_bool_tensors = has_bool_tensors
        self.num_stages = self.grid.pipe_parallel_size
        self.stage_id = self.grid.get_stage_id()
        self.prev_stage, self.next_stage = self.stage_id - 1, self.stage_id + 1
        self.micro_batch_size = self.train_micro_batch_size_per_gpu()
        self.micro_batches = self.gradient_accumulation_steps()
        self.is_pipe_parallel = self.num_stages > 1
        self.data_iterator = None
        self.batch_fn = None
        self.pipe_buffers = {"inputs": [], "labels": [], "outputs": [], "output_tensors": []}
        self.num_pipe_buffers = 0
        self.pipe_recv_buf = None
        self.grad_layer = None
        self.meta_buffer = None
        self.first_output_send = True
        self.first_gradient_send = True
        self.loss = None
        self.total_loss = None
        self.agg_loss = torch.tensor(0.0, device=self.device)
        self.dp_group_loss = torch.tensor(0.0, device=self.device)
        p2p.init_process_groups(self.grid)
        self.loss_model = self.module.loss_fn
        self._force_grad_boundary = False
        self.log_batch_step_id = -1
        if self.optimizer is not None:
            # micro-batch gradients accumulate locally; the DP reduction happens in ReduceGrads
            self.optimizer._require_reduce = False
        log_dist(f"PipelineEngine: stage {self.stage_id}/{self.num_stages} micro_batches={self.micro_batches}",
                 ranks=[-1])

    # ---- public ----------------------------------------------------------------------------------
    def is_first_stage(self):
        return self.stage_id == 0

    def is_last_stage(self):
        return self.stage_id == self.num_stages - 1

    def is_gradient_accumulation_boundary(self):
        return self._force_grad_boundary

    def set_dataiterator(self, iterator):
        if self.is_first_stage() or self.is_last_stage():
            self.training_dataloader = None
            self.data_iterator = iterator

    def set_batch_fn(self, fn):
        self.batch_fn = fn

    def set_dataloader(self, loader):
        """Only the first and last stage read data (inputs / labels)."""
        if self.is_first_stage() or self.is_last_stage():
            self.training_dataloader = loader
            self.data_iterator = iter(self.training_dataloader)

    def set_has_attention_mask(self, value):
        assert isinstance(value, bool)
        self.has_attention_mask = value

    def get_pipeline_parallel_rank(self):
        return self.stage_id

    def get_additional_losses(self):
        """Extra losses a module reported through ``PipelineModule.get_additional_losses`` (aggregated per batch)."""
        return getattr(self, "agg_additional_losses", None)

    def log_for_device(self, *msg):
        print(f"RANK={dist.get_rank()} PIPE-ID={self.stage_id} DATA-ID={self.grid.data_parallel_id} "
              f"MBATCH-ID={getattr(self, 'microbatch_id', 0)} STEP-ID={getattr(self, 'log_batch_step_id', -1)} ::",
              *msg, flush=True)

    def tput_log(self, *msg):
        if self.global_rank == 0 and self.steps_per_print() and self.global_steps % self.steps_per_print() == 0:
            print(*msg)

    def reset_activation_shape(self):
        self.first_output_send = True
        self.pipe_recv_buf = None
        self.grad_layer = None

    def train_batch(self, data_iter=None):
        if not torch._C.is_grad_enabled():
            raise RuntimeError("train_batch() requires gradients enabled. Use eval_batch() instead.")
        if data_iter is not None:
            self.set_dataiterator(data_iter)
        self.module.train()
        self.total_loss = None
        self._compute_loss = True
        sched = schedule.TrainSchedule(micro_batches=self.micro_batches, stages=self.num_stages, stage_id=self.stage_id)
        self._exec_schedule(sched)
        self.agg_train_loss = self._aggregate_total_loss()
        if self.monitor.enabled and self.global_rank == 0:
            self.monitor.write_events([("Train/Samples/train_loss", float(self.agg_train_loss), self.global_samples)])
        return self.agg_train_loss

    def eval_batch(self, data_iter, return_logits=False, compute_loss=True, reduce_output="avg", bcast_loss=True,
                   num_micro_batches=None):
        self.module.eval()
        self._compute_loss = compute_loss
        self.total_loss = None
        self.set_dataiterator(data_iter)
        mb = self.micro_batches if num_micro_batches is None else num_micro_batches
        sched = schedule.InferenceSchedule(micro_batches=mb, stages=self.num_stages, stage_id=self.stage_id)
        dist.barrier()
        with torch.no_grad():
            self._exec_schedule(sched)
        out = None
        if self.is_last_stage():
            out = self._reduce_outputs(self.fwd_outputs, reduce_output, mb)
        if compute_loss and bcast_loss:
            out = self._bcast_pipe_scalar(out)
        self.fwd_outputs = []
        return out

    def forward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def backward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def step(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    # ---- helpers ---------------------------------------------------------------------------------
    def _reduce_outputs(self, outputs, reduce="avg", micro_batches=None):
        if reduce is None or not outputs:
            return outputs
        if torch.is_tensor(outputs[0]):
            return sum(outputs) / len(outputs) if reduce == "avg" else sum(outputs)
        return outputs

    def _bcast_pipe_scalar(self, data, src_rank=None, dtype=torch.float32):
        if src_rank is None:
            src_rank = self.grid.stage_to_global(self.num_stages - 1)
        t = data.clone().detach().to(dtype).to(self.device) if self.global_rank == src_rank else \
            torch.zeros((), dtype=dtype, device=self.device)
        if self.num_stages > 1:
            dist.broadcast(t, src=src_rank, group=self.grid.get_pipe_parallel_group())
        return t

    def _aggregate_total_loss(self):
        if self.is_last_stage():
            loss = self._scale_loss_by_gas(self.total_loss) if self.total_loss is not None else torch.zeros((), device=self.device)
            self.dp_group_loss = loss.clone().detach()
            agg = self.dp_group_loss.clone().float()
            if self.grid.data_parallel_size > 1:
                dist.all_reduce(agg, group=self.grid.get_data_parallel_group())
                agg /= self.grid.data_parallel_size
            losses = torch.stack([self.dp_group_loss.float(), agg])
        else:
            losses = torch.zeros(2, device=self.device)
        if self.num_stages > 1:
            dist.broadcast(losses, src=self.grid.stage_to_global(self.num_stages - 1),
                           group=self.grid.get_pipe_parallel_group())
        self.dp_group_loss = losses[0].clone()
        return losses[1].clone()

    def _next_batch(self):
        batch = next(self.data_iterator)
        if self.batch_fn:
            batch = self.batch_fn(batch)
        return batch

    def _to_dev(self, x):
        if torch.is_tensor(x):
            return x.to(self.device)
        if isinstance(x, (tuple, list)):
            return type(x)(self._to_dev(t) for t in x)
        return x

    def _reserve_pipe_buffers(self, n):
        if self.num_pipe_buffers >= n:
            return
        add = n - self.num_pipe_buffers
        for k in self.pipe_buffers:
            self.pipe_buffers[k].extend([None] * add)
        self.num_pipe_buffers = n

    # ---- instruction handlers ----------------------------------------------------------------------
    def _exec_load_micro_batch(self, buffer_id):
        batch = self._next_batch()
        if self.is_first_stage():
            inp = self._to_dev(batch[0])
            if torch.is_tensor(inp) and inp.is_floating_point():
                inp = inp.clone().detach().requires_grad_(True)
            self.pipe_buffers["inputs"][buffer_id] = inp
        if self.is_last_stage():
            self.pipe_buffers["labels"][buffer_id] = self._to_dev(batch[1])

    def _exec_forward_pass(self, buffer_id):
        self.tput_timer.start()
        inputs = self.pipe_buffers["inputs"][buffer_id]
        outputs = self.module(inputs)
        self.pipe_buffers["outputs"][buffer_id] = outputs
        if self.is_last_stage():
            if self._compute_loss and self.loss_model is not None:
                labels = self.pipe_buffers["labels"][buffer_id]
                self.loss = self.loss_model(outputs, labels)
            else:
                self.loss = outputs
            if torch.is_tensor(self.loss):
                self.fwd_outputs = getattr(self, "fwd_outputs", [])
                self.fwd_outputs.append(self.loss.detach())
                self.total_loss = self.loss.detach().clone() if self.total_loss is None else self.total_loss + self.loss.detach()

    def _exec_backward_pass(self, buffer_id):
        assert self.optimizer is not None
        if self.is_last_stage():
            loss = self._scale_loss_by_gas(self.loss)
            self.optimizer.backward_prologue()
            self.optimizer.scale_if_loss(loss).backward()
            self.optimizer.backward_epilogue()
        else:
            outputs = self.pipe_buffers["outputs"][buffer_id]
            grads = self.grad_layer
            self.optimizer.backward_prologue()
            if isinstance(outputs, tuple):
                outs = [t for t in outputs if t.is_floating_point()]
                torch.autograd.backward(tensors=outs, grad_tensors=list(grads))
            else:
                torch.autograd.backward(tensors=(outputs, ), grad_tensors=(grads, ))
            self.optimizer.backward_epilogue()
        self.pipe_buffers["outputs"][buffer_id] = None

    def _exec_send_activations(self, buffer_id):
        outputs = self.pipe_buffers["outputs"][buffer_id]
        if self.first_output_send:
            self.first_output_send = False
            p2p.send_meta(outputs, self.next_stage)
        for t in ((outputs, ) if torch.is_tensor(outputs) else outputs):
            p2p.send(t.detach(), self.next_stage)

    def _exec_recv_activations(self, buffer_id):
        if self.pipe_recv_buf is None:
            self._recv_is_tuple, self._recv_specs = p2p.recv_meta(self.prev_stage)
            self.pipe_recv_buf = True
        bufs = []
        for dt, shape, rg in self._recv_specs:
            b = torch.empty(shape, dtype=dt, device=self.device)
            p2p.recv(b, self.prev_stage)
            if b.is_floating_point():
                b.requires_grad_(True)
            bufs.append(b)
        self.pipe_buffers["inputs"][buffer_id] = tuple(bufs) if self._recv_is_tuple else bufs[0]

    def _exec_send_grads(self, buffer_id):
        inputs = self.pipe_buffers["inputs"][buffer_id]
        for t in ((inputs, ) if torch.is_tensor(inputs) else inputs):
            if t.is_floating_point():
                g = t.grad if t.grad is not None else torch.zeros_like(t)
                p2p.send(g, self.prev_stage)
        self.pipe_buffers["inputs"][buffer_id] = None

    def _exec_recv_grads(self, buffer_id):
        outputs = self.pipe_buffers["outputs"][buffer_id]
        outs = (outputs, ) if torch.is_tensor(outputs) else tuple(t for t in outputs if t.is_floating_point())
        grads = []
        for t in outs:
            g = torch.empty_like(t)
            p2p.recv(g, self.next_stage)
            grads.append(g)
        self.grad_layer = grads[0] if torch.is_tensor(outputs) else tuple(grads)

    def _exec_reduce_tied_grads(self):
        # tied-weight grads live in the ZeRO flat buffers; reduce the param-level views across the tie group
        for weight, group in self.module.get_tied_weights_and_groups():
            ent = self.optimizer._unit_of_param.get(id(weight))
            if ent is None:
                continue
            u, i = ent
            if u.full_grad is not None:
                view = u.full_grad[u.offsets[i]:u.offsets[i] + u.numels[i]]
                dist.all_reduce(view, group=group)

    def _exec_reduce_grads(self):
        self._force_grad_boundary = True
        self.optimizer.is_gradient_accumulation_boundary = True
        self.optimizer._require_reduce = True
        self.optimizer.reduce_gradients(pipeline_parallel=True)
        self.optimizer._require_reduce = False

    def _exec_optimizer_step(self, lr_kwargs=None):
        self._force_grad_boundary = True
        self._take_model_step(lr_kwargs)
        self._force_grad_boundary = False
        self.micro_steps += self.micro_batches
        self.tput_timer.stop(global_step=True)

    _INSTRUCTION_MAP = {
        schedule.OptimizerStep: _exec_optimizer_step, schedule.ReduceGrads: _exec_reduce_grads,
        schedule.ReduceTiedGrads: _exec_reduce_tied_grads, schedule.LoadMicroBatch: _exec_load_micro_batch,
        schedule.ForwardPass: _exec_forward_pass, schedule.BackwardPass: _exec_backward_pass,
        schedule.SendActivation: _exec_send_activations, schedule.RecvActivation: _exec_recv_activations,
        schedule.SendGrad: _exec_send_grads, schedule.RecvGrad: _exec_recv_grads,
    }

    def _exec_schedule(self, pipe_schedule):
        self._reserve_pipe_buffers(pipe_schedule.num_pipe_buffers())
        self.fwd_outputs = []
        for step_cmds in pipe_schedule:
            for cmd in step_cmds:
                handler = self._INSTRUCTION_MAP.get(type(cmd))
                if handler is None:
                    raise RuntimeError(f"{self.__class__.__name__} does not understand instruction {repr(cmd)}")
                handler(self, **cmd.kwargs)

    # ---- checkpoints: per-layer files (reference pipe/module.py:572-643) ---------------------------
    def module_state_dict(self, exclude_frozen_parameters=False, **kw):
        assert isinstance(self.module, PipelineModule)
        self.module.checkpoint_parallel_write_pipeline = self._config.checkpoint_parallel_write_pipeline
        if getattr(self, "_curr_ckpt_path", None) is not None:
            self.module.save_state_dict(self._curr_ckpt_path, checkpoint_engine=self.checkpoint_engine,
                                        exclude_frozen_params=exclude_frozen_parameters)
            return None
        return super().module_state_dict(exclude_frozen_parameters=exclude_frozen_parameters)

    def load_module_state_dict(self, checkpoint, strict=True, custom_load_fn=None, fetch_z3_params=False):
        if checkpoint is not None and not isinstance(checkpoint, str) and checkpoint.get("module") is not None:
            return super().load_module_state_dict(checkpoint, strict, custom_load_fn)
        if getattr(self, "_curr_ckpt_path", None) is not None:
            self.module.load_state_dir(load_dir=self._curr_ckpt_path, strict=strict,
                                       checkpoint_engine=self.checkpoint_engine)


class PipelineError(Exception):
    pass
