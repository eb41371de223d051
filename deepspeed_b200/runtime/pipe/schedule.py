"""Pipeline instruction streams: 1F1B training, forward-only inference, plain data-parallel.

Behaviour matches the reference schedules (``runtime/pipe/schedule.py``: ``TrainSchedule`` :189-298 emits
``2*(M+S-1)`` steps where a stage runs a forward when step and stage have the same parity and a backward
otherwise; buffers ``max(2, min(S - stage, M))`` :247-256; ``InferenceSchedule`` :131). Implementation here is
table-driven: :func:`one_f_one_b_slot` is the closed form for "which micro-batch, which direction" and the
instruction list for a step is assembled from that slot and the previous one.
"""
from dataclasses import dataclass


class PipeInstruction:
    """An instruction is its class name plus keyword operands (compared by value)."""

    def __init__(self, **operands):
        self.name = type(self).__name__
        self.kwargs = operands
        self.__dict__.update(operands)

    def __repr__(self):
        inner = ", ".join(f"{k}={v!r}" for k, v in self.kwargs.items())
        return f"{self.name}({inner})"

    def __eq__(self, other):
        return type(other) is type(self) and other.kwargs == self.kwargs

    def __hash__(self):
        return hash((self.name, tuple(sorted(self.kwargs.items()))))


class OptimizerStep(PipeInstruction):
    pass


class ReduceGrads(PipeInstruction):
    pass


class ReduceTiedGrads(PipeInstruction):
    pass


class BufferOpInstruction(PipeInstruction):

    def __init__(self, buffer_id, **operands):
        super().__init__(buffer_id=buffer_id, **operands)


LoadMicroBatch = type("LoadMicroBatch", (BufferOpInstruction, ), {})
ForwardPass = type("ForwardPass", (BufferOpInstruction, ), {})
BackwardPass = type("BackwardPass", (BufferOpInstruction, ), {})
SendActivation = type("SendActivation", (BufferOpInstruction, ), {})
RecvActivation = type("RecvActivation", (BufferOpInstruction, ), {})
SendGrad = type("SendGrad", (BufferOpInstruction, ), {})
RecvGrad = type("RecvGrad", (BufferOpInstruction, ), {})


@dataclass(frozen=True)
class Slot:
    micro_batch: int
    forward: bool


def one_f_one_b_slot(step, stage, stages):
    """Closed form of the 1F1B interleave.

    Forward slots: stage ``s`` runs micro-batch ``m`` forward at step ``2m + s`` (same parity as ``s``).
    Backward slots: at step ``2m + 2*stages - 1 - s`` (opposite parity).
    """
    if (step - stage) % 2 == 0:
        return Slot((step - stage) // 2, True)
    return Slot((step - (2 * stages - 1 - stage)) // 2, False)


class PipeSchedule:

    def __init__(self, micro_batches, stages, stage_id):
        self.micro_batches, self.stages, self.stage_id = micro_batches, stages, stage_id
        self.prev_stage, self.next_stage = stage_id - 1, stage_id + 1

    # subclasses yield a list of instructions per step
    def steps(self):
        raise NotImplementedError

    def num_pipe_buffers(self):
        return self.micro_batches

    def _valid_micro_batch(self, mb):
        return 0 <= mb < self.micro_batches

    def _valid_stage(self, s):
        return 0 <= s < self.stages

    def _buffer_idx(self, mb):
        assert self._valid_micro_batch(mb)
        return mb % self.num_pipe_buffers()

    stage = property(lambda self: self.stage_id)
    num_stages = property(lambda self: self.stages)
    num_micro_batches = property(lambda self: self.micro_batches)
    is_first_stage = property(lambda self: self.stage_id == 0)
    is_last_stage = property(lambda self: self.stage_id == self.stages - 1)

    def __iter__(self):
        self._it = self.steps()
        return self

    def __next__(self):
        return next(self._it)


class TrainSchedule(PipeSchedule):

    def num_pipe_buffers(self):
        return max(2, min(self.stages - self.stage_id, self.micro_batches))

    def _step_to_micro_batch(self, step_id):
        slot = one_f_one_b_slot(step_id, self.stage_id, self.stages)
        return slot.micro_batch, slot.forward

    def steps(self):
        n_steps = 2 * (self.micro_batches + self.stages - 1)
        has_prev, has_next = self._valid_stage(self.prev_stage), self._valid_stage(self.next_stage)
        edge = self.is_first_stage or self.is_last_stage
        last = Slot(-1, False)
        for step in range(n_steps):
            cur = one_f_one_b_slot(step, self.stage_id, self.stages)
            live, prev_live = self._valid_micro_batch(cur.micro_batch), self._valid_micro_batch(last.micro_batch)
            buf = self._buffer_idx(cur.micro_batch) if live else None
            pbuf = self._buffer_idx(last.micro_batch) if prev_live else None
            out = []
            if cur.forward:
                # the previous slot was a backward: ship its input-grad upstream, then receive new activations
                if prev_live and has_prev:
                    out.append(SendGrad(pbuf))
                if live and has_prev:
                    out.append(RecvActivation(buf))
                if live and edge:
                    out.append(LoadMicroBatch(buf))
                if live:
                    out.append(ForwardPass(buf))
            else:
                if live and has_next:
                    out.append(RecvGrad(buf))
                if prev_live and has_next:
                    out.append(SendActivation(pbuf))
                if live:
                    out.append(BackwardPass(buf))
            if step == n_steps - 1:
                out += [ReduceTiedGrads(), ReduceGrads(), OptimizerStep()]
            last = cur
            yield out


class InferenceSchedule(PipeSchedule):
    """Forward only; two buffers alternate so a send of micro-batch m-1 overlaps the receive of m."""

    def num_pipe_buffers(self):
        return 2

    def steps(self):
        has_prev, has_next = self._valid_stage(self.prev_stage), self._valid_stage(self.next_stage)
        for step in range(self.micro_batches + self.stages - 1):
            mb = step - self.stage_id
            recv_buf = (step + self.stage_id) % 2
            send_buf = 1 - recv_buf
            out = []
            if (self.is_first_stage or self.is_last_stage) and self._valid_micro_batch(mb):
                out.append(LoadMicroBatch(recv_buf))
            comm = []
            if has_next and self._valid_micro_batch(mb - 1):
                comm.append(SendActivation(send_buf))
            if has_prev and self._valid_micro_batch(mb):
                comm.append(RecvActivation(recv_buf))
            # even stages send first, odd stages receive first (deadlock-free pairing of blocking p2p)
            out += comm if self.stage_id % 2 == 0 else comm[::-1]
            if self._valid_micro_batch(mb):
                out.append(ForwardPass(recv_buf))
            yield out


class DataParallelSchedule(PipeSchedule):

    def num_pipe_buffers(self):
        return 1

    def steps(self):
        for mb in range(self.micro_batches):
            out = [LoadMicroBatch(buffer_id=0), ForwardPass(buffer_id=0), BackwardPass(buffer_id=0)]
            if mb == self.micro_batches - 1:
                out += [ReduceGrads(), OptimizerStep()]
            yield out
