"""Sub-allocator over one pre-allocated flat buffer with compaction (reference
``runtime/zero/contiguous_memory_allocator.py:16``): used for activation-checkpoint and parameter buffers that must
not fragment the CUDA caching allocator. Tensors are views into the buffer; ``release`` frees a range; when no hole
is large enough but total free space suffices the live tensors are compacted (their data moved, views/params re-pointed)."""
import torch


class ContiguousMemoryAllocator:

    def __init__(self, size, dtype, device):
        self.buffer = torch.zeros(size, dtype=dtype, device=device)
        self.size = size
        self.dtype = dtype
        self.free = {0: size}            # address -> length of every hole
        self.live = {}                   # tensor id -> (address, size)
        self.tensors = {}                # tensor id -> view
        self.param_of = {}               # tensor id -> {param id: (param, numel, shape)}
        self.total_free = size
        self.largest = size
        self.max_allocated = 0
        self._count = 0

    # ---- public API -----------------------------------------------------------------------------
    def allocate_tensor(self, size):
        assert size <= self.total_free, "Not enough memory in buffer. Allocation failed"
        if self.largest < size:
            self._defragment()
        addr = next(a for a, n in sorted(self.free.items()) if n >= size)
        hole = self.free.pop(addr)
        if hole > size:
            self.free[addr + size] = hole - size
        view = self.buffer.narrow(0, addr, size)
        tid = id(view)
        self.live[tid], self.tensors[tid], self.param_of[tid] = (addr, size), view, {}
        self.total_free -= size
        self.max_allocated = max(self.max_allocated, self.size - self.total_free)
        self.largest = max(self.free.values(), default=0)
        return view

    def assign_to_param(self, tensor, param, numel, shape):
        tid = id(tensor)
        assert tid in self.live, "No such tensor allocated by the allocator."
        assert tensor.numel() >= numel, "Assert tensor buffer does is not large enough"
        self.param_of[tid][id(param)] = (param, numel, shape)
        param.data = tensor.narrow(0, 0, numel).view(shape)

    def release_tensor(self, tensor):
        self.release_tensor_with_id(id(tensor))

    def release_tensor_with_id(self, tensor_id):
        assert tensor_id in self.live, "Invalid tensor id"
        addr, size = self.live.pop(tensor_id)
        self.tensors.pop(tensor_id)
        for param, _, _ in self.param_of.pop(tensor_id).values():
            param.data = torch.empty(0, dtype=self.dtype, device=self.buffer.device)
        self.total_free += size
        self._free_range(addr, size)

    def print_allocation(self, resolution=200):
        cells = ["." for _ in range(resolution)]
        for addr, size in self.live.values():
            for c in range(int(addr * resolution / self.size), int((addr + size) * resolution / self.size) + 1):
                if c < resolution:
                    cells[c] = "x"
        print("".join(cells))

    # ---- internals ------------------------------------------------------------------------------
    def _free_range(self, addr, size):
        nxt = addr + size
        if nxt in self.free:
            size += self.free.pop(nxt)
        for a, n in list(self.free.items()):
            if a + n == addr:
                self.free.pop(a)
                addr, size = a, n + size
                break
        self.free[addr] = size
        self.largest = max(self.free.values())

    def _defragment(self):
        """Slide every live tensor towards address 0 (in address order), re-pointing views and bound params."""
        cursor = 0
        for tid, (addr, size) in sorted(self.live.items(), key=lambda kv: kv[1][0]):
            if addr != cursor:
                dst = self.buffer.narrow(0, cursor, size)
                src = self.buffer.narrow(0, addr, size)
                if cursor + size > addr:  # overlapping move
                    dst.copy_(src.clone())
                else:
                    dst.copy_(src)
                self.tensors[tid].data = dst
                self.live[tid] = (cursor, size)
                for param, numel, shape in self.param_of[tid].values():
                    param.data = dst.narrow(0, 0, numel).view(shape)
            cursor += size
        self.free = {cursor: self.size - cursor} if cursor < self.size else {}
        self.largest = self.size - cursor
