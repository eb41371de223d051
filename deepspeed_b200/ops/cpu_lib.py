"""ctypes loader of the host runtime library ``libdsb200_cpu.so`` (AVX-512/OpenMP optimizers, async file I/O)."""
import ctypes
import os
import threading

_LIB = None
_LOCK = threading.Lock()
LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libdsb200_cpu.so")


def lib():
    global _LIB
    if _LIB is None:
        with _LOCK:
            if _LIB is None:
                if not os.path.exists(LIB_PATH):
                    from .builder import build_cpu_library
                    build_cpu_library(verbose=False)
                L = ctypes.CDLL(LIB_PATH)
                P, I, L64, F = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float
                L.dsb200_cpu_adam.argtypes = [P, P, I, P, P, P, I, L64, F, F, F, F, F, I, I, I, F]
                L.dsb200_cpu_adam.restype = None
                L.dsb200_cpu_adam_rollback.argtypes = [P, P, I, P, P, L64, F, F, F, F, F, I, I, I, F]
                L.dsb200_cpu_adam_rollback.restype = None
                L.dsb200_cpu_adagrad.argtypes = [P, P, I, P, P, I, L64, F, F, F, F]
                L.dsb200_cpu_adagrad.restype = None
                L.dsb200_cpu_lion.argtypes = [P, P, I, P, P, I, L64, F, F, F, F, F]
                L.dsb200_cpu_lion.restype = None
                L.dsb200_cpu_cast.argtypes = [P, P, I, L64]
                L.dsb200_cpu_cast.restype = None
                L.dsb200_cpu_sumsq.argtypes = [P, L64]
                L.dsb200_cpu_sumsq.restype = ctypes.c_double
                L.dsb200_cpu_num_threads.restype = I
                if hasattr(L, "dsb200_ragged_create"):
                    L.dsb200_ragged_create.argtypes = [I, I, I]
                    L.dsb200_ragged_create.restype = P
                    L.dsb200_ragged_destroy.argtypes = [P]
                    L.dsb200_ragged_free_blocks.argtypes = [P]
                    L.dsb200_ragged_free_blocks.restype = I
                    L.dsb200_ragged_tracked.argtypes = [P]
                    L.dsb200_ragged_tracked.restype = I
                    L.dsb200_ragged_seen.argtypes = [P, L64]
                    L.dsb200_ragged_seen.restype = L64
                    L.dsb200_ragged_blocks_needed.argtypes = [P, L64, I]
                    L.dsb200_ragged_blocks_needed.restype = I
                    L.dsb200_ragged_blocks.argtypes = [P, L64, P, I]
                    L.dsb200_ragged_blocks.restype = I
                    L.dsb200_ragged_flush.argtypes = [P, L64]
                    L.dsb200_ragged_flush.restype = I
                    L.dsb200_ragged_schedule.argtypes = [P, I, P, P, I, P, L64, P, I]
                    L.dsb200_ragged_schedule.restype = I
                if hasattr(L, "dsb200_dc_plan"):
                    D = ctypes.c_double
                    L.dsb200_dc_fit_comm.argtypes = [I, P, P, P]
                    L.dsb200_dc_fit_comm.restype = I
                    L.dsb200_dc_plan.argtypes = [I, P, P, I, P, P, P, D, D, L64, L64, L64, D, D, P, P, P, P]
                    L.dsb200_dc_plan.restype = I
                if hasattr(L, "dsb200_locked_alloc"):
                    L.dsb200_locked_alloc.argtypes = [L64, P]
                    L.dsb200_locked_alloc.restype = P
                    L.dsb200_locked_free.argtypes = [P]
                    L.dsb200_locked_free.restype = I
                    L.dsb200_locked_bytes.restype = L64
                if hasattr(L, "dsb200_shm_create"):
                    L.dsb200_shm_create.argtypes = [ctypes.c_char_p, I, I, L64, I, ctypes.c_double]
                    L.dsb200_shm_create.restype = P
                    L.dsb200_shm_destroy.argtypes = [P]
                    L.dsb200_shm_all_reduce.argtypes = [P, P, L64, I]
                    L.dsb200_shm_all_reduce.restype = I
                L.dsb200_aio_create.argtypes = [L64, I, I, I]
                L.dsb200_aio_create.restype = P
                L.dsb200_aio_destroy.argtypes = [P]
                L.dsb200_aio_pread.argtypes = [P, P, L64, ctypes.c_char_p, L64, I]
                L.dsb200_aio_pread.restype = I
                L.dsb200_aio_pwrite.argtypes = [P, P, L64, ctypes.c_char_p, L64, I]
                L.dsb200_aio_pwrite.restype = I
                L.dsb200_aio_wait.argtypes = [P]
                L.dsb200_aio_wait.restype = L64
                L.dsb200_aligned_alloc.argtypes = [L64, L64]
                L.dsb200_aligned_alloc.restype = P
                L.dsb200_aligned_free.argtypes = [P]
                _LIB = L
    return _LIB
