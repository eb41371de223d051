"""Running total of pinned host memory requested through the framework (reference ``utils/pin_memory_tracker.py``)."""
from .logging import logger


def _fmt_bytes(n):
    for unit in ("B", "KiB", "MiB", "GiB", "TiB"):
        if abs(n) < 1024 or unit == "TiB":
            return f"{n:.2f} {unit}" if unit != "B" else f"{int(n)} B"
        n /= 1024.0


class _PinnedMemoryTracker:

    def __init__(self):
        self.reset()

    def reset(self):
        self.total_bytes = 0
        self.num_allocs = 0
        self.max_alloc = 0

    def track(self, num_bytes):
        self.total_bytes += int(num_bytes)
        self.num_allocs += 1
        self.max_alloc = max(self.max_alloc, int(num_bytes))

    def log_summary(self, tag=""):
        logger.info(f"[pinned memory{' ' + tag if tag else ''}] total={_fmt_bytes(self.total_bytes)} "
                    f"allocs={self.num_allocs} largest={_fmt_bytes(self.max_alloc)}")


_TRACKER = _PinnedMemoryTracker()


def track_pinned_memory(num_bytes):
    _TRACKER.track(num_bytes)


def pinned_memory_summary(tag=""):
    _TRACKER.log_summary(tag)
