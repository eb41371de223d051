"""Reference import path ``deepspeed.ops.op_builder``: builder shims (everything is prebuilt, see ``ops/builder.py``)."""
from ..op_builders import *  # noqa: F401,F403
from ..op_builders import ALL_OPS, _Builder as OpBuilder  # noqa: F401


def get_default_compute_capabilities():
    return "10.0a"


__op_builders__ = [b() for b in ALL_OPS.values()]
