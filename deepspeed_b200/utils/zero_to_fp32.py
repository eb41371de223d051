"""Reference module path ``deepspeed.utils.zero_to_fp32``; the implementation is ``checkpoint/zero_to_fp32.py``."""
from ..checkpoint.zero_to_fp32 import *  # noqa: F401,F403
from ..checkpoint import zero_to_fp32 as _impl

globals().update({k: getattr(_impl, k) for k in dir(_impl) if not k.startswith("__")})

if __name__ == "__main__":
    _impl.main()
