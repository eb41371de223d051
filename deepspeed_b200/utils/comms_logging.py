"""Reference module path of the comms logger (``utils/comms_logging.py``); the implementation is ``comm/logger.py``."""
from ..comm.logger import *  # noqa: F401,F403
from ..comm.logger import CommsLogger, calc_bw_log, get_caller_func  # noqa: F401
