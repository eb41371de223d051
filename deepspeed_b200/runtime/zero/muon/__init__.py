from ....ops.muon import *  # noqa: F401,F403
