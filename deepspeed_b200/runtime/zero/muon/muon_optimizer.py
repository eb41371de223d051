"""Reference import path ``deepspeed.runtime.zero.muon.muon_optimizer``."""
from ....ops.muon import MuonWithAuxAdam, adam_update  # noqa: F401
