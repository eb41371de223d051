"""Reference import path ``deepspeed.runtime.zero.muon.original_muon``; implementation in ``ops/muon.py``."""
from ....ops.muon import (Muon, MuonWithAuxAdam, SingleDeviceMuon, SingleDeviceMuonWithAuxAdam, adam_update,  # noqa
                          muon_update, zeropower_via_gram_newtonschulz, zeropower_via_newtonschulz5, NS_METHODS)
