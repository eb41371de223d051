from ...sequence.ulysses_sp import *  # noqa: F401,F403
from ...sequence.ulysses_sp import (SequenceTiledCompute, TiledMLP, UlyssesSPAttentionHF,  # noqa: F401
                                    UlyssesSPDataLoaderAdapter)
