from .transformer import DominoTransformer, DominoTransformerLayer, DominoUtil  # noqa: F401
from .async_linear import DominoAsyncColumnParallelLinear, RowParallelLinearNoComm  # noqa: F401
