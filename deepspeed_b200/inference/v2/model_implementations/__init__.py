from .arch import ArchSpec, GenericDecoder  # noqa: F401
from .policies import SPECS, REFERENCE_FAMILIES, build_hf_model, load_hf_weights  # noqa: F401
