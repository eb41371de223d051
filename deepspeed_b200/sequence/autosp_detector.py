"""Find the attention modules of a (multimodal) HuggingFace model for AutoSP (reference
``sequence/autosp_detector.py``): ViT-encoder attention, LLM-decoder attention and the vision->language projection,
recognised by class / module name. Extend the registries below for new architectures."""
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import torch.nn as nn

# ViT attention class -> does its token sequence start with a [CLS] token
VIT_ATTENTION = {
    "ViTAttention": True,
    "CLIPAttention": True,
    "SiglipAttention": False,
    "InternVisionAttention": False,
    "Qwen2VLVisionAttention": False,
    "Idefics2VisionAttention": False,
    "PaliGemmaVisionAttention": False,
}
LLM_ATTENTION = {"LlamaAttention", "MistralAttention", "Qwen2Attention", "InternLM2Attention", "GemmaAttention",
                 "Phi3Attention", "GPTNeoXAttention", "FalconAttention", "MptAttention"}
VISION_PROJECTION_KEYWORDS = ("visual_projection", "mm_projector", "vision_proj", "multi_modal_projector",
                              "img_projection")
# names kept for code written against the reference module
_VIT_ATTN_CLASSNAMES = set(VIT_ATTENTION)
_VIT_HAS_CLS_TOKEN = VIT_ATTENTION
_LLM_ATTN_CLASSNAMES = LLM_ATTENTION
_VISION_PROJ_KEYWORDS = VISION_PROJECTION_KEYWORDS


@dataclass
class SPModelInfo:
    vit_attn_modules: List[Tuple[str, nn.Module]] = field(default_factory=list)
    llm_attn_modules: List[Tuple[str, nn.Module]] = field(default_factory=list)
    vision_projection_module: Optional[Tuple[str, nn.Module]] = None  # the outermost match only


def detect_model_sp_info(model: nn.Module) -> SPModelInfo:
    info = SPModelInfo()
    for name, module in model.named_modules():
        cls = type(module).__name__
        if cls in VIT_ATTENTION or cls in _VIT_ATTN_CLASSNAMES:
            info.vit_attn_modules.append((name, module))
        elif cls in LLM_ATTENTION:
            info.llm_attn_modules.append((name, module))
        if info.vision_projection_module is None and any(k in name for k in VISION_PROJECTION_KEYWORDS):
            info.vision_projection_module = (name, module)
    return info
