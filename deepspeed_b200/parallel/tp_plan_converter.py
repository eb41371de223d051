"""Hugging Face ``tp_plan`` -> AutoTP layer specs (reference ``module_inject/tp_plan_converter.py`` and
``runtime/tensor_parallel/config.py::_get_hf_tp_plan``).

A plan maps wildcard module paths to a style: ``colwise`` (output-sharded), ``colwise_rep`` /
``colwise_gather_output`` (output-sharded, result all-gathered), ``rowwise`` (input-sharded, result all-reduced), ``rowwise_split_input`` (the same, the whole input is sliced
locally first),
``replicated_with_grad_allreduce`` (kept whole, gradient summed over the group). A plan with any other style is
rejected as a whole: applying half of a column/row pair would compute garbage."""
from typing import Dict, List, Optional

from .autotp_config import PartitionType, TPLayerSpec

SUPPORTED_STYLES = {"colwise", "colwise_rep", "colwise_gather_output", "rowwise", "rowwise_split_input",
                    "replicated_with_grad_allreduce"}


def get_hf_tp_plan(model) -> Optional[Dict[str, str]]:
    """The model's plan: ``model._tp_plan`` (already prefixed paths) or ``config.base_model_tp_plan`` (paths relative
    to the backbone)."""
    plan = getattr(model, "_tp_plan", None)
    if not plan:
        plan = getattr(getattr(model, "config", None), "base_model_tp_plan", None)
    return dict(plan) if plan else None


class TPPlanConverter:

    @staticmethod
    def convert(hf_tp_plan: Dict[str, str]) -> Optional[List[TPLayerSpec]]:
        if not hf_tp_plan:
            return None
        bad = sorted({s for s in hf_tp_plan.values() if str(s).lower() not in SUPPORTED_STYLES})
        if bad:
            raise ValueError(f"HuggingFace tp_plan contains unsupported partition style(s): {bad}. The plan is "
                             "rejected as a whole; provide an explicit 'tensor_parallel.partition_config' instead.")
        specs = []
        for pattern, style in hf_tp_plan.items():
            style = style.lower()
            rx = TPPlanConverter._wildcard_to_regex(pattern)
            rx += "$" if rx.endswith(r"\.weight") else r"\.weight$"
            if style.startswith("colwise"):
                specs.append(TPLayerSpec([rx], PartitionType.COLUMN, gather_output=style != "colwise"))
            elif style == "rowwise":
                specs.append(TPLayerSpec([rx], PartitionType.ROW))
            elif style == "rowwise_split_input":  # follows a gathered column layer (fused QKV / gate-up projections)
                specs.append(TPLayerSpec([rx], PartitionType.ROW, split_input=True))
            else:
                specs.append(TPLayerSpec([rx], PartitionType.SKIP, grad_allreduce=True))
        return specs

    @staticmethod
    def _wildcard_to_regex(pattern: str) -> str:
        return ".*" + pattern.replace(".", r"\.").replace("*", r".*")
