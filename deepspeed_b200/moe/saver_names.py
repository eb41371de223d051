import os


def expert_ckpt_name(ckpt_dir, layer_id, expert_id, tag, mp_rank=0):
    """Reference ``engine.py:4244 _get_expert_ckpt_name``; ``layer_id`` <= -1 gives the legacy layer-agnostic name."""
    tag_dir = "" if tag is None else str(tag)
    if layer_id <= -1:
        return os.path.join(ckpt_dir, tag_dir, f"expert_{expert_id}_mp_rank_{mp_rank:02d}_model_states.pt")
    return os.path.join(ckpt_dir, tag_dir, f"layer_{layer_id}_expert_{expert_id}_mp_rank_{mp_rank:02d}_model_states.pt")
