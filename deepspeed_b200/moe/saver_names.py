import os


def expert_ckpt_name(ckpt_dir, layer_id, expert_id, tag, mp_rank=0):
    return os.path.join(ckpt_dir, str(tag), f"layer_{layer_id}_expert_{expert_id}_mp_rank_{mp_rank:02d}_model_states.pt")
