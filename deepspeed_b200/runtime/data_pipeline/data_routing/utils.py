"""Pure-PyTorch gather/scatter used by model code that applies random-LTD by hand (reference
``runtime/data_pipeline/data_routing/utils.py``)."""
import torch


def bsh_decoder_gather(reserved_length, hidden_states, mask):
    B, S, _ = hidden_states.shape
    keep = torch.rand(B, S, device=hidden_states.device).argsort(dim=1)[:, :reserved_length].sort(dim=1).values
    part = torch.gather(hidden_states, 1, keep[:, :, None].expand(-1, -1, hidden_states.shape[2]))
    return part, list(keep), mask[:, :, :reserved_length, :reserved_length]


def bsh_decoder_scatter(hidden_states, part_hidden_states, rand_list):
    for k, rows in enumerate(rand_list):
        hidden_states[k, rows, :] = part_hidden_states[k]
    return hidden_states
