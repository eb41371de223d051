"""Vision-language boundary of a sequence-parallel multimodal model (reference ``sequence/autosp_fusion.py``).

The ViT side produces visual tokens sharded over the SP ranks, the LLM side expects the *fused* (text + visual)
sequence sharded uniformly. The adapter wraps the projection layer: project the local visual tokens, all-gather them
(differentiably), splice them into the text embeddings the way the architecture does, pad the fused length to a
multiple of the SP size and hand every rank its slice.

  LlavaFusionAdapter     each image placeholder token is replaced by a chunk of visual tokens (length changes)
  InternVLFusionAdapter  every IMG_CONTEXT token is replaced 1:1 by a visual token (length preserved)
  Qwen2VLFusionAdapter   tokens strictly between vision_start / vision_end markers are replaced 1:1
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import comm as dist
from .autosp_vit import _GatherPatches

_DEFAULT_IMAGE_TOKEN_ID = -200


class ModalityFusionSPAdapter(nn.Module):

    def __init__(self, projection: nn.Module, process_group, image_token_id: int = _DEFAULT_IMAGE_TOKEN_ID):
        super().__init__()
        self.projection = projection
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group)
        self.image_token_id = image_token_id

    def _gather_visual(self, visual_embeds):
        n = torch.tensor([visual_embeds.shape[1]], dtype=torch.long, device=visual_embeds.device)
        lens = [torch.zeros_like(n) for _ in range(self.world_size)]
        dist.all_gather(lens, n, group=self.process_group)
        return _GatherPatches.apply(visual_embeds, self.process_group, [int(t) for t in lens])

    def _shard(self, fused):
        pad = (-fused.shape[1]) % self.world_size
        if pad:
            fused = F.pad(fused, (0, 0, 0, pad))
        per = fused.shape[1] // self.world_size
        r = dist.get_rank(self.process_group)
        return fused[:, r * per:(r + 1) * per].contiguous()

    def forward(self, visual_features, text_embeds, input_ids):
        """visual_features [b, local visual tokens, vit hidden]; text_embeds [b, text len, lm hidden] (full);
        input_ids [b, text len]. Returns this rank's slice [b, fused len / sp, lm hidden]."""
        full_visual = self._gather_visual(self.projection(visual_features))
        return self._shard(self._splice_visual_into_text(text_embeds, full_visual, input_ids))

    def _splice_visual_into_text(self, text_embeds, visual_embeds, input_ids):
        raise NotImplementedError(f"{type(self).__name__}._splice_visual_into_text is not implemented. Subclass "
                                  "ModalityFusionSPAdapter and override this method to match your model's "
                                  "prepare_inputs_embeds logic.")


class LlavaFusionAdapter(ModalityFusionSPAdapter):

    def _splice_visual_into_text(self, text_embeds, visual_embeds, input_ids):
        rows = []
        for b in range(text_embeds.shape[0]):
            where = (input_ids[b] == self.image_token_id).nonzero(as_tuple=True)[0].tolist()
            if not where:
                rows.append(text_embeds[b])
                continue
            chunks = torch.chunk(visual_embeds[b], len(where), dim=0)
            pieces, prev = [], 0
            for chunk, pos in zip(chunks, where):
                pieces.append(text_embeds[b, prev:pos])
                pieces.append(chunk)
                prev = pos + 1  # the placeholder itself is dropped
            pieces.append(text_embeds[b, prev:])
            rows.append(torch.cat(pieces, dim=0))
        width = max(r.shape[0] for r in rows)
        return torch.stack([F.pad(r, (0, 0, 0, width - r.shape[0])) for r in rows])


def _scatter_one_to_one(text_embeds, visual_embeds, mask):
    """Replace the positions where ``mask`` is set, in order, by the leading visual tokens of the same sample."""
    order = (mask.long().cumsum(dim=1) - 1).clamp(min=0)  # index of the visual token for each masked position
    order = order.clamp(max=visual_embeds.shape[1] - 1)
    picked = torch.gather(visual_embeds, 1, order.unsqueeze(-1).expand(-1, -1, visual_embeds.shape[-1]))
    return torch.where(mask.unsqueeze(-1), picked.to(text_embeds.dtype), text_embeds)


class InternVLFusionAdapter(ModalityFusionSPAdapter):

    def _splice_visual_into_text(self, text_embeds, visual_embeds, input_ids):
        return _scatter_one_to_one(text_embeds, visual_embeds, input_ids == self.image_token_id)


class Qwen2VLFusionAdapter(ModalityFusionSPAdapter):

    def __init__(self, projection, process_group, vision_start_token_id, vision_end_token_id):
        super().__init__(projection, process_group, image_token_id=vision_start_token_id)
        self.vision_start_token_id, self.vision_end_token_id = vision_start_token_id, vision_end_token_id

    def _splice_visual_into_text(self, text_embeds, visual_embeds, input_ids):
        starts = (input_ids == self.vision_start_token_id).long().cumsum(dim=1)
        ends = (input_ids == self.vision_end_token_id).long().cumsum(dim=1)
        inside = (starts > ends) & (input_ids != self.vision_start_token_id)  # strictly between a start and its end
        return _scatter_one_to_one(text_embeds, visual_embeds, inside)
