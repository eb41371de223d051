"""MoE: capacity (DS-MoE ``MoE``) and dropless (AutoEP) layers; expert parallel (ep=2) matches ep=1
(reference tests/unit/v1/moe/*)."""
import pytest
import torch

from common import run_distributed


def test_topk_gating_capacity_and_aux_loss():
    from deepspeed_b200.moe.sharded_moe import topkgating
    torch.manual_seed(0)
    logits = torch.randn(64, 4)
    l_aux, idx, slot, w, cap, counts = topkgating(logits, 2, capacity_factor=1.0, min_capacity=4)
    assert cap == 32 and idx.shape == (64, 2) and int(counts.sum()) == 128
    assert (slot[slot >= 0] < cap).all() and float(l_aux) > 0
    # slots are unique per expert
    for e in range(4):
        s = slot[(idx == e) & (slot >= 0)]
        assert s.numel() == s.unique().numel()


def test_ds_moe_layer_equals_dense_expert_when_single_expert():
    from deepspeed_b200.moe import MoE
    torch.manual_seed(0)
    expert = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.GELU(), torch.nn.Linear(32, 16))
    moe = MoE(16, expert, num_experts=1, ep_size=1, k=1, capacity_factor=4.0, min_capacity=64)
    moe.deepspeed_moe._set_ep_group(None)
    x = torch.randn(2, 8, 16)
    out, l_aux, counts = moe(x)
    ref = moe.deepspeed_moe.experts.deepspeed_experts[0](x)
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-6)


def _autoep(ep):
    import deepspeed_b200 as ds
    from deepspeed_b200.models.mixtral import MixtralConfig, MixtralForCausalLM
    from deepspeed_b200.utils import groups
    torch.manual_seed(0)
    cfg = MixtralConfig.tiny(dtype=torch.float32, gradient_checkpointing=False, ep_size=1, router_aux_loss_coef=0.0)
    full = MixtralForCausalLM(cfg)
    sd = full.state_dict()
    rank = ds.comm.get_rank()
    if ep > 1:
        cfg2 = MixtralConfig.tiny(dtype=torch.float32, gradient_checkpointing=False, ep_size=ep, router_aux_loss_coef=0.0)
        model = MixtralForCausalLM(cfg2)
        nl = cfg.num_local_experts // ep
        for k in list(sd.keys()):
            if ".experts.w" in k:
                sd[k] = sd[k][rank * nl:(rank + 1) * nl].clone()
        model.load_state_dict(sd)
        groups.initialize()
        model.set_deepspeed_parallelism()
    else:
        model = full
    torch.manual_seed(7)
    ids = torch.randint(0, 512, (2, 16))  # same tokens on every rank
    loss = model(ids, labels=ids)
    loss.backward()
    g = model.layers[0].block_sparse_moe.experts.w2.grad.clone()
    gr = model.layers[0].block_sparse_moe.router.weight.grad.clone()
    return float(loss), g, gr, rank


def test_autoep_expert_parallel_matches_single():
    single = run_distributed(_autoep, 1, args=(1, ))[0]
    ep = run_distributed(_autoep, 2, args=(2, ))
    for loss, g, gr, rank in ep:
        assert abs(loss - single[0]) < 1e-4
        nl = single[1].shape[0] // 2
        # every rank saw the same batch, so each local expert accumulated both ranks' (identical) tokens: 2x
        torch.testing.assert_close(g, 2 * single[1][rank * nl:(rank + 1) * nl], rtol=1e-3, atol=1e-5)
        torch.testing.assert_close(gr, single[2], rtol=1e-3, atol=1e-5)


def _dense_top2(logits, capacity_factor, min_capacity, noise):
    """Dense restatement of the reference top-2 gate (moe/sharded_moe.py:341-432): combine weights [S, E, C]."""
    import math
    import torch.nn.functional as F
    S, E = logits.shape
    gates = F.softmax(logits, dim=1)
    i1 = gates.argmax(1)
    m1 = F.one_hot(i1, E)
    noisy = logits + noise if noise is not None else logits
    i2 = noisy.masked_fill(m1.bool(), float("-inf")).argmax(1)
    m2 = F.one_hot(i2, E)
    l1 = torch.cumsum(m1, 0) - 1
    l2 = torch.cumsum(m2, 0) - 1 + m1.sum(0, keepdim=True)
    l_aux = (gates.mean(0) * m1.float().mean(0)).mean() * E * E
    cap = max(int(math.ceil(S / E * capacity_factor * 2)), min_capacity)
    m1 = m1 * (l1 < cap)
    m2 = m2 * (l2 < cap)
    s1, s2 = (l1 * m1).sum(1), (l2 * m2).sum(1)
    g1, g2 = (gates * m1).sum(1), (gates * m2).sum(1)
    den = (g1 + g2).clamp(min=torch.finfo(gates.dtype).eps)
    comb = torch.zeros(S, E, cap)
    for s in range(S):
        if m1[s].any():
            comb[s, i1[s], s1[s]] += g1[s] / den[s]
        if m2[s].any():
            comb[s, i2[s], s2[s]] += g2[s] / den[s]
    return l_aux, comb, cap


@pytest.mark.parametrize("sampling", [False, True])
def test_top2_gating_matches_reference_rules(sampling):
    from deepspeed_b200.moe import sharded_moe as sm
    torch.manual_seed(7)
    S, E = 64, 4
    logits = torch.randn(S, E) * 2.0
    logits[:, 0] += 1.5  # skew towards expert 0 so that the capacity limit drops tokens
    noise = None
    if sampling:
        torch.manual_seed(99)
        noise = sm.gumbel_rsample(logits.shape, device=logits.device)
        torch.manual_seed(99)  # top2gating draws the same noise
    l_aux, idx, slot, w, cap, counts = sm.top2gating(logits.clone(), 0.5, 4, True, None, sampling)
    want_aux, want, want_cap = _dense_top2(logits.clone(), 0.5, 4, noise)
    assert cap == want_cap
    torch.testing.assert_close(l_aux, want_aux)
    got = torch.zeros(S, E, cap)
    for s in range(S):
        for j in range(2):
            if slot[s, j] >= 0:
                got[s, idx[s, j], slot[s, j]] += w[s, j]
    assert (slot < 0).any(), "test must exercise dropped tokens"
    torch.testing.assert_close(got, want)


def _fused_plan_roundtrip(top_k, n_experts):
    """CPU emulation of the fused MoE dispatch / combine kernels driven by ``build_plan``: every token row must reach the
    right expert rank at the planned row, and the combine must find it again (index math of moe/fused_dispatch.py)."""
    import deepspeed_b200 as ds
    from deepspeed_b200.moe.fused_dispatch import build_plan
    dist = ds.comm
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(10 + rank)
    T, H = 37 + 5 * rank, 8
    x = torch.randn(T, H)
    scores = torch.rand(T, n_experts)
    top_w, top_i = scores.topk(top_k, dim=-1)
    plan = build_plan(top_i, n_experts, world, rank, None)
    E_local = n_experts // world
    # what the kernels read from the peers: token rows, expert-sorted token lists
    Ts = [None] * world
    dist.all_gather_object(Ts, (x, plan.sorted_tok, plan.segs.clone(), plan.n_rows))
    # ---- dispatch (expert rank pulls) ----
    expert_in = torch.zeros(plan.n_rows, H)
    for (src, s0, d0, n) in plan.segs.tolist():
        for i in range(n):
            expert_in[d0 + i] = Ts[src][0][int(Ts[src][1][s0 + i])]
    assert sum(plan.local_counts) == plan.n_rows
    # rows of local expert el are contiguous and really were routed to it
    off = 0
    for el, cnt in enumerate(plan.local_counts):
        off += cnt
    # ---- "experts": tag each row with its local expert id so the combine can be checked ----
    tag = torch.zeros(plan.n_rows)
    off = 0
    for el, cnt in enumerate(plan.local_counts):
        tag[off:off + cnt] = rank * E_local + el
        off += cnt
    outs = [None] * world
    dist.all_gather_object(outs, (expert_in, tag))
    # ---- combine (token owner pulls) ----
    for t in range(T):
        for j in range(top_k):
            r, row = int(plan.slot_rank[t * top_k + j]), int(plan.slot_row[t * top_k + j])
            assert torch.equal(outs[r][0][row], x[t]), (t, j)          # my token row sits where the plan says
            assert int(outs[r][1][row]) == int(top_i[t, j]), (t, j)    # ... inside the segment of the chosen expert
    return True


@pytest.mark.parametrize("top_k,n_experts", [(2, 4), (1, 3), (2, 6)])
def test_fused_moe_plan_index_math(top_k, n_experts):
    world = 3 if n_experts % 3 == 0 else 2
    assert all(run_distributed(_fused_plan_roundtrip, world, args=(top_k, n_experts)))


def _moe_tp_drop():
    """tp=2 (one replica): both ranks hold the same tokens; the DS-MoE layer sends half of every expert's capacity
    through the experts on each rank and gathers the halves - the result and the gradients equal the tp=1 layer."""
    import deepspeed_b200 as ds
    from deepspeed_b200.moe import MoE
    from deepspeed_b200.moe.mappings import drop_tokens, gather_tokens
    from deepspeed_b200.utils import groups
    torch.manual_seed(0)
    expert = torch.nn.Sequential(torch.nn.Linear(16, 32), torch.nn.GELU(), torch.nn.Linear(32, 16))
    moe = MoE(16, expert, num_experts=2, ep_size=1, k=1, capacity_factor=2.0, min_capacity=8, use_rts=False)
    x = torch.randn(2, 8, 16)
    moe.deepspeed_moe._set_ep_group(None)
    xr = x.clone().requires_grad_(True)
    want, _, _ = moe(xr)
    want.square().sum().backward()
    gw = [p.grad.clone() for p in moe.parameters()]
    gx = xr.grad.clone()
    for p in moe.parameters():
        p.grad = None
    groups.initialize(tp_size=2, force=True)
    assert groups._get_model_parallel_world_size() == 2
    t = torch.arange(8.0).view(1, 8, 1).requires_grad_(True)
    half = drop_tokens(t, dim=1)
    assert half.shape == (1, 4, 1) and float(half[0, 0, 0]) == 4.0 * ds.comm.get_rank()
    back = gather_tokens(half, dim=1)
    assert torch.equal(back, t)
    back.sum().backward()
    assert torch.equal(t.grad, torch.ones_like(t))
    xt = x.clone().requires_grad_(True)
    got, _, _ = moe(xt)
    got.square().sum().backward()
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(xt.grad, gx, rtol=1e-5, atol=1e-6)
    # expert weights saw half of the tokens on each rank: the halves add up to the tp=1 gradient
    for p, g in zip(moe.parameters(), gw):
        if "experts" in [n for n, q in moe.named_parameters() if q is p][0]:
            tot = p.grad.clone()
            ds.comm.all_reduce(tot)
            torch.testing.assert_close(tot, g, rtol=1e-4, atol=1e-6)
    return True


def test_ds_moe_drops_duplicate_tokens_across_tensor_parallel_ranks():
    assert all(run_distributed(_moe_tp_drop, 2))
