"""Decode CUDA-graph cache of the hybrid engine (runtime/hybrid_engine_graph.py): dispatch logic with a recording
backend on CPU, real graphs on a GPU."""
import pytest
import torch

from deepspeed_b200.runtime.hybrid_engine_graph import (DecodeGraphCache, decode_steps_from_generate_kwargs,
                                                        validate_cuda_graph_support)


class _Recorder:
    """Backend that "captures" by keeping the closure and "replays" by calling it into the same output tensor."""

    def __init__(self):
        self.captures = self.replays = 0

    def capture(self, fn, pool):
        self.captures += 1
        out = fn()
        return (fn, out), out, "pool"

    def replay(self, graph):
        self.replays += 1
        fn, out = graph
        out.copy_(fn())


class _Toy:
    """Static-cache decoder stand-in: writes ``input_ids`` at ``cache_position`` and returns the running sum."""

    def __init__(self):
        self.cache = torch.zeros(2, 16)
        self.calls = 0

    def forward(self, input_ids=None, cache_position=None, past_key_values=None, use_cache=True):
        self.calls += 1
        if input_ids.shape[1] == 1:
            self.cache.index_copy_(1, cache_position, input_ids.float())
        else:
            self.cache[:, :input_ids.shape[1]] = input_ids.float()
        return self.cache.sum(1, keepdim=True).clone()


def test_one_graph_serves_every_position_with_device_side_position():
    toy, rec = _Toy(), _Recorder()
    g = DecodeGraphCache(toy.forward, max_positions=8, backend=rec)
    kv = object()
    g.begin_sequence(5)
    out = g(input_ids=torch.tensor([[1, 2, 3], [4, 5, 6]]), past_key_values=kv)  # prompt: eager
    assert rec.captures == 0 and out.flatten().tolist() == [6.0, 15.0]
    for step in range(5):
        out = g(input_ids=torch.full((2, 1), step + 1), cache_position=torch.tensor([3 + step]), past_key_values=kv)
    assert rec.captures == 1 and rec.replays == 5 and g.captured_positions == 1
    assert out.flatten().tolist() == [6.0 + 15, 15.0 + 15]
    with pytest.raises(RuntimeError, match="exceeded the declared length"):
        g(input_ids=torch.ones(2, 1, dtype=torch.long), cache_position=torch.tensor([9]), past_key_values=kv)
    # a new cache object (new generate call) or a new batch shape is another signature
    g.begin_sequence(2)
    g(input_ids=torch.ones(2, 1, dtype=torch.long), cache_position=torch.tensor([0]), past_key_values=object())
    assert rec.captures == 2


def test_growing_mask_gets_a_graph_per_position_and_eviction():
    calls = []

    def fwd(input_ids=None, attention_mask=None):
        calls.append(attention_mask.shape[1])
        return attention_mask.sum(1, keepdim=True).float()

    rec = _Recorder()
    g = DecodeGraphCache(fwd, max_positions=3, backend=rec)
    for _ in range(2):  # the second sequence replays the first one's graphs
        g.begin_sequence(3)
        for p in range(3):
            out = g(input_ids=torch.ones(1, 1, dtype=torch.long), attention_mask=torch.ones(1, 4 + p))
            assert float(out) == 4 + p
    assert rec.captures == 3 and g.captured_positions == 3 and rec.replays == 6
    g.begin_sequence(4)  # longer than max_positions: eager
    n = len(calls)
    g(input_ids=torch.ones(1, 1, dtype=torch.long), attention_mask=torch.ones(1, 9))
    assert rec.captures == 3 and len(calls) == n + 1
    g.begin_sequence(None)
    g(input_ids=torch.ones(1, 1, dtype=torch.long), attention_mask=torch.ones(1, 9))
    assert rec.captures == 3
    g.invalidate()
    assert g.captured_positions == 0


def test_capture_failure_disables_graphs_for_good():

    class Bad(_Recorder):

        def capture(self, fn, pool):
            raise RuntimeError("no graphs today")

    toy = _Toy()
    g = DecodeGraphCache(toy.forward, max_positions=4, backend=Bad())
    g.begin_sequence(2)
    out = g(input_ids=torch.ones(2, 1, dtype=torch.long), cache_position=torch.tensor([0]))
    assert out.flatten().tolist() == [1.0, 1.0] and toy.calls == 1
    g.begin_sequence(2)
    assert g._use_graphs is False


def test_length_and_support_rules():
    assert decode_steps_from_generate_kwargs({"max_new_tokens": 8, "min_new_tokens": 8}) == 7
    assert decode_steps_from_generate_kwargs({"max_new_tokens": 8}) is None
    assert decode_steps_from_generate_kwargs({"max_new_tokens": 8, "min_new_tokens": 4}) is None

    class C:
        release_inference_cache, inference_tp_size = False, 1

    if not torch.cuda.is_available():
        assert "CUDA" in validate_cuda_graph_support(C(), 0)


@pytest.mark.gpu
def test_real_cuda_graphs_match_eager_decode():
    dev = torch.device("cuda")
    torch.manual_seed(0)
    emb = torch.nn.Embedding(50, 64).to(dev).half()
    lin = torch.nn.Linear(64, 50).to(dev).half()
    cache = torch.zeros(2, 32, 64, device=dev, dtype=torch.half)

    def fwd(input_ids=None, cache_position=None):
        h = emb(input_ids)
        if input_ids.shape[1] == 1:
            cache.index_copy_(1, cache_position, h)
        else:
            cache[:, :input_ids.shape[1]] = h
        valid = (torch.arange(32, device=dev) <= cache_position.max()).to(h.dtype)[None, :, None]
        return lin((cache * valid).sum(1))

    def run(graphs):
        cache.zero_()
        f = DecodeGraphCache(fwd, 16) if graphs else fwd
        if graphs:
            f.begin_sequence(6)
        ids = torch.tensor([[1, 2, 3], [4, 5, 6]], device=dev)
        logits = f(input_ids=ids, cache_position=torch.arange(3, device=dev))
        toks = []
        for step in range(6):
            nxt = logits.argmax(-1, keepdim=True)
            toks.append(nxt.clone())
            logits = f(input_ids=nxt, cache_position=torch.tensor([3 + step], device=dev))
        return torch.cat(toks, 1), logits.clone(), f

    t_e, l_e, _ = run(False)
    t_g, l_g, f = run(True)
    assert torch.equal(t_e, t_g) and torch.equal(l_e, l_g)
    assert f.captured_positions == 1 and f.replays == 6
    assert validate_cuda_graph_support(type("C", (), {"release_inference_cache": False, "inference_tp_size": 1})(),
                                       3) is not None
