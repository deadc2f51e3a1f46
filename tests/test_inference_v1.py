"""init_inference: AutoTP parser finds the all-reduce linears; TP=2 output equals the un-sharded model (gloo)."""
import torch
from torch import nn

from common import run_distributed


class Attn(nn.Module):

    def __init__(self, h, heads):
        super().__init__()
        self.num_heads, self.hidden_size = heads, h
        self.q_proj, self.k_proj, self.v_proj, self.o_proj = (nn.Linear(h, h) for _ in range(4))

    def forward(self, x):
        B, S, _ = x.shape
        d = self.q_proj.out_features // self.num_heads
        q, k, v = (p(x).view(B, S, self.num_heads, d).transpose(1, 2) for p in (self.q_proj, self.k_proj, self.v_proj))
        o = torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True)
        return self.o_proj(o.transpose(1, 2).reshape(B, S, -1))


class MLP(nn.Module):

    def __init__(self, h):
        super().__init__()
        self.gate_proj, self.up_proj, self.down_proj = nn.Linear(h, 2 * h), nn.Linear(h, 2 * h), nn.Linear(2 * h, h)

    def forward(self, x):
        return self.down_proj(torch.nn.functional.silu(self.gate_proj(x)) * self.up_proj(x))


class Block(nn.Module):

    def __init__(self, h, heads):
        super().__init__()
        self.input_layernorm, self.self_attn = nn.LayerNorm(h), Attn(h, heads)
        self.post_attention_layernorm, self.mlp = nn.LayerNorm(h), MLP(h)

    def forward(self, x):
        x = x + self.self_attn(self.input_layernorm(x))
        return x + self.mlp(self.post_attention_layernorm(x))


class TinyDecoderModel(nn.Module):

    def __init__(self, h=32, heads=4, n=2, vocab=50):
        super().__init__()
        self.embed = nn.Embedding(vocab, h)
        self.layers = nn.ModuleList([Block(h, heads) for _ in range(n)])
        self.norm = nn.LayerNorm(h)
        self.lm_head = nn.Linear(h, vocab, bias=False)

    def forward(self, ids):
        x = self.embed(ids)
        for l in self.layers:
            x = l(x)
        return self.lm_head(self.norm(x))


def test_tp_parser_finds_allreduce_linears():
    from deepspeedsyclsupport_b200.module_inject.auto_tp import AutoTP
    policy = AutoTP.tp_parser(TinyDecoderModel())
    assert len(policy) == 1 and policy[0][0] is Block
    assert set(policy[0][1]) == {"self_attn.o_proj", "mlp.down_proj"}


def test_inference_config_aliases_and_deprecation():
    from deepspeedsyclsupport_b200.inference.config import DeepSpeedInferenceConfig
    c = DeepSpeedInferenceConfig(mp_size=2, dtype="bf16", kernel_inject=True, max_tokens=77)
    assert c.tensor_parallel.tp_size == 2 and c.dtype == torch.bfloat16 and c.replace_with_kernel_inject and c.max_out_tokens == 77


def _tp_body():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200.module_inject.layers import LinearAllreduce, LinearLayer
    torch.manual_seed(0)
    model = TinyDecoderModel().eval()
    ids = torch.randint(0, 50, (2, 9))
    with torch.no_grad():
        ref = model(ids)
    engine = deepspeed.init_inference(model, tensor_parallel={"tp_size": 2}, dtype=torch.float32)
    blk = engine.module.layers[0]
    assert isinstance(blk.self_attn.o_proj, LinearAllreduce) and isinstance(blk.self_attn.q_proj, LinearLayer)
    assert blk.self_attn.q_proj.weight.shape == (16, 32) and blk.self_attn.o_proj.weight.shape == (32, 16)
    assert blk.self_attn.num_heads == 2
    with torch.no_grad():
        out = engine(ids)
    assert torch.allclose(out, ref, atol=1e-5), (out - ref).abs().max()


def test_init_inference_autotp_two_ranks():
    run_distributed(_tp_body, world_size=2)
