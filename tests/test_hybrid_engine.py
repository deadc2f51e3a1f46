"""Hybrid engine: generate() under ZeRO-3 (pinned gather) with LoRA folding, then keep training (gloo, 2 ranks)."""
import torch
from torch import nn

from common import run_distributed


class LoRALinear(nn.Module):
    """DeepSpeed-Chat LinearLayer_LoRA convention."""

    def __init__(self, i, o, r=2):
        super().__init__()
        self.weight = nn.Parameter(torch.randn(o, i) * 0.1)
        self.lora_right_weight = nn.Parameter(torch.randn(i, r) * 0.1)
        self.lora_left_weight = nn.Parameter(torch.randn(r, o) * 0.1)
        self.lora_scaling = 0.5
        self.fused = False

    def forward(self, x):
        y = torch.nn.functional.linear(x, self.weight)
        if not getattr(self, "_probe_fused", False):
            y = y + (x @ self.lora_right_weight @ self.lora_left_weight) * self.lora_scaling
        return y


class TinyLM(nn.Module):

    def __init__(self, vocab=17, h=8):
        super().__init__()
        self.emb = nn.Embedding(vocab, h)
        self.layers = nn.ModuleList([LoRALinear(h, h) for _ in range(2)])
        self.head = nn.Linear(h, vocab)

    def logits(self, ids):
        x = self.emb(ids)
        for l in self.layers:
            x = x + torch.tanh(l(x))
        return self.head(x)

    def forward(self, ids, labels):
        return torch.nn.functional.cross_entropy(self.logits(ids).flatten(0, 1), labels.flatten())

    def generate(self, ids, max_new_tokens=4):
        # inside generate the engine has folded LoRA into `weight`: the explicit LoRA branch must be skipped
        for l in self.layers:
            l._probe_fused = True
        try:
            for _ in range(max_new_tokens):
                nxt = self.logits(ids)[:, -1].argmax(-1, keepdim=True)
                ids = torch.cat([ids, nxt], 1)
        finally:
            for l in self.layers:
                l._probe_fused = False
        return ids


def _body():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200.runtime.hybrid_engine import DeepSpeedHybridEngine
    from deepspeedsyclsupport_b200.runtime.zero.param_store import ZeroParamStatus
    torch.manual_seed(0)
    model = TinyLM()
    ref = TinyLM()
    ref.load_state_dict(model.state_dict())
    cfg = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}, "hybrid_engine": {"enabled": True, "pin_parameters": True}}
    engine, _, _, _ = deepspeed.initialize(model=model, config=cfg)
    assert isinstance(engine, DeepSpeedHybridEngine)
    prompt = torch.tensor([[1, 2, 3], [4, 5, 6]])
    engine.eval()
    out = engine.module.generate(prompt, max_new_tokens=3)
    with torch.no_grad():
        want = ref.eval()
        ids = prompt
        for _ in range(3):
            ids = torch.cat([ids, want.logits(ids)[:, -1].argmax(-1, keepdim=True)], 1)
    assert torch.equal(out, ids), (out, ids)
    assert not engine.is_lora_fused
    assert all(u.status == ZeroParamStatus.NOT_AVAILABLE for u in engine.optimizer.units if not u.persistent)
    engine.train()
    labels = torch.randint(0, 17, (2, 3))
    l0 = None
    for _ in range(3):
        loss = engine(prompt, labels)
        engine.backward(loss)
        engine.step()
        l0 = l0 if l0 is not None else loss.item()
    assert loss.item() < l0


def test_hybrid_engine_generate_then_train():
    run_distributed(_body, world_size=2)
