"""Compression library (reference tests: tests/unit/compression/test_compression.py)."""
import torch
from torch import nn


class Net(nn.Module):

    def __init__(self):
        super().__init__()
        self.layer = nn.ModuleList([nn.ModuleDict({"attn_out": nn.Linear(16, 16), "fc1": nn.Linear(16, 32), "fc2": nn.Linear(32, 16)}) for _ in range(2)])
        self.emb = nn.Embedding(10, 16)

    def forward(self, ids):
        x = self.emb(ids)
        for l in self.layer:
            x = x + l["attn_out"](x)
            x = x + l["fc2"](torch.relu(l["fc1"](x)))
        return x


def _cfg():
    return {"compression_training": {
        "weight_quantization": {"shared_parameters": {"enabled": True, "quantize_weight_in_forward": True, "quantize_groups": 1, "schedule_offset": 2,
                                                      "quantization_type": "symmetric"},
                                "different_groups": {"wq1": {"params": {"start_bits": 8, "target_bits": 4, "quantization_period": 1}, "modules": ["attn_out"]}}},
        "activation_quantization": {"shared_parameters": {"enabled": True, "quantization_type": "asymmetric", "range_calibration": "dynamic", "schedule_offset": 0},
                                    "different_groups": {"aq1": {"params": {"bits": 8}, "modules": ["attn_out"]}}},
        "sparse_pruning": {"shared_parameters": {"enabled": True, "method": "l1", "schedule_offset": 0, "schedule_offset_end": 100},
                           "different_groups": {"sp1": {"params": {"dense_ratio": 0.5}, "modules": ["attn_out"]}}},
        "row_pruning": {"shared_parameters": {"enabled": True, "method": "l1", "schedule_offset": 1},
                        "different_groups": {"rp1": {"params": {"dense_ratio": 0.5}, "modules": ["fc1"], "related_modules": [["fc2"]]}}},
    }}


def test_quantizers_against_formulas():
    from deepspeedsyclsupport_b200.compression.utils import AsymQuantizer, BinaryQuantizer, SymQuantizer, TernaryQuantizer, TopKBinarizer
    torch.manual_seed(0)
    x = torch.randn(4, 32)
    q = SymQuantizer.apply(x, 4, None, None, 4)
    for g in range(4):
        scale = 2 * x[g].abs().max() / 16
        assert torch.allclose(q[g], (x[g] / scale).round().clamp(-8, 7) * scale, atol=1e-6)
    q = AsymQuantizer.apply(x, 8, None, None, 1)
    assert (q - x).abs().max() <= (x.max() - x.min()) / 256 + 1e-6
    assert TernaryQuantizer.apply(x, 2, None, None, 2).unique().numel() <= 6
    b = BinaryQuantizer.apply(x, 1, None, None, 1)
    assert torch.allclose(b.abs(), x.abs().mean().expand_as(b))
    m = TopKBinarizer.apply(x.abs(), 0.25, False)
    assert int(m.sum()) == 32 and x.abs()[m.bool()].min() >= x.abs()[~m.bool()].max()
    xr = x.clone().requires_grad_()
    SymQuantizer.apply(xr, 4, None, None, 1).sum().backward()
    assert torch.equal(xr.grad, torch.ones_like(x))  # straight-through


def test_init_schedule_and_clean():
    from deepspeedsyclsupport_b200.compression import init_compression, redundancy_clean
    from deepspeedsyclsupport_b200.compression.basic_layer import Embedding_Compress, LinearLayer_Compress
    from deepspeedsyclsupport_b200.compression.config import get_compression_config
    from deepspeedsyclsupport_b200.compression.scheduler import compression_scheduler
    torch.manual_seed(0)
    model = Net()
    cfg = _cfg()
    init_compression(model, cfg)
    l0 = model.layer[0]
    assert isinstance(l0["attn_out"], LinearLayer_Compress) and isinstance(model.emb, Embedding_Compress)
    assert l0["attn_out"].sparse_pruning_method == "l1" and l0["fc1"].row_pruning_method == "l1" and l0["attn_out"].weight.target_bits == 4
    sched = compression_scheduler(model, get_compression_config(cfg))
    sched.step(step_zero_check=True)
    assert l0["attn_out"].sparse_pruning_enabled and l0["attn_out"].activation_quantization_enabled
    assert not l0["fc1"].row_pruning_enabled and not l0["attn_out"].weight_quantization_enabled
    sched.step()
    assert l0["fc1"].row_pruning_enabled
    sched.step()
    assert l0["attn_out"].weight_quantization_enabled and sched.weight_quantization_enabled
    for l in model.layer:
        l["attn_out"].activation_quantization_enabled = False  # keep the before/after comparison exact
    ids = torch.randint(0, 10, (3, 5))
    model.eval()
    with torch.no_grad():
        y_masked = model(ids)
    redundancy_clean(model, cfg)
    assert l0["fc1"].out_features == 16 and l0["fc2"].in_features == 16 and l0["fc1"].weight.shape == (16, 16)
    assert (l0["attn_out"].weight == 0).float().mean() >= 0.45
    assert l0["attn_out"].weight.unique().numel() <= 16  # 4-bit grid
    with torch.no_grad():
        y_clean = model(ids)
    assert torch.allclose(y_masked, y_clean, atol=1e-5), (y_masked - y_clean).abs().max()


def test_student_initialization_and_head_pruning():
    from deepspeedsyclsupport_b200.compression.basic_layer import LinearLayer_Compress
    from deepspeedsyclsupport_b200.compression.compress import student_initialization
    torch.manual_seed(0)
    teacher, student = nn.ModuleDict({"h": nn.ModuleList([nn.Linear(4, 4) for _ in range(4)]), "head": nn.Linear(4, 2)}), \
        nn.ModuleDict({"h": nn.ModuleList([nn.Linear(4, 4) for _ in range(2)]), "head": nn.Linear(4, 2)})
    cfg = {"compression_training": {"layer_reduction": {"enabled": True, "keep_number_layer": 2, "module_name_prefix": "h", "teacher_layer": [1, 3],
                                                        "other_module_name": ["head"]}}}
    student_initialization(student, teacher, cfg)
    assert torch.equal(student["h"][0].weight, teacher["h"][1].weight) and torch.equal(student["h"][1].weight, teacher["h"][3].weight)
    assert torch.equal(student["head"].bias, teacher["head"].bias)
    o = LinearLayer_Compress(8, 8)
    o.enable_head_pruning(0.5, "topk", num_heads=4)
    o.head_pruning_enabled = True
    x = torch.randn(2, 8)
    y = o(x)
    mask = o.get_mask("head").view(-1).bool()
    assert int(mask.sum()) == 2
    col_mask = mask.repeat_interleave(2)
    assert torch.allclose(y, torch.nn.functional.linear(x * col_mask, o.weight, o.bias), atol=1e-6)
    kept = o.fix_head_pruning_helper(dim_reduction=True)
    assert o.weight.shape == (8, 4) and torch.equal(kept.view(-1), mask)
