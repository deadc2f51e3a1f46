"""State-dict factory (TP re-sharding of Megatron checkpoints), elastic agent, sparse tensor, OnDevice, misc glue."""
import os
import sys

import pytest
import torch


def _megatron_sd(h=8, heads=2, tp=1, rank=0, ver=0, seed=0):
    g = torch.Generator().manual_seed(seed)
    full = {
        "transformer.layers.0.attention.query_key_value.weight": torch.randn(3 * h, h, generator=g),
        "transformer.layers.0.attention.query_key_value.bias": torch.randn(3 * h, generator=g),
        "transformer.layers.0.attention.dense.weight": torch.randn(h, h, generator=g),
        "transformer.layers.0.attention.dense.bias": torch.randn(h, generator=g),
        "transformer.layers.0.mlp.dense_h_to_4h.weight": torch.randn(4 * h, h, generator=g),
        "transformer.layers.0.mlp.dense_h_to_4h.bias": torch.randn(4 * h, generator=g),
        "transformer.layers.0.mlp.dense_4h_to_h.weight": torch.randn(h, 4 * h, generator=g),
        "transformer.layers.0.mlp.dense_4h_to_h.bias": torch.randn(h, generator=g),
        "transformer.layers.0.input_layernorm.weight": torch.randn(h, generator=g),
        "word_embeddings.weight": torch.randn(16, h, generator=g),
    }
    if tp == 1:
        return full
    out = {}
    for k, v in full.items():
        if "query_key_value" in k:  # version 0: q, k, v thirds each split across ranks
            q, kk, vv = v.chunk(3, 0)
            out[k] = torch.cat([t.chunk(tp, 0)[rank] for t in (q, kk, vv)], 0)
        elif "attention.dense.weight" in k or "dense_4h_to_h.weight" in k:
            out[k] = v.chunk(tp, 1)[rank]
        elif "dense_h_to_4h" in k or "word_embeddings" in k:
            out[k] = v.chunk(tp, 0)[rank]
        else:
            out[k] = v
    return out


def test_megatron_sd_loader_merge_and_split(tmp_path):
    from deepspeedsyclsupport_b200.runtime.state_dict_factory import SDLoaderFactory
    full = _megatron_sd()
    files = []
    for r in range(2):
        p = str(tmp_path / f"mp_rank_{r:02d}_model_states.pt")
        torch.save({"module": _megatron_sd(tp=2, rank=r), "checkpoint_version": 0, "mp_world_size": 2}, p)
        files.append(p)
    loader = SDLoaderFactory.get_sd_loader(files, None, "Megatron")
    _, sd, (scales, merge_count) = loader.load(mp_world_size=1, mp_rank=0)
    assert merge_count == 2 and scales is None
    for k, v in full.items():
        assert torch.equal(sd["module"][k], v), k
    for r in range(4):  # 2 files -> 4 ranks: split
        _, sd4, _ = loader.load(mp_world_size=4, mp_rank=r)
        want = _megatron_sd(tp=4, rank=r)
        for k, v in want.items():
            assert torch.equal(sd4["module"][k], v), (r, k)
    _, sdq, (scales, _) = loader.load(mp_world_size=1, mp_rank=0, quantize=True, quantize_bits=8, quantize_groups=2)
    assert sdq["module"]["transformer.layers.0.attention.dense.weight"].dtype == torch.int8 and scales is not None


def test_elastic_agent_restarts_failed_worker(tmp_path):
    from deepspeedsyclsupport_b200.elasticity.elastic_agent import DSElasticAgent, assign_ranks, pick_members
    assert assign_ranks([3, 0, 5], 3, 2) == ([2, 3], 6)
    assert pick_members([2, 0, 1], min_nodes=2, max_nodes=2) == [0, 1] and pick_members([4], 2, 4) is None
    marker = tmp_path / "attempts.txt"
    script = tmp_path / "worker.py"
    script.write_text("import os, sys\n"
                      f"open({str(marker)!r}, 'a').write(os.environ['TORCHELASTIC_RESTART_COUNT'] + ' ' + os.environ['RANK'] + ' ' + os.environ['WORLD_SIZE'] + '\\n')\n"
                      "sys.exit(1 if os.environ['TORCHELASTIC_RESTART_COUNT'] == '0' else 0)\n")
    agent = DSElasticAgent([sys.executable, str(script)], dict(os.environ), nproc_per_node=2, max_restarts=3, monitor_interval=0.2)
    assert agent.run() == 0
    lines = sorted(marker.read_text().split("\n")[:-1])
    assert lines == ["0 0 2", "0 1 2", "1 0 2", "1 1 2"]
    agent = DSElasticAgent([sys.executable, "-c", "import sys; sys.exit(3)"], dict(os.environ), max_restarts=1, monitor_interval=0.1)
    assert agent.run() == 1


def test_sparse_tensor_roundtrip():
    from deepspeedsyclsupport_b200.runtime.sparse_tensor import SparseTensor
    g = torch.zeros(10, 4)
    g[[1, 7]] = torch.randn(2, 4)
    st = SparseTensor(g)
    assert st.indices.tolist() == [1, 7] and torch.equal(st.to_dense(), g)
    st.add(SparseTensor(g))
    assert torch.equal(st.to_dense(), 2 * g)
    assert torch.equal(st.to_coo_tensor().to_dense(), 2 * g)


def test_on_device_and_glue():
    from deepspeedsyclsupport_b200.utils.init_on_device import OnDevice
    with OnDevice(dtype=torch.bfloat16, device="meta"):
        m = torch.nn.Linear(4, 4)
    assert m.weight.is_meta and m.weight.dtype == torch.bfloat16
    assert not torch.nn.Linear(2, 2).weight.is_meta
    from deepspeedsyclsupport_b200 import git_version_info as gv
    assert gv.version and isinstance(gv.installed_ops, dict) and "version" in gv.torch_info
    from deepspeedsyclsupport_b200.pydantic_v1 import BaseModel  # noqa: F401
    from deepspeedsyclsupport_b200.nebula.config import DeepSpeedNebulaConfig
    assert DeepSpeedNebulaConfig({"nebula": {"enabled": True, "persistent_time_interval": 5}}).persistent_time_interval == 5


@pytest.mark.parametrize("bits,group_dim", [(8, 1), (4, 1), (8, 0)])
def test_post_init_weight_quantization(bits, group_dim):
    """weight_quantization.post_init_quant: matching Linear / Embedding modules become int8 / int4 drop-ins whose outputs stay
    within the quantization error of the original; tied weights stay tied (one quantized copy)."""
    from deepspeedsyclsupport_b200.inference.quantization import (QuantizedEmbedding, QuantizedLinear,
                                                                   _init_group_wise_weight_quantization)

    class Tiny(torch.nn.Module):

        def __init__(self):
            super().__init__()
            self.embed = torch.nn.Embedding(64, 64)
            self.attn = torch.nn.ModuleDict({"q_proj": torch.nn.Linear(64, 64), "keep": torch.nn.Linear(64, 64)})
            self.head = torch.nn.Linear(64, 64, bias=False)
            self.head.weight = self.embed.weight  # tied

        def forward(self, ids):
            return self.head(self.attn["keep"](self.attn["q_proj"](self.embed(ids))))

    torch.manual_seed(0)
    m = Tiny().eval()
    ids = torch.randint(0, 64, (3, 9))
    ref = m(ids)
    spec = {"num_bits": bits, "group_size": 32, "group_dim": group_dim, "symmetric": False}
    cfg = {"weight_quantization": {"post_init_quant": {"q_proj": spec, "embed": spec, "head": spec}}}
    _init_group_wise_weight_quantization(m, cfg)
    assert isinstance(m.attn["q_proj"], QuantizedLinear) and isinstance(m.embed, QuantizedEmbedding) and isinstance(m.head, QuantizedLinear)
    assert isinstance(m.attn["keep"], torch.nn.Linear)  # no key matched
    assert m.head.qweight.data_ptr() == m.embed.qweight.data_ptr()  # tied weights share one payload
    assert m.embed.qweight.numel() == 64 * 64 * bits // 8
    out = m(ids)
    tol = 0.02 if bits == 8 else 0.25
    assert (out - ref).abs().max() < tol * ref.abs().max()


class _FakeUNet(torch.nn.Module):
    in_channels = 4

    def __init__(self):
        super().__init__()
        self.conv = torch.nn.Conv2d(4, 4, 3, padding=1)
        self.proj = torch.nn.Linear(8, 4)

    def forward(self, sample, t, enc, return_dict=True):
        return self.conv(sample) * t.view(-1, 1, 1, 1) + self.proj(enc).mean(1)[:, :, None, None]


class _FakeVAE(torch.nn.Module):

    def __init__(self):
        super().__init__()
        self.d = torch.nn.Conv2d(4, 3, 1)
        self.e = torch.nn.Conv2d(3, 4, 1)

    def decode(self, x, return_dict=True):
        return self.d(x)

    def encode(self, x, return_dict=True):
        return self.e(x)


def test_diffusers_wrappers_pass_through_on_host():
    from deepspeedsyclsupport_b200.model_implementations import DSVAE, DSClipEncoder, DSUNet
    torch.manual_seed(0)
    unet, vae = _FakeUNet(), _FakeVAE()
    w = DSUNet(unet, enable_cuda_graph=True)
    x, t, e = torch.randn(2, 4, 8, 8), torch.tensor([0.5, 0.25]), torch.randn(2, 5, 8)
    assert torch.allclose(w(x, t, e), unet(x, t, e)) and w.in_channels == 4 and w.fwd_count == 1
    assert all(not p.requires_grad for p in w.parameters())
    v = DSVAE(vae)
    assert torch.allclose(v.decode(x), vae.decode(x)) and v.encode(torch.randn(1, 3, 4, 4)).shape == torch.Size([1, 4, 4, 4])
    enc = DSClipEncoder(torch.nn.Linear(8, 8))
    assert torch.allclose(enc(e), enc.enc(e))
    m = DSClipEncoder._build_causal_attention_mask(2, 4, torch.float32)
    assert m.shape == (2, 1, 4, 4) and m[0, 0, 0, 1] < -1e30 and m[0, 0, 1, 0] == 0


def _fallback_allreduce_body():
    import deepspeedsyclsupport_b200 as deepspeed
    from deepspeedsyclsupport_b200 import comm as dist

    class Net(torch.nn.Module):

        def __init__(self):
            super().__init__()
            self.emb = torch.nn.Embedding(20, 8, sparse=True)
            self.lin = torch.nn.Linear(8, 4)

        def forward(self, ids):
            return self.lin(self.emb(ids)).sum()

    torch.manual_seed(0)
    net = Net()
    engine_net = Net()  # the engine's own parameters have their gradients consumed by the flat store: use a free-standing copy
    eng, _, _, _ = deepspeed.initialize(model=engine_net, optimizer=torch.optim.SGD(engine_net.parameters(), lr=0.1),
                                        config={"train_micro_batch_size_per_gpu": 2, "sparse_gradients": True})
    rank, world = dist.get_rank(), dist.get_world_size()
    ids = torch.tensor([[1, 3], [3, 5]]) + rank  # different rows per rank
    net.zero_grad()
    loss = net(ids)
    loss.backward()
    assert net.emb.weight.grad.is_sparse
    dense_before = [net.emb.weight.grad.to_dense().clone(), net.lin.weight.grad.clone()]
    gathered = [[torch.zeros_like(t) for _ in range(world)] for t in dense_before]
    for t, g in zip(dense_before, gathered):
        dist.all_gather(g, t)
    want = [sum(g) / world for g in gathered]
    lin_bias_before = net.lin.bias.grad.clone()
    eng.buffered_allreduce_fallback(grads=[(n, p.grad) for n, p in net.named_parameters()])
    assert torch.allclose(net.emb.weight.grad.to_dense(), want[0], atol=1e-6)
    assert torch.allclose(net.lin.weight.grad, want[1], atol=1e-6)
    # variable-length gather and a few getters of the reference surface
    parts = eng.sparse_all_gather(torch.arange(rank + 2).float())
    assert [p.numel() for p in parts] == [r + 2 for r in range(world)]
    assert eng.eigenvalue_max_iter() == 100 and eng.zero_partial_offload() == 1.0 and eng.dump_state() is False
    assert eng.is_map_style_dataset([1, 2]) and not eng.is_iterable_style_dataset([1, 2])
    eng.update_optimizer_step(7)


def test_engine_fallback_allreduce_dense_and_sparse_two_ranks():
    from common import run_distributed
    run_distributed(_fallback_allreduce_body, world_size=2)


def test_activation_checkpointing_config_object_and_cupy_packing():
    from deepspeedsyclsupport_b200.runtime.activation_checkpointing.config import DeepSpeedActivationCheckpointingConfig
    from deepspeedsyclsupport_b200.runtime.compression.cupy import CupyBackend
    c = DeepSpeedActivationCheckpointingConfig({"activation_checkpointing": {"partition_activations": True, "number_checkpoints": 4}})
    assert c.partition_activations is True and c.number_checkpoints == 4 and c.cpu_checkpointing is False and c.profile is False
    d = DeepSpeedActivationCheckpointingConfig({})
    assert d.partition_activations is False and d.number_checkpoints is None
    be = CupyBackend()
    signs = torch.tensor([1, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 0, 0, 0, 0], dtype=torch.bool)
    chunks = be.compress_by_chunk(be.torch2cupy(signs), 2)
    assert [int(c) for c in chunks] == [0b10000001, 0b11110000]
    assert torch.equal(be.unpackbits(torch.cat(chunks)), signs) and be.cupy2torch(chunks[0]) is chunks[0]
