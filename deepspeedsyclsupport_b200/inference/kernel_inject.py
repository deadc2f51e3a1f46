"""`init_inference(model, replace_with_kernel_inject=True)`: run a HuggingFace causal LM on this framework's kernels.

The reference swaps every transformer block of a supported architecture for a fused `DeepSpeedTransformerInference`
module through per-architecture policy containers (module_inject/containers/*, ops/transformer/inference/*,
csrc/transformer/inference).  Here the fused kernels already exist as one generic ragged transformer
(inference/v2/model_implementations/ragged_transformer.py: fused norm / QKV GEMM + RoPE + KV-append / paged attention /
gated MLP, blocked KV cache), so injection means: map the HF module's weights onto that implementation once (no copy of
the checkpoint files, tensor-parallel sharding included) and route `generate()` through it with continuous batching of
the prompts.  Supported `config.model_type`s: llama (incl. attention biases = internlm), mistral, mixtral, opt, falcon,
phi, gpt2, gptj, gpt_neox.  (The reference's bloom / gpt-neo containers need ALiBi / windowed attention in the paged kernel:
not implemented; those models run through AutoTP instead.)"""
from types import SimpleNamespace

import torch
from torch import nn

from ..utils.logging import log_dist

SUPPORTED = ("llama", "mistral", "internlm", "mixtral", "opt", "falcon", "phi", "gpt2", "gptj", "gpt_neox")


def can_inject(model):
    cfg = getattr(model, "config", None)
    return cfg is not None and getattr(cfg, "model_type", None) in SUPPORTED


class KernelInjectedCausalLM(nn.Module):
    """Drop-in for the wrapped HF model's inference surface: `generate()` and a last-position `forward()`."""

    def __init__(self, hf_model, dtype=torch.bfloat16, tp_size=1, max_context=None, max_batch_tokens=None, max_sequences=None):
        super().__init__()
        from .v2 import RaggedInferenceEngineConfig, build_engine_from_state_dict
        cfg = hf_model.config
        self.config = cfg
        hf_cfg = cfg.to_dict() if hasattr(cfg, "to_dict") else dict(cfg.__dict__)
        max_context = max_context or int(getattr(cfg, "max_position_embeddings", 2048))
        ecfg = RaggedInferenceEngineConfig(
            tensor_parallel={"tp_size": tp_size},
            state_manager={"max_ragged_batch_size": max_batch_tokens or min(max_context, 2048), "max_ragged_sequence_count": max_sequences or 64,
                           "max_context": max_context, "max_tracked_sequences": max(2 * (max_sequences or 64), 128)})
        sd = {k: v.detach() for k, v in hf_model.state_dict().items()}
        self.engine = build_engine_from_state_dict(cfg.model_type, hf_cfg, sd, ecfg, dtype=dtype)
        self._next_uid = 0
        log_dist(f"kernel injection: {cfg.model_type} mapped onto the ragged transformer ({dtype}, tp={tp_size})", ranks=[0])

    # ------------------------------------------------------------------ helpers
    def _uids(self, n):
        out = list(range(self._next_uid, self._next_uid + n))
        self._next_uid += n
        return out

    def _put_chunked(self, uids, seqs):
        """Prefill prompts of arbitrary length through the engine's token budget; returns last-token logits [n, V]."""
        budget = self.engine._config.state_manager.max_ragged_batch_size
        logits = [None] * len(uids)
        for i, (uid, toks) in enumerate(zip(uids, seqs)):
            for start in range(0, len(toks), budget):
                out = self.engine.put([uid], [toks[start:start + budget]])
            logits[i] = out[0]
        return torch.stack(logits)

    @staticmethod
    def _sample(logits, do_sample, temperature, top_k, top_p):
        if not do_sample:
            return logits.argmax(-1)
        logits = logits.float() / max(temperature, 1e-5)
        if top_k and top_k > 0:
            kth = logits.topk(min(top_k, logits.shape[-1]), dim=-1).values[..., -1:]
            logits = logits.masked_fill(logits < kth, float("-inf"))
        if top_p and top_p < 1.0:
            sorted_logits, idx = logits.sort(dim=-1, descending=True)
            probs = sorted_logits.softmax(-1)
            drop = probs.cumsum(-1) - probs > top_p
            sorted_logits = sorted_logits.masked_fill(drop, float("-inf"))
            logits = torch.full_like(logits, float("-inf")).scatter(-1, idx, sorted_logits)
        return torch.multinomial(logits.softmax(-1), 1).squeeze(-1)

    # ------------------------------------------------------------------ public surface
    @torch.no_grad()
    def forward(self, input_ids, attention_mask=None, **kwargs):
        """Logits of the LAST position of every sequence, shape [B, 1, V] (stateless: the KV blocks are released)."""
        seqs = self._strip_padding(input_ids, attention_mask)
        uids = self._uids(len(seqs))
        logits = self._put_chunked(uids, seqs)
        for uid in uids:
            self.engine.flush(uid)
        return SimpleNamespace(logits=logits.unsqueeze(1))

    @staticmethod
    def _strip_padding(input_ids, attention_mask):
        if attention_mask is None:
            return [row for row in input_ids]
        return [row[mask.bool()] for row, mask in zip(input_ids, attention_mask)]

    @torch.no_grad()
    def generate(self, input_ids=None, attention_mask=None, max_new_tokens=None, max_length=None, do_sample=False, temperature=1.0, top_k=0,
                 top_p=1.0, eos_token_id=None, pad_token_id=None, **kwargs):
        """Greedy / sampled decoding with the blocked KV cache; all sequences of the batch decode together (ragged batch),
        finished ones leave the batch.  Returns [B, prompt + new] padded with pad_token_id (default eos or 0)."""
        if input_ids is None:
            input_ids = kwargs.pop("inputs")
        seqs = self._strip_padding(input_ids, attention_mask)
        B = len(seqs)
        if max_new_tokens is None:
            max_new_tokens = (max_length - max(len(s) for s in seqs)) if max_length else 20
        eos = set(eos_token_id) if isinstance(eos_token_id, (list, tuple)) else ({eos_token_id} if eos_token_id is not None else set())
        uids = self._uids(B)
        logits = self._put_chunked(uids, seqs)
        outs = [s.tolist() for s in seqs]
        live = list(range(B))
        for _ in range(max_new_tokens):
            nxt = self._sample(logits, do_sample, temperature, top_k, top_p)
            still, toks = [], []
            for j, i in enumerate(live):
                t = int(nxt[j])
                outs[i].append(t)
                if t in eos:
                    self.engine.flush(uids[i])
                else:
                    still.append(i)
                    toks.append(nxt[j:j + 1])
            live = still
            if not live or _ == max_new_tokens - 1:
                break
            logits = self.engine.put([uids[i] for i in live], toks)
        for i in live:
            self.engine.flush(uids[i])
        pad = pad_token_id if pad_token_id is not None else (next(iter(eos)) if eos else 0)
        width = max(len(o) for o in outs)
        out = torch.full((B, width), pad, dtype=input_ids.dtype, device=input_ids.device)
        for i, o in enumerate(outs):
            out[i, :len(o)] = torch.tensor(o, dtype=input_ids.dtype, device=input_ids.device)
        return out
