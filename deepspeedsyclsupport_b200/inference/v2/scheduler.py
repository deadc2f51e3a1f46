"""Dynamic SplitFuse scheduling on top of InferenceEngineV2: long prompts are split into chunks and fused with the
decode tokens of running sequences so that every forward pass runs at (about) the same token budget.  The reference
describes the policy (blogs/deepspeed-fastgen/README.md:57-105) but implements it in DeepSpeed-MII; it is provided
here so the engine is usable stand-alone."""
from collections import deque

import torch


class _Request:

    def __init__(self, uid, prompt, max_new_tokens):
        self.uid, self.prompt, self.max_new_tokens = uid, prompt, max_new_tokens
        self.pos = 0  # prompt tokens already fed
        self.generated = []
        self.next_token = None

    @property
    def prefill_done(self):
        return self.pos >= self.prompt.numel()

    @property
    def done(self):
        return len(self.generated) >= self.max_new_tokens


class SplitFuseScheduler:

    def __init__(self, engine, token_budget=None, sample_fn=None):
        self.engine = engine
        self.budget = token_budget or engine._config.state_manager.max_ragged_batch_size
        self.sample = sample_fn or (lambda logits: logits.argmax(-1))
        self.waiting, self.running, self.finished = deque(), [], {}

    def add_request(self, uid, prompt_tokens, max_new_tokens=16):
        self.waiting.append(_Request(uid, torch.as_tensor(prompt_tokens).reshape(-1), max_new_tokens))

    def step(self):
        """Compose one ragged batch: all running decodes (1 token each) + prompt chunks up to the token budget."""
        uids, toks, reqs = [], [], []
        budget = self.budget
        max_seqs = self.engine._config.state_manager.max_ragged_sequence_count
        free_blocks = self.engine.free_blocks
        for r in self.running:
            if r.prefill_done and budget > 0 and len(uids) < max_seqs:
                n, blocks = self.engine.query(r.uid, 1, free_blocks)
                if n == 1:
                    uids.append(r.uid); toks.append(r.next_token.reshape(1)); reqs.append((r, 1)); budget -= 1; free_blocks -= blocks
        candidates = [r for r in self.running if not r.prefill_done] + list(self.waiting)
        for r in candidates:
            if budget <= 0 or len(uids) >= max_seqs:
                break
            want = min(budget, r.prompt.numel() - r.pos)
            n, blocks = self.engine.query(r.uid, want, free_blocks)
            if n <= 0:
                continue
            uids.append(r.uid); toks.append(r.prompt[r.pos:r.pos + n]); reqs.append((r, n)); budget -= n; free_blocks -= blocks
            if r in self.waiting:
                self.waiting.remove(r)
                self.running.append(r)
        if not uids:
            return 0
        logits = self.engine.put(uids, toks)
        nxt = self.sample(logits)
        for (r, n), tok in zip(reqs, nxt):
            if not r.prefill_done:
                r.pos += n
                if r.prefill_done:
                    r.next_token = tok.detach().cpu()
                    r.generated.append(int(tok))
            else:
                r.next_token = tok.detach().cpu()
                r.generated.append(int(tok))
        for r in list(self.running):
            if r.prefill_done and r.done:
                self.running.remove(r)
                self.finished[r.uid] = r.generated
                self.engine.flush(r.uid)
        return len(uids)

    def run_to_completion(self, max_steps=100000):
        steps = 0
        while (self.waiting or self.running) and steps < max_steps:
            if self.step() == 0:
                break
            steps += 1
        return self.finished
