from .dropping_utils import GatherTokens, ScatterTokens, bert_sample_tokens, gpt_sample_tokens, token_gather, token_scatter_, token_sort_
