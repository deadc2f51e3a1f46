"""Splitting of fused QKV weights for tensor parallelism (parity: deepspeed/module_inject/fusedqkv_utils.py):
a fused [3*h, in] (or [h_q + 2 h_kv, in]) projection must be sliced per q/k/v block so every rank keeps matching heads."""
import torch


def split_by_qkvlist_and_refuse(qkv_list, split_size, split_dim=0, cat_dim=0):
    qkv_split_list = [torch.split(mat, split_size, dim=split_dim) for mat in qkv_list]
    return [torch.cat([qkv_s[i] for qkv_s in qkv_split_list], dim=cat_dim) for i in range(len(qkv_split_list[0]))]


def require_tp_fused_qkvw(name, mp_size):
    fused_qkvw_name_list = ["qkv_proj", "query_key_value", "attn.Wqkv", "self_attn.W_pack", "c_attn"]
    if mp_size == 1:
        return False
    return any(f in name for f in fused_qkvw_name_list)


def prepare_tp_fused_qkvw(module_str, src, mp_size, gpu_index, num_q_heads=None, num_kv_heads=None, head_dim=None):
    """Return this rank's slice of a fused QKV weight/bias `src` ([out, in] or [out])."""
    if src is None:
        return None
    out = src.shape[0]
    if num_q_heads is not None and num_kv_heads is not None and head_dim is not None and num_q_heads != num_kv_heads:
        q, k, v = torch.split(src, [num_q_heads * head_dim, num_kv_heads * head_dim, num_kv_heads * head_dim], dim=0)
        return torch.cat([q.chunk(mp_size, 0)[gpu_index], k.chunk(mp_size, 0)[gpu_index], v.chunk(mp_size, 0)[gpu_index]], 0)
    if "bloom" in module_str.lower() or "glm" in module_str.lower():
        # [heads, 3, d] interleaved: slicing heads keeps q,k,v of the same head together
        return src.chunk(mp_size, 0)[gpu_index]
    q, k, v = src.chunk(3, dim=0)
    return torch.cat([q.chunk(mp_size, 0)[gpu_index], k.chunk(mp_size, 0)[gpu_index], v.chunk(mp_size, 0)[gpu_index]], 0)
