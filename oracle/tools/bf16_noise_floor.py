"""ORACLE tool: how far can ANY bf16-activation pipeline be from the fp32 forward?

Re-runs the fp32 CPU forward with only the bf16 roundings the engine performs (GEMM inputs xn /
attn / act, q, k, v, P, final xn) switched on one at a time and all together.  Measured
(wide-qwen3, 2 layers, d=1024): all roundings -> rms 0.0048, max 0.026 of a logit std 0.64; the
GPU engine measures rms 0.004 — i.e. it sits on the rounding floor.
    python -m oracle.tools.bf16_noise_floor
"""
import torch, math, numpy as np, sys
from byzantine_consensus_llm_agents_b200.model import ModelConfig, rope_table
from oracle.model import build_hf_model, full_logits
# python -m oracle.tools.bf16_noise_floor [d_model n_heads n_kv_heads d_ff n_layers]
_a = [int(x) for x in sys.argv[1:6]] + [1024, 16, 8, 3072, 2][len(sys.argv[1:6]):]
D, H, HK, F, L = _a
G = H // HK
cfg = ModelConfig("wide-qwen3", n_layers=L, d_model=D, n_heads=H, n_kv_heads=HK, d_ff=F, vocab=8192, tie_embeddings=True)
hf = build_hf_model(cfg, 0); sd = hf.state_dict()
rng = np.random.default_rng(0); ids = rng.integers(0, 7000, size=200).tolist()
ref = full_logits(hf, ids)
cs = rope_table(cfg, 'cpu')
def rms(x,w,eps=1e-6): return x*torch.rsqrt(x.pow(2).mean(-1,keepdim=True)+eps)*w
def fwd(points):
    rb = lambda name,x: x.bfloat16().float() if name in points else x
    T=len(ids); h = sd['model.embed_tokens.weight'][ids].float()
    pos = torch.arange(T)
    cos = torch.cat([cs[pos,:64],cs[pos,:64]],-1)[:,None]; sin = torch.cat([cs[pos,64:],cs[pos,64:]],-1)[:,None]
    for l in range(cfg.n_layers):
        p=f'model.layers.{l}.'
        xn = rb('xn1', rms(h, sd[p+'input_layernorm.weight']))
        q = (xn@sd[p+'self_attn.q_proj.weight'].T).view(T,H,128); k=(xn@sd[p+'self_attn.k_proj.weight'].T).view(T,HK,128); v=(xn@sd[p+'self_attn.v_proj.weight'].T).view(T,HK,128)
        q = rms(q, sd[p+'self_attn.q_norm.weight']); k = rms(k, sd[p+'self_attn.k_norm.weight'])
        rot = lambda x: torch.cat([-x[...,64:], x[...,:64]],-1)
        q = rb('q', q*cos+rot(q)*sin); k = rb('k', k*cos+rot(k)*sin); v = rb('v', v)
        kx = k.repeat_interleave(G,1); vx = v.repeat_interleave(G,1)
        s = torch.einsum('thd,shd->hts', q, kx)/math.sqrt(128)
        s = s.masked_fill(torch.triu(torch.ones(T,T,dtype=bool),1)[None], float('-inf'))
        P = torch.softmax(s,-1)
        o = torch.einsum('hts,shd->thd', rb('P',P), vx) 
        o = rb('attn', o.reshape(T,-1))
        h = h + o@sd[p+'self_attn.o_proj.weight'].T
        xn = rb('xn2', rms(h, sd[p+'post_attention_layernorm.weight']))
        g = xn@sd[p+'mlp.gate_proj.weight'].T; u = xn@sd[p+'mlp.up_proj.weight'].T
        act = rb('act', torch.nn.functional.silu(g)*u)
        h = h + act@sd[p+'mlp.down_proj.weight'].T
    xl = rb('xn_last', rms(h, sd['model.norm.weight']))
    return xl@sd['model.embed_tokens.weight'].T
print("logit std %.4f" % ref.std().item())
base = fwd(set()); print("self-consistency", (base-ref).abs().max().item())
allp = ['xn1','q','k','v','P','attn','xn2','act','xn_last']
for pset in [[p] for p in allp]+[allp]:
    d = (fwd(set(pset))-ref); print(pset, "rms %.5f max %.4f"%(d.pow(2).mean().sqrt().item(), d.abs().max().item()))
