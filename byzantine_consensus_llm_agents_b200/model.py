"""Model descriptions and weight packing for the engine.

Architectures are the ones the reference's MODEL_PRESETS name (config.py:20-25) with the
dimensions of SURVEY.md §8d.  Weights are random-init (there are no checkpoints offline):
N(0, 0.02) linears/embeddings, norm weights 1 — transformers' default init — either taken from
a transformers state_dict (parity tests: same values as the CPU oracle) or generated directly
on the GPU (bench).  All matrices are bf16; norm weights and the RoPE table are fp32.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import torch

from . import lib as L


@dataclass(frozen=True)
class ModelConfig:
    name: str
    n_layers: int
    d_model: int
    n_heads: int
    n_kv_heads: int
    d_ff: int
    vocab: int
    head_dim: int = 128
    rms_eps: float = 1e-6
    rope_theta: float = 1e6
    qk_norm: bool = True          # Qwen3 per-head q/k RMSNorm; Mistral has none
    tie_embeddings: bool = False
    family: str = "qwen3"         # chat-template / tokenizer family
    max_pos: int = 8192

    @property
    def qkv_dim(self) -> int:
        return (self.n_heads + 2 * self.n_kv_heads) * self.head_dim

    @property
    def kv_bytes_per_token(self) -> int:
        return 2 * self.n_layers * self.n_kv_heads * self.head_dim * 2

    def shard(self, tp_size: int) -> "ModelConfig":
        """This rank's view under tensor parallelism: heads, d_ff and vocabulary divided."""
        if tp_size == 1:
            return self
        for what, n in (("n_heads", self.n_heads), ("n_kv_heads", self.n_kv_heads), ("vocab", self.vocab)):
            if n % tp_size:
                raise ValueError(f"{what}={n} is not divisible by tensor_parallel_size={tp_size}")
        if (self.d_ff // 16) % tp_size:
            raise ValueError(f"d_ff={self.d_ff} cannot be split {tp_size} ways in 16-row blocks")
        return ModelConfig(**{**self.__dict__, "n_heads": self.n_heads // tp_size,
                              "n_kv_heads": self.n_kv_heads // tp_size, "d_ff": self.d_ff // tp_size,
                              "vocab": self.vocab // tp_size, "name": f"{self.name}/tp{tp_size}"})

    def layer_weight_bytes(self) -> int:
        d, f = self.d_model, self.d_ff
        per = self.qkv_dim * d + d * self.n_heads * self.head_dim + 2 * f * d + d * f
        return 2 * per

    def decode_weight_bytes(self) -> int:
        """bf16 bytes streamed per decode step: all layer matrices + lm_head (SURVEY §8d col 9)."""
        return self.n_layers * self.layer_weight_bytes() + 2 * self.vocab * self.d_model

    def prefill_gemm_flops_per_token(self) -> int:
        return self.n_layers * self.layer_weight_bytes()  # 2 FLOP per bf16 weight = bytes


PRESETS = {
    "qwen3-0.6b": ModelConfig("qwen3-0.6b", 28, 1024, 16, 8, 3072, 151936, tie_embeddings=True),
    "qwen3-8b": ModelConfig("qwen3-8b", 36, 4096, 32, 8, 12288, 151936),
    "qwen3-14b": ModelConfig("qwen3-14b", 40, 5120, 40, 8, 17408, 151936),
    "qwen3-32b": ModelConfig("qwen3-32b", 64, 5120, 64, 8, 25600, 151936),
    # Mistral-Small-Instruct-2409's config.json: rms_norm_eps 1e-05 (Qwen3's is 1e-06)
    "mistral-22b": ModelConfig("mistral-22b", 56, 6144, 48, 8, 16384, 32768, rms_eps=1e-5, qk_norm=False,
                               family="mistral"),
}

_HF_NAMES = {
    "qwen/qwen3-0.6b": "qwen3-0.6b",
    "qwen/qwen3-8b": "qwen3-8b",
    "qwen/qwen3-14b": "qwen3-14b",
    "qwen/qwen3-32b": "qwen3-32b",
    "mistralai/mistral-small-instruct-2409": "mistral-22b",
}


def config_for(model_name: str) -> ModelConfig:
    key = model_name.lower()
    if key in PRESETS:
        return PRESETS[key]
    if key in _HF_NAMES:
        return PRESETS[_HF_NAMES[key]]
    raise KeyError(f"no architecture known for model {model_name!r}")


def config_from_dir(model_dir: str) -> ModelConfig:
    """ModelConfig from a local HuggingFace model directory's config.json (the file vLLM reads when
    the reference names a model, vllm_agent.py:126-144).  Dense Qwen3 / Qwen2 / Llama / Mistral-style
    decoders with head_dim 128 are what the kernels support."""
    import json
    from pathlib import Path
    c = json.loads((Path(model_dir) / "config.json").read_text())
    mt = str(c.get("model_type", "")).lower()
    heads = int(c["num_attention_heads"])
    head_dim = int(c.get("head_dim") or c["hidden_size"] // heads)
    if head_dim != 128:
        raise ValueError(f"{model_dir}: head_dim {head_dim} is not supported (the attention kernels are built for 128)")
    if c.get("num_local_experts") or c.get("num_experts"):
        raise ValueError(f"{model_dir}: mixture-of-experts models are not supported")
    theta = c.get("rope_theta") or (c.get("rope_parameters") or {}).get("rope_theta") or 1e6
    return ModelConfig(name=Path(model_dir).name, n_layers=int(c["num_hidden_layers"]), d_model=int(c["hidden_size"]),
                       n_heads=heads, n_kv_heads=int(c.get("num_key_value_heads") or heads),
                       d_ff=int(c["intermediate_size"]), vocab=int(c["vocab_size"]), head_dim=head_dim,
                       rms_eps=float(c.get("rms_norm_eps", 1e-6)), rope_theta=float(theta),
                       qk_norm=mt.startswith("qwen3"), tie_embeddings=bool(c.get("tie_word_embeddings", False)),
                       family="mistral" if ("mistral" in mt or "llama" in mt) else "qwen3",
                       max_pos=min(int(c.get("max_position_embeddings", 8192)), 32768))


def rope_table(cfg: ModelConfig, device) -> torch.Tensor:
    """[max_pos, 128] fp32: cos(pos * inv_freq)[0:64] | sin(...)[0:64], computed as transformers
    does (modeling_qwen3.py Qwen3RotaryEmbedding: inv_freq fp32, outer product fp32)."""
    dim = cfg.head_dim
    inv_freq = 1.0 / (cfg.rope_theta ** (torch.arange(0, dim, 2, dtype=torch.int64).float() / dim))
    pos = torch.arange(cfg.max_pos, dtype=torch.float32)
    freqs = pos[:, None] * inv_freq[None, :]
    return torch.cat([freqs.cos(), freqs.sin()], dim=1).contiguous().to(device)


def _i64(x: int) -> int:
    """Two's-complement view of a 64-bit constant (torch has no uint64 arithmetic)."""
    x &= 0xFFFFFFFFFFFFFFFF
    return x - (1 << 64) if x >= (1 << 63) else x


def _lsr(x: torch.Tensor, k: int) -> torch.Tensor:
    """Logical shift right of int64 lanes (torch's >> is arithmetic)."""
    return (x >> k) & ((1 << (64 - k)) - 1)


def _splitmix64(x: torch.Tensor) -> torch.Tensor:
    x = x + _i64(0x9E3779B97F4A7C15)
    x = (x ^ _lsr(x, 30)) * _i64(0xBF58476D1CE4E5B9)
    x = (x ^ _lsr(x, 27)) * _i64(0x94D049BB133111EB)
    return x ^ _lsr(x, 31)


def hashed_normal(rows: int, cols: int, seed: int, tensor_id: int, device, std: float = 0.02,
                  row0: int = 0, n_rows: int = None) -> torch.Tensor:
    """bf16 [n_rows, cols] slice (rows row0..) of a pseudo-normal N(0, std) matrix that is a pure
    function of (seed, tensor_id, row, col) and BIT-IDENTICAL on every device: element (r, c) hashes
    the counter r*cols+c with splitmix64 under two keys, sums the eight 16-bit fields (Irwin-Hall,
    integer arithmetic only), and scales the centred integer by one fp32 multiply.  Lets a CPU
    oracle stream a full-depth model layer by layer over exactly the weights a GPU holds, without
    either side storing or shipping them (tests/test_gpu_model.py full-depth check)."""
    n_rows = rows - row0 if n_rows is None else n_rows
    out = torch.empty(n_rows, cols, dtype=torch.bfloat16, device=device)
    key = _i64(seed * 0xD1342543DE82EF95 + tensor_id * 0xA0761D6478BD642F + 0x2545F4914F6CDD1D)
    scale = std / (8.0 * (2.0 ** 32 - 1.0) / 12.0) ** 0.5      # eight uniform 16-bit fields
    chunk = max(1, (1 << 24) // cols)
    for r in range(0, n_rows, chunk):
        m = min(chunk, n_rows - r)
        idx = torch.arange((row0 + r) * cols, (row0 + r + m) * cols, dtype=torch.int64, device=device)
        tot = torch.zeros_like(idx)
        for k in (0, 1):
            h = _splitmix64(idx ^ (key + k * _i64(0x632BE59BD9B4E019)))
            for sh in (0, 16, 32, 48):
                tot += _lsr(h, sh) & 0xFFFF
        tot -= 4 * 0xFFFF
        out[r:r + m] = (tot.to(torch.float32) * scale).view(m, cols).to(torch.bfloat16)
    return out


# HF parameter names of one decoder layer and the ids hashed_normal draws them under
_HASHED_LAYER_TENSORS = ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj", "self_attn.o_proj",
                         "mlp.gate_proj", "mlp.up_proj", "mlp.down_proj")


def hashed_state_dict_layer(cfg: "ModelConfig", layer: int, seed: int, device) -> dict:
    """transformers-named bf16 weights of decoder layer `layer` under the hashed init."""
    d, hd = cfg.d_model, cfg.head_dim
    shapes = {"self_attn.q_proj": (cfg.n_heads * hd, d), "self_attn.k_proj": (cfg.n_kv_heads * hd, d),
              "self_attn.v_proj": (cfg.n_kv_heads * hd, d), "self_attn.o_proj": (d, cfg.n_heads * hd),
              "mlp.gate_proj": (cfg.d_ff, d), "mlp.up_proj": (cfg.d_ff, d), "mlp.down_proj": (d, cfg.d_ff)}
    return {name: hashed_normal(*shapes[name], seed, 16 * (layer + 1) + j, device)
            for j, name in enumerate(_HASHED_LAYER_TENSORS)}


def hashed_embed(cfg: "ModelConfig", seed: int, device, which: str = "embed") -> torch.Tensor:
    return hashed_normal(cfg.vocab, cfg.d_model, seed, 1 if which == "embed" else 2, device)


def interleave_gate_up(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """[F,d],[F,d] -> [2F,d] with rows laid out [16 gate | 16 up] repeating (GEMM SwiGLU epilogue)."""
    f, d = gate.shape
    assert f % 16 == 0
    return torch.stack([gate.view(f // 16, 16, d), up.view(f // 16, 16, d)], dim=1).reshape(2 * f, d)


@dataclass
class Weights:
    cfg: ModelConfig
    embed: torch.Tensor
    lm_head: torch.Tensor
    final_norm: torch.Tensor
    ln1: list = field(default_factory=list)
    ln2: list = field(default_factory=list)
    w_qkv: list = field(default_factory=list)
    w_o: list = field(default_factory=list)
    w_gate_up: list = field(default_factory=list)
    w_down: list = field(default_factory=list)
    q_norm: list = field(default_factory=list)
    k_norm: list = field(default_factory=list)
    cos_sin: torch.Tensor = None

    @staticmethod
    def from_hf_state_dict(cfg: ModelConfig, sd: dict, device) -> "Weights":
        def bf(t):
            return t.detach().to(torch.bfloat16).contiguous().to(device)

        def f32(t):
            return t.detach().float().contiguous().to(device)

        embed = bf(sd["model.embed_tokens.weight"])
        lm = embed if cfg.tie_embeddings or "lm_head.weight" not in sd else bf(sd["lm_head.weight"])
        w = Weights(cfg, embed, lm, f32(sd["model.norm.weight"]))
        for l in range(cfg.n_layers):
            w._append_layer(sd, f"model.layers.{l}.", device)
        w.cos_sin = rope_table(cfg, device)
        return w

    def _append_layer(self, sd, p: str, device) -> None:
        """Pack one decoder layer's transformers-named tensors (prefix p) into the engine's layout."""
        def bf(t):
            return t.detach().to(torch.bfloat16).contiguous().to(device)

        def f32(t):
            return t.detach().float().contiguous().to(device)

        self.ln1.append(f32(sd[p + "input_layernorm.weight"]))
        self.ln2.append(f32(sd[p + "post_attention_layernorm.weight"]))
        self.w_qkv.append(bf(torch.cat([sd[p + "self_attn.q_proj.weight"],
                                        sd[p + "self_attn.k_proj.weight"],
                                        sd[p + "self_attn.v_proj.weight"]], dim=0)))
        self.w_o.append(bf(sd[p + "self_attn.o_proj.weight"]))
        self.w_gate_up.append(bf(interleave_gate_up(sd[p + "mlp.gate_proj.weight"],
                                                    sd[p + "mlp.up_proj.weight"])))
        self.w_down.append(bf(sd[p + "mlp.down_proj.weight"]))
        if self.cfg.qk_norm:
            self.q_norm.append(f32(sd[p + "self_attn.q_norm.weight"]))
            self.k_norm.append(f32(sd[p + "self_attn.k_norm.weight"]))

    @staticmethod
    def from_safetensors_dir(cfg: ModelConfig, model_dir: str, device) -> "Weights":
        """Checkpoint of a local HuggingFace model directory (model.safetensors, or the shards listed in
        model.safetensors.index.json), read tensor by tensor and packed exactly as from_hf_state_dict
        packs a state dict — never more than one decoder layer in host memory."""
        import json
        from pathlib import Path

        from safetensors import safe_open
        d = Path(model_dir)
        index = d / "model.safetensors.index.json"
        if index.exists():
            where = json.loads(index.read_text())["weight_map"]
        else:
            with safe_open(str(d / "model.safetensors"), "pt") as f:
                where = {k: "model.safetensors" for k in f.keys()}
        handles = {}

        def get(name):
            fn = where[name]
            if fn not in handles:
                handles[fn] = safe_open(str(d / fn), "pt")
            return handles[fn].get_tensor(name)

        class Lazy(dict):
            def __missing__(self, key):
                return get(key)

            def __contains__(self, key):
                return key in where

        # top-level tensors, then one layer at a time through the same packing code
        sd = Lazy()
        w = Weights.from_hf_state_dict(ModelConfig(**{**cfg.__dict__, "n_layers": 0}), sd, device)
        w.cfg = cfg
        for l in range(cfg.n_layers):
            w._append_layer(sd, f"model.layers.{l}.", device)
        return w

    def shard(self, rank: int, size: int) -> "Weights":
        """Column-parallel qkv / gate_up, row-parallel o / down, vocab-parallel lm_head; norms, the
        embedding and the RoPE table are replicated (what vLLM does for the reference when
        tensor_parallel_size > 1).  Returns contiguous copies of this rank's slices."""
        if size == 1:
            return self
        c, lc = self.cfg, self.cfg.shard(size)
        hq, hkv, f, hd = lc.n_heads, lc.n_kv_heads, lc.d_ff, c.head_dim
        qo, ko, vo = 0, c.n_heads * hd, (c.n_heads + c.n_kv_heads) * hd
        w = Weights(lc, self.embed, self.lm_head[rank * lc.vocab:(rank + 1) * lc.vocab].contiguous(),
                    self.final_norm, cos_sin=self.cos_sin)
        for l in range(c.n_layers):
            qkv = self.w_qkv[l]
            w.w_qkv.append(torch.cat([qkv[qo + rank * hq * hd: qo + (rank + 1) * hq * hd],
                                      qkv[ko + rank * hkv * hd: ko + (rank + 1) * hkv * hd],
                                      qkv[vo + rank * hkv * hd: vo + (rank + 1) * hkv * hd]]).contiguous())
            w.w_o.append(self.w_o[l][:, rank * hq * hd:(rank + 1) * hq * hd].contiguous())
            w.w_gate_up.append(self.w_gate_up[l][2 * rank * f: 2 * (rank + 1) * f].contiguous())
            w.w_down.append(self.w_down[l][:, rank * f:(rank + 1) * f].contiguous())
        w.ln1, w.ln2, w.q_norm, w.k_norm = self.ln1, self.ln2, self.q_norm, self.k_norm
        return w

    @staticmethod
    def random_init_sharded(cfg: ModelConfig, device, seed: int, rank: int, size: int) -> "Weights":
        """random_init(...).shard(rank, size) without ever holding more than one full matrix:
        every rank draws the same full tensors in the same order and keeps its slice."""
        if size == 1:
            return Weights.random_init(cfg, device, seed)
        g = torch.Generator(device=device)
        g.manual_seed(seed)
        lc = cfg.shard(size)
        hd = cfg.head_dim

        def lin(n, k):
            t = torch.empty(n, k, dtype=torch.bfloat16, device=device)
            t.normal_(0.0, 0.02, generator=g)
            return t

        ones = lambda n: torch.ones(n, dtype=torch.float32, device=device)
        embed = lin(cfg.vocab, cfg.d_model)
        full_head = embed if cfg.tie_embeddings else lin(cfg.vocab, cfg.d_model)
        w = Weights(lc, embed, full_head[rank * lc.vocab:(rank + 1) * lc.vocab].contiguous(), ones(cfg.d_model))
        del full_head
        one = Weights(ModelConfig(**{**cfg.__dict__, "n_layers": 1}), embed, embed, w.final_norm)
        for _ in range(cfg.n_layers):
            w.ln1.append(ones(cfg.d_model))
            w.ln2.append(ones(cfg.d_model))
            one.w_qkv, one.w_o = [lin(cfg.qkv_dim, cfg.d_model)], [lin(cfg.d_model, cfg.n_heads * hd)]
            one.w_gate_up, one.w_down = [lin(2 * cfg.d_ff, cfg.d_model)], [lin(cfg.d_model, cfg.d_ff)]
            sh = one.shard(rank, size)
            w.w_qkv += sh.w_qkv
            w.w_o += sh.w_o
            w.w_gate_up += sh.w_gate_up
            w.w_down += sh.w_down
            if cfg.qk_norm:
                w.q_norm.append(ones(hd))
                w.k_norm.append(ones(hd))
        w.cos_sin = rope_table(cfg, device)
        return w

    @staticmethod
    def hashed_init(cfg: ModelConfig, device, seed: int = 0) -> "Weights":
        """Full model under the device-independent hashed init (hashed_normal): norm weights 1, every
        matrix a pure function of (seed, tensor, row, col).  Generated on the device, one layer at a
        time; a CPU process can regenerate any layer bit for bit (oracle/tools/make_full_depth_golden.py)."""
        ones = lambda n: torch.ones(n, dtype=torch.float32, device=device)
        embed = hashed_embed(cfg, seed, device)
        lm = embed if cfg.tie_embeddings else hashed_embed(cfg, seed, device, "lm_head")
        w = Weights(cfg, embed, lm, ones(cfg.d_model))
        for l in range(cfg.n_layers):
            sd = hashed_state_dict_layer(cfg, l, seed, device)
            w.ln1.append(ones(cfg.d_model))
            w.ln2.append(ones(cfg.d_model))
            w.w_qkv.append(torch.cat([sd["self_attn.q_proj"], sd["self_attn.k_proj"], sd["self_attn.v_proj"]]))
            w.w_o.append(sd["self_attn.o_proj"])
            w.w_gate_up.append(interleave_gate_up(sd["mlp.gate_proj"], sd["mlp.up_proj"]).contiguous())
            w.w_down.append(sd["mlp.down_proj"])
            if cfg.qk_norm:
                w.q_norm.append(ones(cfg.head_dim))
                w.k_norm.append(ones(cfg.head_dim))
            del sd
        w.cos_sin = rope_table(cfg, device)
        return w

    @staticmethod
    def random_init(cfg: ModelConfig, device, seed: int = 0) -> "Weights":
        """transformers-default init generated on the device (bench: 14B would need 59 GB of
        host fp32 otherwise).  Same distribution, different stream than the CPU oracle."""
        g = torch.Generator(device=device)
        g.manual_seed(seed)

        def lin(n, k):
            t = torch.empty(n, k, dtype=torch.bfloat16, device=device)
            t.normal_(0.0, 0.02, generator=g)
            return t

        ones = lambda n: torch.ones(n, dtype=torch.float32, device=device)
        embed = lin(cfg.vocab, cfg.d_model)
        lm = embed if cfg.tie_embeddings else lin(cfg.vocab, cfg.d_model)
        w = Weights(cfg, embed, lm, ones(cfg.d_model))
        for _ in range(cfg.n_layers):
            w.ln1.append(ones(cfg.d_model))
            w.ln2.append(ones(cfg.d_model))
            w.w_qkv.append(lin(cfg.qkv_dim, cfg.d_model))
            w.w_o.append(lin(cfg.d_model, cfg.n_heads * cfg.head_dim))
            w.w_gate_up.append(lin(2 * cfg.d_ff, cfg.d_model))
            w.w_down.append(lin(cfg.d_model, cfg.d_ff))
            if cfg.qk_norm:
                w.q_norm.append(ones(cfg.head_dim))
                w.k_norm.append(ones(cfg.head_dim))
        w.cos_sin = rope_table(cfg, device)
        return w

    def nbytes(self) -> int:
        seen, total = set(), 0
        for t in [self.embed, self.lm_head, *self.w_qkv, *self.w_o, *self.w_gate_up, *self.w_down]:
            if t.data_ptr() not in seen:
                seen.add(t.data_ptr())
                total += t.numel() * t.element_size()
        return total


def _ptr_array(tensors):
    arr = (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])
    return arr


class ModelHandle:
    """Host-side bcg_model struct (include/bcg_b200.h) pointing at `Weights` + per-layer KV."""

    def __init__(self, w: Weights, kv_cache_layers: list, page_size: int = 16, tp_size: int = 1):
        cfg = w.cfg
        self.w = w
        self.kv = kv_cache_layers
        self._keep = []
        m = L.Model()
        m.n_layers, m.d_model, m.n_heads, m.n_kv_heads = cfg.n_layers, cfg.d_model, cfg.n_heads, cfg.n_kv_heads
        m.head_dim, m.d_ff, m.vocab = cfg.head_dim, cfg.d_ff, cfg.vocab
        m.rms_eps, m.page_size = cfg.rms_eps, page_size
        m.embed, m.lm_head, m.final_norm = w.embed.data_ptr(), w.lm_head.data_ptr(), w.final_norm.data_ptr()
        for name in ("ln1", "ln2", "w_qkv", "w_o", "w_gate_up", "w_down"):
            arr = _ptr_array(getattr(w, name))
            self._keep.append(arr)
            setattr(m, name, C.cast(arr, C.POINTER(C.c_void_p)))
        if cfg.qk_norm:
            for name in ("q_norm", "k_norm"):
                arr = _ptr_array(getattr(w, name))
                self._keep.append(arr)
                setattr(m, name, C.cast(arr, C.POINTER(C.c_void_p)))
        arr = _ptr_array(kv_cache_layers)
        self._keep.append(arr)
        m.kv_cache = C.cast(arr, C.POINTER(C.c_void_p))
        m.cos_sin = w.cos_sin.data_ptr()
        m.tp_size = tp_size
        self.struct = m
