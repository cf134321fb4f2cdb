"""CPU-side checks: the C-ABI library loads and exports every symbol include/bcg_b200.h declares
(no compute calls), the ctypes structs mirror the header, and the engine's host logic (page pool,
prefix-cache chain, tokenizer, seeds) behaves."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from byzantine_consensus_llm_agents_b200 import lib as L

ROOT = Path(__file__).resolve().parent.parent
HEADER = (ROOT / "include" / "bcg_b200.h").read_text()


@pytest.fixture(scope="module")
def so():
    return L.load(build_if_missing=True)


def test_every_declared_symbol_is_exported(so):
    declared = set(re.findall(r"^\s*(?:int|long long|const char\*)\s+(bcg_\w+)\s*\(", HEADER, flags=re.M))
    assert len(declared) >= 14
    assert declared == set(L.EXPORTED), declared ^ set(L.EXPORTED)
    for name in declared:
        assert hasattr(so, name), name
    assert so.bcg_abi_version() == int(re.search(r"#define BCG_ABI_VERSION (\d+)", HEADER).group(1))


def test_header_cites_reference_or_says_no_counterpart():
    # every entry point's comment names what it replaces (a vLLM op the reference exercises, a
    # reference file:line) or states that the reference has no counterpart
    blocks = re.split(r"\n\n", HEADER)
    for b in blocks:
        m = re.search(r"\bint (bcg_(?!abi_version|last_error|gemm_splits_for)\w+)\s*\(", b)
        if m and "/*" in b:
            assert re.search(r"SURVEY|vllm|Stands in|counterpart|reference", b), m.group(1)


def test_host_only_entry_points_work_without_gpu(so):
    assert so.bcg_gemm_splits_for(10, 34816, 5120) == 1          # 272 tiles fill the 296 work slots
    # Qwen3-14B decode: every projection is ONE wave of one-item CTAs on 2 x 148 slots
    for n, k in [(7168, 5120), (5120, 5120), (34816, 5120), (5120, 17408)]:
        items = -(-n // 128) * so.bcg_gemm_splits_for(10, n, k)
        assert 0.9 * 296 <= items <= 296, (n, k, items)
    assert so.bcg_gemm_splits_for(10, 5120, 5120) > 1
    assert so.bcg_gemm_splits_for(300, 5120, 5120) == 1          # prefill never splits
    # splits depend on (N, K) only -> batch-invariant decode arithmetic
    for n, k in [(7168, 5120), (5120, 17408), (1024, 3072)]:
        assert len({so.bcg_gemm_splits_for(m, n, k) for m in (1, 7, 16, 64, 128)}) == 1


def test_argument_errors_are_reported_not_crashes(so):
    rc = so.bcg_gemm_bf16_tn(None, None, None, 4, 128, 64, 64, 64, 128, 0, 1, 0, None)
    assert rc == 1 and b"bcg_gemm_bf16_tn" in so.bcg_last_error()
    rc = so.bcg_embed_rmsnorm(None, None, None, None, None, 1, 1024, 1e-6, None)
    assert rc == 1
    a = L.SamplerArgs()
    assert so.bcg_mask_sample(C.byref(a), None) == 1
    rc = so.bcg_decode_attn_paged_static(None, None, None, 1, None, None, 128, 64, 8, 16, 32, 0.1, None)
    assert rc == 4 and b"16 MiB" in so.bcg_last_error()


def test_ctypes_structs_mirror_header():
    def fields(struct_name):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (struct_name, struct_name), HEADER, flags=re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            for part in decl.split(","):
                names.append(re.findall(r"(\w+)\s*$", part.strip())[0])
        return names
    assert fields("bcg_sampler_args") == [f[0] for f in L.SamplerArgs._fields_]
    assert fields("bcg_model") == [f[0] for f in L.Model._fields_]
    assert fields("bcg_workspace") == [f[0] for f in L.Workspace._fields_]


def test_missing_library_is_loud(monkeypatch, tmp_path):
    monkeypatch.setattr(L, "_lib", None)
    monkeypatch.setattr(L, "LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(L.BcgError, match="no CPU fallback"):
        L.load()


def test_product_path_never_imports_oracle():
    pkg = ROOT / "byzantine_consensus_llm_agents_b200"
    for p in pkg.rglob("*.py"):
        src = p.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), p


# ------------------------------------------------------------------ host logic
def test_page_pool_prefix_cache_and_eviction():
    from byzantine_consensus_llm_agents_b200.engine import PagePool
    pool = PagePool(8)
    h1 = PagePool.chain(b"\0" * 16, list(range(16)))
    h2 = PagePool.chain(h1, list(range(16, 32)))
    assert h1 != h2 and PagePool.chain(b"\0" * 16, list(range(16))) == h1
    a, b = pool.alloc(), pool.alloc()
    pool.publish(a, h1)
    pool.publish(b, h2)
    pool.release(a)
    pool.release(b)
    assert pool.lookup(h1) == a and pool.lookup(h2) == b
    assert len(pool.free) == 6                      # cached pages are not on the free list
    pool.acquire(a)                                 # a second request now shares page a
    got = [pool.alloc() for _ in range(6)]
    assert a not in got and b not in got
    c = pool.alloc()                                # pool empty -> evict the unreferenced cached page b
    assert c == b and pool.lookup(h2) is None and pool.lookup(h1) == a
    with pytest.raises(MemoryError):
        for _ in range(8):
            pool.alloc()
    pool.release(a)
    pool.reset_cache()
    assert pool.lookup(h1) is None


def test_row_seed_matches_oracle_restatement():
    from byzantine_consensus_llm_agents_b200.engine import row_seed
    from oracle.engine import row_seed as oracle_seed
    seen = set()
    for base in (0, 7, 12345):
        for call in range(4):
            for row in range(5):
                s = row_seed(base, call, row)
                assert s == oracle_seed(base, call, row)
                assert 0 <= s[0] < 2 ** 32 and 0 <= s[1] < 2 ** 32
                seen.add(s)
    assert len(seen) == 60


def test_tokenizer_roundtrip_specials_and_geometry():
    from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer
    tk = get_tokenizer("qwen3", 151936)
    assert tk.special_ids["<|endoftext|>"] == 151643
    assert tk.special_ids["<|im_start|>"] == 151644
    assert tk.special_ids["<|im_end|>"] == 151645 == tk.eos_id
    text = "<|im_start|>user\nValue: 17 — naïve café 中文 {\"a\": [1, 2]} /no_think<|im_end|>\n"
    ids = tk.encode(text)
    assert ids[0] == 151644 and 151645 in ids
    assert tk.decode(ids, skip_special_tokens=False) == text
    assert tk.decode(ids) == text.replace("<|im_start|>", "").replace("<|im_end|>", "")
    assert tk.decode([151700, 151935]) == ""        # unused ids at the top decode to nothing
    blob, offs = tk.byte_table()
    assert offs.shape == (151937,) and offs[-1] == blob.shape[0]
    assert len({tk.token_bytes[i] for i in range(*tk.regular_range)}) == tk.n_regular   # all distinct
    # determinism: a second build gives the same vocabulary
    from byzantine_consensus_llm_agents_b200.tokenizer import SyntheticTokenizer
    tk2 = SyntheticTokenizer("qwen3", 151936)
    assert tk2.token_bytes == tk.token_bytes
    m = get_tokenizer("mistral", 32768)
    assert m.encode("<s>[INST] hi [/INST]")[:3] == [m.bos_id, m.special_ids["<s>"], m.special_ids["[INST]"]]


def test_reference_prompts_tokenise_at_realistic_density():
    import gzip
    import json
    from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer
    tk = get_tokenizer("qwen3", 151936)
    rounds = json.load(gzip.open(ROOT / "tests" / "golden" / "bench_prompts_qwen3_8h2b.json.gz", "rt"))["rounds"]
    s, u, _ = rounds[0]["decide"]["prompts"][0]
    n = len(tk.encode(s + u))
    assert 3.0 <= len(s + u) / n <= 5.5             # ~4 chars/token, ~1.7k tokens
    assert 1200 <= n <= 2300


def test_cached_encoder_returns_the_tokenizers_ids_and_evicts_oldest_first():
    """Retries resubmit identical prompts: the engine's prompt memo must be invisible in the ids."""
    from byzantine_consensus_llm_agents_b200.tokenizer import CachedEncoder, get_tokenizer
    tk = get_tokenizer("qwen3", 151936)
    enc = CachedEncoder(tk, capacity=3)
    prompts = [f"<|im_start|>user\nround {i}: propose a value /no_think<|im_end|>\n" for i in range(5)]
    for p in prompts[:3]:
        assert enc(p) == tk.encode(p, True)
    assert (enc.hits, enc.misses) == (0, 3)
    assert enc(prompts[0]) is enc(prompts[0]) and enc.hits == 2     # same shared list, no re-encode
    enc(prompts[3])                                                  # evicts prompts[0], the oldest
    assert enc.misses == 4
    enc(prompts[1])
    assert enc.hits == 3
    enc(prompts[0])
    assert enc.misses == 5 and enc(prompts[0]) == tk.encode(prompts[0], True)
    m = get_tokenizer("mistral", 32768)
    assert CachedEncoder(m)("hi")[0] == m.bos_id                     # add_special_tokens=True, like vLLM
    enc.clear()
    assert enc(prompts[1]) == tk.encode(prompts[1], True) and enc.misses == 6


def test_gemm_weight_interleave_layout():
    import torch
    from byzantine_consensus_llm_agents_b200.model import interleave_gate_up
    g = torch.arange(64).float()[:, None].repeat(1, 2)
    u = -torch.arange(64).float()[:, None].repeat(1, 2)
    w = interleave_gate_up(g, u)
    assert w.shape == (128, 2)
    assert w[:16, 0].tolist() == list(range(16)) and w[16:32, 0].tolist() == [-float(i) for i in range(16)]
    assert w[32:48, 0].tolist() == list(range(16, 32))


def test_model_presets_reproduce_published_parameter_counts():
    from byzantine_consensus_llm_agents_b200.model import PRESETS

    def params(c):
        per = c.qkv_dim * c.d_model + c.d_model * c.n_heads * 128 + 3 * c.d_ff * c.d_model + 2 * c.d_model \
            + (256 if c.qk_norm else 0)
        emb = c.vocab * c.d_model * (1 if c.tie_embeddings else 2)
        return c.n_layers * per + emb + c.d_model
    want = {"qwen3-0.6b": 0.596e9, "qwen3-8b": 8.191e9, "qwen3-14b": 14.768e9, "qwen3-32b": 32.762e9,
            "mistral-22b": 22.247e9}
    for k, v in want.items():
        assert abs(params(PRESETS[k]) - v) / v < 2e-3, (k, params(PRESETS[k]))
    assert abs(PRESETS["qwen3-14b"].decode_weight_bytes() - 27.98e9) < 0.02e9
    assert PRESETS["qwen3-14b"].kv_bytes_per_token == 163840
