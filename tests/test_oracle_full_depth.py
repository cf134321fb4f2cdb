"""The layer-streamed fp32 forward that produces the full-depth golden logits
(oracle/tools/make_full_depth_golden.py) against the ordinary whole-model transformers forward, on
models small enough to hold: pins the streaming, the hashed weight generator's slicing, and the
fixture's own consistency.  CPU only."""
import glob
from pathlib import Path

import numpy as np
import pytest
import torch

from byzantine_consensus_llm_agents_b200 import model as M
from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer
from oracle.model import build_hf_model_hashed, full_logits
from oracle.tools.make_full_depth_golden import case_inputs, streamed_logits

TINY_QWEN = M.ModelConfig("tiny-qwen3", n_layers=3, d_model=256, n_heads=4, n_kv_heads=2, d_ff=512, vocab=4096)
TINY_MISTRAL = M.ModelConfig("tiny-mistral", n_layers=2, d_model=384, n_heads=6, n_kv_heads=1, d_ff=768,
                             vocab=4096, rms_eps=1e-5, qk_norm=False, family="mistral")


@pytest.mark.parametrize("cfg", [TINY_QWEN, TINY_MISTRAL], ids=lambda c: c.name)
def test_streamed_forward_equals_whole_model_forward(cfg):
    tk = get_tokenizer(cfg.family, cfg.vocab)
    prompt, forced = case_inputs(tk, 90, 4, seed=3)
    ids = prompt + forced[:-1]
    want = full_logits(build_hf_model_hashed(cfg, seed=5), ids)[len(prompt) - 1:]
    got = streamed_logits(cfg, ids, len(prompt) - 1, weight_seed=5)
    assert got.shape == want.shape
    assert (got - want).abs().max().item() <= 2e-5      # same fp32 code, different matmul blocking


def test_hashed_normal_is_sliceable_and_normal_enough():
    full = M.hashed_normal(300, 256, 9, 33, "cpu")
    assert torch.equal(full[100:117], M.hashed_normal(300, 256, 9, 33, "cpu", row0=100, n_rows=17))
    assert not torch.equal(full, M.hashed_normal(300, 256, 9, 34, "cpu"))
    assert not torch.equal(full, M.hashed_normal(300, 256, 10, 33, "cpu"))
    x = full.float()
    assert abs(x.mean().item()) < 5e-4 and abs(x.std().item() - 0.02) < 5e-4
    kurt = (x ** 4).mean().item() / x.var().item() ** 2
    assert 2.6 < kurt < 3.1


@pytest.mark.parametrize("path", sorted(glob.glob(str(Path(__file__).parent / "golden" / "full_depth_*.npz"))),
                         ids=lambda p: Path(p).stem)
def test_full_depth_fixture_is_self_consistent(path):
    z = np.load(path)
    logits = z["logits_f16"].astype(np.float32)
    n_layers, _, seed, prompt_len, steps = z["meta"].tolist()
    cfg = M.PRESETS[str(z["model"])]
    assert logits.shape == (steps, cfg.vocab) and n_layers <= cfg.n_layers
    tk = get_tokenizer(cfg.family, cfg.vocab)
    prompt, forced = case_inputs(tk, prompt_len, steps, seed)
    assert prompt == z["prompt"].tolist() and forced == z["forced"].tolist()
    # the fp16 copy and the exact top-8 agree
    for i in range(steps):
        assert np.abs(logits[i, z["top_ids"][i]] - z["top_vals"][i]).max() <= 4e-3
        assert int(logits[i].argmax()) in z["top_ids"][i][:2].tolist()
