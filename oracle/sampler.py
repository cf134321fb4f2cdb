"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement of the sampler the reference reaches through vLLM, in numpy:

  * logits -> fp32                                   vllm/v1/sample/sampler.py:91
  * grammar bitmask applied before sampling          vllm/v1/structured_output/utils.py:44-122
    (int32 [B, ceil(V/32)], bit = 1 allowed, LSB first; xgrammar.allocate_token_bitmask layout)
  * greedy if temperature < 1e-5 else logits / T     vllm/v1/sample/sampler.py:223-236
  * probs = softmax(logits); argmax(probs / q), q ~ Exp(1)
                                                     vllm/v1/sample/ops/topk_topp_sampler.py:129,395-416

vLLM draws q from torch's global CUDA generator, which no other implementation can
reproduce.  The engine therefore defines its own counter-based stream —
Philox4x32-10 keyed by a per-row seed, counter (i >> 2, step, 0, 0) — and evaluates the same
exponential race in the log domain:

    argmax_i softmax(l/T)_i / q_i  ==  argmax_i ( l_i / T - log(q_i) ),   q_i = -log(u_i)

`rlog` uses only IEEE-754 single-precision +, *, / (each rounded to nearest, never fused), so
this file and csrc/sampler.cu produce bit-identical scores and therefore identical token ids.
`sample_prob_domain` is the literal vLLM formulation (softmax then divide) on the same q, used
by the tests to show the two agree whenever the top-2 race margin is not a rounding tie.

Parity status: vLLM's sampler has no golden vectors in the reference repository (the
reference has no tests at all) -> "parity unpinned" against upstream for T > 0; greedy is pinned
by definition (argmax, lowest index on ties == torch.argmax on CPU).
"""
from __future__ import annotations

import numpy as np

F = np.float32
_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0: int, k1: int):
    """Vectorised Philox4x32-10.  c* are uint32 arrays (broadcastable), k* python ints."""
    c0 = np.asarray(c0, dtype=np.uint64)
    c1 = np.broadcast_to(np.asarray(c1, dtype=np.uint64), c0.shape).copy()
    c2 = np.broadcast_to(np.asarray(c2, dtype=np.uint64), c0.shape).copy()
    c3 = np.broadcast_to(np.asarray(c3, dtype=np.uint64), c0.shape).copy()
    c0 = c0.copy()
    k0 &= 0xFFFFFFFF
    k1 &= 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK32
        n0 = hi1 ^ c1 ^ np.uint64(k0)
        n2 = hi0 ^ c3 ^ np.uint64(k1)
        c0, c1, c2, c3 = n0, lo1, n2, lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return (c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32))


def rlog(x: np.ndarray) -> np.ndarray:
    """Reproducible natural log for positive normal float32 (mirrors rlog() in sampler.cu)."""
    x = np.asarray(x, dtype=F)
    bits = x.view(np.uint32)
    e = (bits >> np.uint32(23)).astype(np.int32) - 127
    m = ((bits & np.uint32(0x007FFFFF)) | np.uint32(0x3F800000)).view(F)
    big = m > F(1.41421356)
    m = np.where(big, m * F(0.5), m).astype(F)
    e = e + big.astype(np.int32)
    f = (m + F(-1.0)).astype(F)
    s = (f / (F(2.0) + f).astype(F)).astype(F)
    z = (s * s).astype(F)
    p = (F(0.22222222) + (z * F(0.18181818)).astype(F)).astype(F)
    p = (F(0.2857143) + (z * p).astype(F)).astype(F)
    p = (F(0.4) + (z * p).astype(F)).astype(F)
    p = (F(0.6666667) + (z * p).astype(F)).astype(F)
    r = ((F(2.0) * s).astype(F) + (s * (z * p).astype(F)).astype(F)).astype(F)
    return ((e.astype(F) * F(0.6931472)).astype(F) + r).astype(F)


def uniforms(V: int, step: int, seed0: int, seed1: int) -> np.ndarray:
    """u_i = (k_i + 0.5) * 2^-23 for i in [0, V) (V % 4 == 0)."""
    assert V % 4 == 0
    c0 = np.arange(V // 4, dtype=np.uint32)
    r = philox4x32_10(c0, np.uint32(step), np.uint32(0), np.uint32(0), seed0, seed1)
    rnd = np.stack(r, axis=1).reshape(-1)
    k = (rnd >> np.uint32(9)).astype(F)
    return ((k + F(0.5)).astype(F) * F(1.1920929e-07)).astype(F)


def race_scores(logits: np.ndarray, temperature: float, step: int, seed0: int, seed1: int):
    logits = np.asarray(logits, dtype=F)
    u = uniforms(logits.shape[0], step, seed0, seed1)
    q = (-rlog(u)).astype(F)
    return ((logits / F(temperature)).astype(F) + (-rlog(q)).astype(F)).astype(F), q


def unpack_mask(mask_row: np.ndarray, V: int) -> np.ndarray:
    """int32 words, LSB first -> bool [V]."""
    w = np.asarray(mask_row).view(np.uint32)
    bits = (w[:, None] >> np.arange(32, dtype=np.uint32)[None, :]) & np.uint32(1)
    return bits.reshape(-1)[:V].astype(bool)


def _argmax_low_index(score: np.ndarray) -> int:
    # np.argmax returns the first maximal index, which is the tie rule of the kernel
    return int(np.argmax(score))


def sample(logits, mask_row, temperature: float, step: int, seed0: int, seed1: int) -> int:
    """The engine's canonical sampler for one row.  Returns -1 if nothing is allowed."""
    logits = np.asarray(logits, dtype=F)
    V = logits.shape[0]
    allowed = np.ones(V, bool) if mask_row is None else unpack_mask(mask_row, V)
    if not allowed.any():
        return -1
    if temperature < 1e-5:
        score = logits.copy()
    else:
        score, _ = race_scores(logits, temperature, step, seed0, seed1)
    score = np.where(allowed, score, F(-np.inf))
    return _argmax_low_index(score)


def sample_prob_domain(logits, mask_row, temperature: float, step: int, seed0: int, seed1: int):
    """vLLM's literal formulation on the same q: softmax(l/T) / q -> argmax (float64 reference).
    Returns (token, relative margin between best and second best)."""
    logits = np.asarray(logits, dtype=np.float64)
    V = logits.shape[0]
    allowed = np.ones(V, bool) if mask_row is None else unpack_mask(mask_row, V)
    x = np.where(allowed, logits / temperature, -np.inf)
    x = x - x.max()
    p = np.exp(x)
    p /= p.sum()
    u = uniforms(V, step, seed0, seed1).astype(np.float64)
    q = -np.log(u)
    r = p / q
    order = np.argsort(-r)
    margin = (r[order[0]] - r[order[1]]) / r[order[0]] if V > 1 else 1.0
    return int(order[0]), float(margin)
