"""CPU checks of the sampler oracle itself (oracle/sampler.py)."""
import numpy as np

from oracle import sampler as O


def test_rlog_accuracy():
    x = np.concatenate([np.float32(2.0) ** np.arange(-24, 5, dtype=np.float32),
                        np.random.default_rng(0).uniform(1e-7, 17.0, 20000).astype(np.float32)])
    got = O.rlog(x).astype(np.float64)
    ref = np.log(x.astype(np.float64))
    assert np.max(np.abs(got - ref) / np.maximum(1e-3, np.abs(ref))) < 1e-6


def test_philox_known_answer():
    # Random123 known-answer vectors for philox4x32-10
    r = O.philox4x32_10(np.array([0], np.uint32), 0, 0, 0, 0, 0)
    assert [int(v[0]) for v in r] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    r = O.philox4x32_10(np.array([0xffffffff], np.uint32), 0xffffffff, 0xffffffff, 0xffffffff,
                        0xffffffff, 0xffffffff)
    assert [int(v[0]) for v in r] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    r = O.philox4x32_10(np.array([0x243f6a88], np.uint32), 0x85a308d3, 0x13198a2e, 0x03707344,
                        0xa4093822, 0x299f31d0)
    assert [int(v[0]) for v in r] == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_uniforms_open_interval():
    u = O.uniforms(4096, 3, 123, 456)
    assert u.dtype == np.float32 and (u > 0).all() and (u < 1).all()


def test_log_domain_equals_vllm_prob_domain():
    """argmax(l/T - log q) == argmax(softmax(l/T)/q) except at rounding-level ties."""
    rng = np.random.default_rng(1)
    V, agree, checked = 2048, 0, 0
    for trial in range(200):
        logits = (rng.standard_normal(V) * 0.64).astype(np.float32)
        words = rng.integers(0, 2 ** 32, size=V // 32, dtype=np.uint64).astype(np.uint32).view(np.int32)
        T = float(rng.choice([0.3, 0.5, 1.0]))
        a = O.sample(logits, words, T, trial, 11, 22)
        b, margin = O.sample_prob_domain(logits, words, T, trial, 11, 22)
        if margin > 1e-4:
            checked += 1
            agree += (a == b)
    assert checked > 150 and agree == checked


def test_temperature_sampling_is_unbiased():
    logits = np.array([0.0, 1.0, 2.0, -1.0] + [-30.0] * 28, np.float32)
    T = 0.5
    p = np.exp(logits.astype(np.float64) / T)
    p /= p.sum()
    n = 4000
    counts = np.zeros(32)
    for s in range(n):
        counts[O.sample(logits, None, T, s, 7, 9)] += 1
    assert np.abs(counts[:4] / n - p[:4]).max() < 0.03


def test_greedy_and_mask():
    logits = np.array([1.0, 5.0, 5.0, 2.0] * 8, np.float32)
    assert O.sample(logits, None, 0.0, 0, 0, 0) == 1
    mask = np.array([0b1001], np.uint32).view(np.int32)
    assert O.sample(logits, mask, 0.0, 0, 0, 0) == 3
    assert O.sample(logits, np.array([0], np.int32), 0.0, 0, 0, 0) == -1
