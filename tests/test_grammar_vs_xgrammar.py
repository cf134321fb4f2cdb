"""The engine's schema -> DFA -> token-mask compiler against xgrammar (the oracle), bit for bit,
for every state of every grammar the reference can emit."""
import json

import numpy as np
import pytest

xgr = pytest.importorskip("xgrammar")

from byzantine_consensus_llm_agents_b200 import grammar as G  # noqa: E402
from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer  # noqa: E402
from oracle.grammar_xgr import XgrOracle, shortest_prefixes  # noqa: E402
from oracle.sampler import unpack_mask  # noqa: E402


def ref_schemas(vmin=0, vmax=50):
    """The four schemas of bcg_agents.py:590-599, 651-658, 1083-1092, 1155-1162."""
    honest_decide = {"type": "object", "properties": {
        "internal_strategy": {"type": "string"},
        "value": {"type": "integer", "minimum": vmin, "maximum": vmax},
        "public_reasoning": {"type": "string"}},
        "required": ["internal_strategy", "value", "public_reasoning"], "additionalProperties": False}
    byz_decide = {"type": "object", "properties": {
        "internal_strategy": {"type": "string"},
        "value": {"anyOf": [{"type": "integer", "minimum": vmin, "maximum": vmax},
                            {"type": "string", "enum": ["abstain"]}]},
        "public_reasoning": {"type": "string"}},
        "required": ["internal_strategy", "value"], "additionalProperties": False}
    honest_vote = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue"]}},
                   "required": ["decision"], "additionalProperties": False}
    byz_vote = {"type": "object", "properties": {"decision": {"type": "string", "enum": ["stop", "continue", "abstain"]}},
                "required": ["decision"], "additionalProperties": False}
    return {"honest_decide": honest_decide, "byz_decide": byz_decide, "honest_vote": honest_vote,
            "byz_vote": byz_vote}


@pytest.fixture(scope="module")
def small():
    tk = get_tokenizer("qwen3", 8192)
    return tk, XgrOracle(tk)


def _compare_all_states(tk, oracle, schema):
    g = G.compile_schema(schema, tk)
    pre = shortest_prefixes(g.trans)
    assert all(p is not None for p in pre)
    for s, p in enumerate(pre):
        want = oracle.mask_after_bytes(schema, p)
        got = g.masks[s]
        if not np.array_equal(want, got):
            w, h = unpack_mask(want, tk.vocab_size), unpack_mask(got, tk.vocab_size)
            diff = np.nonzero(w != h)[0]
            raise AssertionError(
                f"state {s} after {p!r}: {diff.size} tokens differ, e.g. "
                f"{[(int(i), tk.token_bytes[i], bool(w[i]), bool(h[i])) for i in diff[:8]]}")
    return g


@pytest.mark.parametrize("name", ["honest_decide", "byz_decide", "honest_vote", "byz_vote"])
def test_reference_schemas_all_states_small_vocab(small, name):
    tk, oracle = small
    g = _compare_all_states(tk, oracle, ref_schemas()[name])
    assert g.n_states < 400


@pytest.mark.parametrize("vrange", [(0, 100), (10, 20), (-5, 5), (0, 9), (7, 1234)])
def test_value_range_variants(small, vrange):
    """value_range is CLI-configurable in the reference (main.py:1005-1030)."""
    tk, oracle = small
    _compare_all_states(tk, oracle, ref_schemas(*vrange)["honest_decide"])


def test_language_membership(small):
    tk, _ = small
    g = G.compile_schema(ref_schemas()["honest_decide"], tk)
    ok = ['{"internal_strategy":"a","value":0,"public_reasoning":"b"}',
          '{ "internal_strategy" : "x\\n\\u00e9 é中" ,\n\t"value": 50, "public_reasoning": "" }',
          '{"internal_strategy":"","value":49,"public_reasoning":"\\\\"}']
    bad = ['{"internal_strategy":"a","value":51,"public_reasoning":"b"}',
           '{"internal_strategy":"a","value":00,"public_reasoning":"b"}',
           '{"internal_strategy":"a","value":5}',
           '{"value":5,"internal_strategy":"a","public_reasoning":"b"}',
           '{"internal_strategy":"a\nb","value":5,"public_reasoning":"b"}',
           '{"internal_strategy":"a","value":5,"public_reasoning":"b"} ']
    for s in ok:
        st = g.walk(0, s.encode())
        assert st >= 0 and g.accept[st], s
        json.loads(s)
    for s in bad:
        st = g.walk(0, s.encode())
        assert st < 0 or not g.accept[st], s


def test_optional_property_and_abstain(small):
    tk, _ = small
    g = G.compile_schema(ref_schemas()["byz_decide"], tk)
    for s in ['{"internal_strategy":"a","value":"abstain"}',
              '{"internal_strategy":"a","value":7,"public_reasoning":"z"}']:
        st = g.walk(0, s.encode())
        assert st >= 0 and g.accept[st]
    st = g.walk(0, b'{"internal_strategy":"a","value":"stop"}')
    assert st < 0


def test_unsupported_schema_is_loud(small):
    tk, _ = small
    with pytest.raises(G.UnsupportedSchema):
        G.compile_schema({"type": "object", "properties": {"a": {"type": "string"}}}, tk)
    with pytest.raises(G.UnsupportedSchema):
        G.compile_schema({"type": "string", "pattern": "a+"}, tk)


def test_random_walks_full_vocab():
    """Full 151,936-token vocabulary: random token walks, mask compared at every step."""
    tk = get_tokenizer("qwen3", 151936)
    oracle = XgrOracle(tk)
    rng = np.random.default_rng(0)
    for name, schema in ref_schemas().items():
        g = G.compile_schema(schema, tk)
        tb = G.grammar_token_bytes(tk)
        for walk in range(3):
            m = oracle.matcher(schema)
            bm = xgr.allocate_token_bitmask(1, tk.vocab_size)
            state = 0
            for step in range(40):
                m.fill_next_token_bitmask(bm, 0)
                want = bm[0].numpy()
                assert np.array_equal(want, g.masks[state]), (name, walk, step)
                allowed = np.nonzero(unpack_mask(want, tk.vocab_size))[0]
                tok = int(rng.choice(allowed))
                assert m.accept_token(tok)
                if tok == tk.stop_ids[0]:
                    assert g.accept[state] and m.is_terminated()
                    break
                state = g.walk(state, tb[tok])
                assert state >= 0
