"""ORACLE (test infrastructure only — never imported by the product path).

Grammar oracle = xgrammar 0.2.0 (third-party, installed in this image; vLLM's default
structured-output backend, vllm/config/structured_outputs.py:21) driven exactly the way vLLM
drives it for the reference's `GuidedDecodingParams(json=schema)`:

    TokenizerInfo(vocab)                                  backend_xgrammar.py:60-69
    GrammarCompiler.compile_json_schema(json, any_whitespace=True)   :81-83, 116-122
    GrammarMatcher.fill_next_token_bitmask / accept_token / is_terminated   :148-167, 191-192

The reference repository pins nothing here (requirements.txt:1 is `vllm>=0.6.0`) and holds no
golden masks, so the pin is "xgrammar 0.2.0 as installed" — outputs of the real library run in
this container, recorded under tests/golden/grammar_masks.npz by
oracle/tools/make_grammar_golden.py.
"""
from __future__ import annotations

import json

import numpy as np


def tokenizer_info(tokenizer):
    import xgrammar as xgr
    ev = tokenizer.encoded_vocab_for_xgrammar()
    last = max(max(tokenizer.special_ids.values()), tokenizer.regular_range[1] - 1)
    ev = ev[:last + 1]   # HF get_vocab() has no entries for the unused ids at the top
    return xgr.TokenizerInfo(ev, vocab_type=xgr.VocabType.BYTE_LEVEL, vocab_size=tokenizer.vocab_size,
                             stop_token_ids=[tokenizer.stop_ids[0]])


class XgrOracle:
    def __init__(self, tokenizer):
        import xgrammar as xgr
        self.xgr = xgr
        self.tk = tokenizer
        self.info = tokenizer_info(tokenizer)
        self.compiler = xgr.GrammarCompiler(self.info)
        self._compiled = {}
        # single-byte token ids (byte-level alphabet)
        self.byte_token = {}
        for i in range(*tokenizer.regular_range):
            b = tokenizer.token_bytes[i]
            if len(b) == 1 and b[0] not in self.byte_token:
                self.byte_token[b[0]] = i

    def compiled(self, schema: dict):
        key = json.dumps(schema)
        if key not in self._compiled:
            self._compiled[key] = self.compiler.compile_json_schema(key, any_whitespace=True)
        return self._compiled[key]

    def matcher(self, schema: dict):
        return self.xgr.GrammarMatcher(self.compiled(schema))

    def mask_after_bytes(self, schema: dict, prefix: bytes) -> np.ndarray:
        """int32 [ceil(V/32)] bitmask after feeding `prefix` one byte-token at a time."""
        m = self.matcher(schema)
        for b in prefix:
            if not m.accept_token(self.byte_token[b]):
                raise AssertionError(f"xgrammar rejected byte {b!r} of prefix {prefix!r}")
        bm = self.xgr.allocate_token_bitmask(1, self.tk.vocab_size)
        m.fill_next_token_bitmask(bm, 0)
        return bm[0].numpy().copy()

    def mask_after_tokens(self, schema: dict, tokens) -> np.ndarray:
        m = self.matcher(schema)
        for t in tokens:
            assert m.accept_token(int(t)), t
        bm = self.xgr.allocate_token_bitmask(1, self.tk.vocab_size)
        m.fill_next_token_bitmask(bm, 0)
        return bm[0].numpy().copy()


def shortest_prefixes(trans: np.ndarray) -> list:
    """BFS over the byte DFA: for every state, one shortest byte string that reaches it."""
    S = trans.shape[0]
    pre = [None] * S
    pre[0] = b""
    queue = [0]
    qi = 0
    while qi < len(queue):
        s = queue[qi]
        qi += 1
        for b in range(256):
            t = int(trans[s, b])
            if t >= 0 and pre[t] is None:
                pre[t] = pre[s] + bytes([b])
                queue.append(t)
    return pre
