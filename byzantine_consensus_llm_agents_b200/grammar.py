"""JSON-schema -> byte-level DFA -> token-level mask tables.

The reference asks vLLM for `GuidedDecodingParams(json=schema)` (vllm_agent.py:318, 423);
vLLM's default backend compiles the schema with xgrammar (`any_whitespace=True`) and, every
decode step, runs a pushdown matcher on the HOST to fill an int32 bitmask
(backend_xgrammar.py:191-192).  All schemas the reference produces (bcg_agents.py:590-599,
651-658, 1083-1092, 1155-1162) denote REGULAR languages, so the engine compiles them once to a
minimal byte-level DFA and precomputes, per DFA state, the exact token bitmask
(int32 [n_states, ceil(V/32)], bit = 1 allowed, LSB first — xgrammar's layout).  The decode loop
then needs no host work: the sampler kernel indexes the mask row by state and advances the
state over the sampled token's bytes on the device.

Grammar text is xgrammar 0.2.0's `Grammar.from_json_schema(schema, any_whitespace=True)`:
    ws      ::= [ \\n\\t]*
    string  ::= '"' ( [^\\0-\\x1f"\\\\] | '\\\\' ( ["\\\\/bfnrt] | 'u' HEX{4} ) )* '"'
    object  ::= '{' ws '"k1"' ws ':' ws v1 ( ws ',' ws '"k2"' ws ':' ws v2 )... ws '}'
                (properties in declaration order, optional ones may be skipped)
    integer ::= canonical decimal within [minimum, maximum]
tests/test_grammar_vs_xgrammar.py checks every reachable state's mask row bit-for-bit against
xgrammar's matcher (the oracle).
"""
from __future__ import annotations

import json
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np

WS_BYTES = b" \n\t"


# ------------------------------------------------------------------------- regex AST
def lit(b: bytes):
    return ("lit", bytes(b))


def cls(byte_values):
    m = np.zeros(256, bool)
    m[list(byte_values)] = True
    return ("cls", m)


def seq(*xs):
    return ("seq", list(xs))


def alt(*xs):
    return ("alt", list(xs))


def star(x):
    return ("star", x)


def opt(x):
    return ("alt", [("eps",), x])


EPS = ("eps",)


def _range(lo, hi):
    return cls(range(lo, hi + 1))


def ws():
    return star(cls(WS_BYTES))


def utf8_char_class():
    """[^\\0-\\x1f"\\\\] over code points as UTF-8 byte sequences, the way xgrammar 0.2.0 lowers a
    negated character class to bytes (probed against the library, see
    tests/test_grammar_vs_xgrammar.py): STRUCTURAL UTF-8 — lead byte 110xxxxx + 1, 1110xxxx + 2,
    11110xxx + 3 continuation bytes [80-BF] — without the overlong / surrogate / > U+10FFFF
    exclusions of strict UTF-8 (so C0, C1, F5-F7 are accepted as lead bytes)."""
    ascii_ok = [b for b in range(0x20, 0x80) if b not in (0x22, 0x5C)]
    cont = _range(0x80, 0xBF)
    return alt(
        cls(ascii_ok),
        seq(_range(0xC0, 0xDF), cont),
        seq(_range(0xE0, 0xEF), cont, cont),
        seq(_range(0xF0, 0xF7), cont, cont, cont),
    )


def json_string():
    hexd = cls(b"0123456789abcdefABCDEF")
    esc = seq(lit(b"\\"), alt(cls(b'"\\/bfnrt'), seq(lit(b"u"), hexd, hexd, hexd, hexd)))
    return seq(lit(b'"'), star(alt(utf8_char_class(), esc)), lit(b'"'))


def _digits_any(n):
    return seq(*[_range(0x30, 0x39) for _ in range(n)]) if n else EPS


def _ge(d: str):
    if not d:
        return EPS
    c = int(d[0])
    alts = [seq(lit(d[0].encode()), _ge(d[1:]))]
    if c < 9:
        alts.append(seq(_range(0x30 + c + 1, 0x39), _digits_any(len(d) - 1)))
    return alt(*alts)


def _le(d: str):
    if not d:
        return EPS
    c = int(d[0])
    alts = [seq(lit(d[0].encode()), _le(d[1:]))]
    if c > 0:
        alts.append(seq(_range(0x30, 0x30 + c - 1), _digits_any(len(d) - 1)))
    return alt(*alts)


def _fixed_len_between(lo: str, hi: str):
    assert len(lo) == len(hi)
    if not lo:
        return EPS
    l0, h0 = int(lo[0]), int(hi[0])
    if l0 == h0:
        return seq(lit(lo[0].encode()), _fixed_len_between(lo[1:], hi[1:]))
    alts = [seq(lit(lo[0].encode()), _ge(lo[1:]))]
    if h0 - l0 >= 2:
        alts.append(seq(_range(0x30 + l0 + 1, 0x30 + h0 - 1), _digits_any(len(lo) - 1)))
    alts.append(seq(lit(hi[0].encode()), _le(hi[1:])))
    return alt(*alts)


def _nonneg_between(a: int, b: int):
    """canonical decimals of a <= n <= b, 0 <= a <= b (no leading zeros)."""
    alts = []
    for length in range(len(str(a)), len(str(b)) + 1):
        lo = max(a, 0 if length == 1 else 10 ** (length - 1))
        hi = min(b, 10 ** length - 1)
        if lo <= hi:
            alts.append(_fixed_len_between(str(lo).zfill(length), str(hi)))
    return alt(*alts)


def json_integer(minimum: Optional[int] = None, maximum: Optional[int] = None):
    if minimum is None and maximum is None:
        return alt(lit(b"0"), seq(opt(lit(b"-")), _range(0x31, 0x39), star(_range(0x30, 0x39))))
    if minimum is None or maximum is None:
        raise UnsupportedSchema("integer with only one of minimum/maximum")
    if minimum > maximum:
        raise UnsupportedSchema("integer minimum > maximum")
    alts = []
    if maximum >= 0:
        alts.append(_nonneg_between(max(minimum, 0), maximum))
    if minimum < 0:
        alts.append(seq(lit(b"-"), _nonneg_between(max(1, -maximum), -minimum)))
    return alt(*alts)


def json_number():
    digits = seq(_range(0x30, 0x39), star(_range(0x30, 0x39)))
    intpart = alt(lit(b"0"), seq(_range(0x31, 0x39), star(_range(0x30, 0x39))))
    return seq(opt(lit(b"-")), intpart, opt(seq(lit(b"."), digits)),
               opt(seq(cls(b"eE"), opt(cls(b"+-")), digits)))


class UnsupportedSchema(ValueError):
    pass


def schema_to_ast(schema: dict):
    """The subset of JSON Schema the reference emits (plus the obvious neighbours)."""
    if "anyOf" in schema:
        return alt(*[schema_to_ast(s) for s in schema["anyOf"]])
    if "enum" in schema:
        return alt(*[lit(json.dumps(v).encode()) for v in schema["enum"]])
    if "const" in schema:
        return lit(json.dumps(schema["const"]).encode())
    t = schema.get("type")
    if t == "string":
        if any(k in schema for k in ("pattern", "format", "minLength", "maxLength")):
            raise UnsupportedSchema("string constraints other than enum/const")
        return json_string()
    if t == "integer":
        return json_integer(schema.get("minimum"), schema.get("maximum"))
    if t == "number":
        if "minimum" in schema or "maximum" in schema:
            raise UnsupportedSchema("bounded number")
        return json_number()
    if t == "boolean":
        return alt(lit(b"true"), lit(b"false"))
    if t == "null":
        return lit(b"null")
    if t == "array":
        if "items" not in schema or any(k in schema for k in ("minItems", "maxItems", "prefixItems")):
            raise UnsupportedSchema("array without plain `items`")
        item = schema_to_ast(schema["items"])
        return alt(seq(lit(b"["), ws(), item, star(seq(ws(), lit(b","), ws(), item)), ws(), lit(b"]")),
                   seq(lit(b"["), ws(), lit(b"]")))
    if t == "object":
        props = schema.get("properties", {})
        if schema.get("additionalProperties", True) is not False:
            raise UnsupportedSchema("object with additionalProperties (not a regular language here)")
        required = set(schema.get("required", []))
        names = list(props)
        entries = [seq(lit(json.dumps(k).encode()), ws(), lit(b":"), ws(), schema_to_ast(props[k]))
                   for k in names]
        is_req = [k in required for k in names]
        n = len(names)

        def after(i):  # at least one property already emitted; properties i.. remain
            alts = []
            for j in range(i, n):
                alts.append(seq(ws(), lit(b","), ws(), entries[j], after(j + 1)))
                if is_req[j]:
                    break
            else:
                alts.append(EPS)
            return alt(*alts) if len(alts) > 1 else alts[0]

        firsts = []
        all_optional = True
        for j in range(n):
            firsts.append(seq(entries[j], after(j + 1)))
            if is_req[j]:
                all_optional = False
                break
        body = []
        if firsts:
            body.append(seq(lit(b"{"), ws(), alt(*firsts) if len(firsts) > 1 else firsts[0], ws(),
                            lit(b"}")))
        if all_optional:
            body.append(seq(lit(b"{"), ws(), lit(b"}")))
        return alt(*body) if len(body) > 1 else body[0]
    raise UnsupportedSchema(f"schema type {t!r}")


# ------------------------------------------------------------------------- NFA -> DFA
class _NFA:
    def __init__(self):
        self.eps: List[List[int]] = []
        self.trans: List[List[Tuple[np.ndarray, int]]] = []

    def new(self) -> int:
        self.eps.append([])
        self.trans.append([])
        return len(self.eps) - 1

    def build(self, node, s: int) -> int:
        """Wire `node` starting at state s; returns its end state."""
        kind = node[0]
        if kind == "eps":
            return s
        if kind == "lit":
            for b in node[1]:
                t = self.new()
                m = np.zeros(256, bool)
                m[b] = True
                self.trans[s].append((m, t))
                s = t
            return s
        if kind == "cls":
            t = self.new()
            self.trans[s].append((node[1], t))
            return t
        if kind == "seq":
            for x in node[1]:
                s = self.build(x, s)
            return s
        if kind == "alt":
            end = self.new()
            for x in node[1]:
                a = self.new()
                self.eps[s].append(a)
                e = self.build(x, a)
                self.eps[e].append(end)
            return end
        if kind == "star":
            a = self.new()
            end = self.new()
            self.eps[s].append(a)
            e = self.build(node[1], a)
            self.eps[e].append(a)
            self.eps[a].append(end)
            return end
        raise ValueError(kind)


def _closure(nfa: _NFA, states) -> frozenset:
    stack, seen = list(states), set(states)
    while stack:
        s = stack.pop()
        for t in nfa.eps[s]:
            if t not in seen:
                seen.add(t)
                stack.append(t)
    return frozenset(seen)


def ast_to_dfa(ast) -> Tuple[np.ndarray, np.ndarray]:
    """-> (trans int16 [S,256] with -1 = reject, accept bool [S]); state 0 is the start;
    trimmed (every state reaches an accepting state) and minimised."""
    nfa = _NFA()
    start = nfa.new()
    end = nfa.build(ast, start)
    d0 = _closure(nfa, [start])
    ids: Dict[frozenset, int] = {d0: 0}
    order = [d0]
    rows = []
    i = 0
    while i < len(order):
        cur = order[i]
        i += 1
        # byte -> set of targets
        tgt: List[set] = [set() for _ in range(256)]
        for s in cur:
            for mask, t in nfa.trans[s]:
                for b in np.nonzero(mask)[0]:
                    tgt[b].add(t)
        row = np.full(256, -1, np.int32)
        cache: Dict[frozenset, int] = {}
        for b in range(256):
            if not tgt[b]:
                continue
            key = frozenset(tgt[b])
            if key not in cache:
                c = _closure(nfa, key)
                if c not in ids:
                    ids[c] = len(order)
                    order.append(c)
                cache[key] = ids[c]
            row[b] = cache[key]
        rows.append(row)
    trans = np.stack(rows)
    accept = np.array([end in st for st in order], bool)

    # trim states that cannot reach an accepting state
    S = len(order)
    live = accept.copy()
    changed = True
    while changed:
        changed = False
        for s in range(S):
            if not live[s]:
                t = trans[s]
                if live[t[t >= 0]].any():
                    live[s] = True
                    changed = True
    trans = np.where((trans >= 0) & live[np.maximum(trans, 0)], trans, -1)

    # Moore partition refinement
    part = np.where(accept, 1, 0)
    part[~live] = -1
    while True:
        sig = {}
        new = np.full(S, -1, np.int64)
        for s in range(S):
            if not live[s]:
                continue
            key = (int(part[s]), tuple(np.where(trans[s] >= 0, part[np.maximum(trans[s], 0)], -1).tolist()))
            if key not in sig:
                sig[key] = len(sig)
            new[s] = sig[key]
        if len(sig) == len(set(part[live].tolist())):
            part = new
            break
        part = new
    # renumber: start state's class first, then BFS order for determinism
    n_cls = int(part.max()) + 1
    rep = [-1] * n_cls
    for s in range(S):
        if live[s] and rep[part[s]] < 0:
            rep[part[s]] = s
    remap = {int(part[0]): 0}
    queue = [int(part[0])]
    qi = 0
    while qi < len(queue):
        c = queue[qi]
        qi += 1
        for b in range(256):
            t = trans[rep[c], b]
            if t >= 0 and int(part[t]) not in remap:
                remap[int(part[t])] = len(remap)
                queue.append(int(part[t]))
    n = len(remap)
    if n > 32767:
        # the device tables (and the sampler's DFA walk) hold states as int16
        raise UnsupportedSchema(f"schema compiles to {n} DFA states; the device tables hold at most 32767")
    out = np.full((n, 256), -1, np.int16)
    acc = np.zeros(n, bool)
    for c, i2 in remap.items():
        row = trans[rep[c]]
        out[i2] = [remap[int(part[t])] if t >= 0 else -1 for t in row]
        acc[i2] = accept[rep[c]]
    return out, acc


# ------------------------------------------------------------------------- token tables
@dataclass
class CompiledGrammar:
    schema_key: str
    trans: np.ndarray       # int16 [S, 256]
    accept: np.ndarray      # bool [S]
    masks: np.ndarray       # int32 [S, ceil(V/32)]
    n_states: int

    def walk(self, state: int, data: bytes) -> int:
        for b in data:
            if state < 0:
                return -1
            state = int(self.trans[state, b])
        return state


def grammar_token_bytes(tokenizer) -> List[bytes]:
    """Bytes of every token AS THE GRAMMAR SEES THEM: vLLM hands xgrammar the HF vocabulary,
    in which special tokens are ordinary strings ("<|im_start|>" can legally appear inside a
    JSON string); the stop token is never text; ids past the tokenizer's last token do not
    exist for the grammar."""
    out = list(tokenizer.token_bytes)
    for s, i in tokenizer.special_ids.items():
        out[i] = s.encode()
    for i in tokenizer.stop_ids[:1]:
        out[i] = b""
    return out


def token_masks(trans: np.ndarray, accept: np.ndarray, tok_bytes: List[bytes], vocab_size: int,
                stop_ids) -> np.ndarray:
    """bool-exact token bitmask per DFA state."""
    S = trans.shape[0]
    V = vocab_size
    lens = np.fromiter((len(b) for b in tok_bytes), np.int64, count=V)
    maxlen = int(lens.max()) if V else 0
    mat = np.zeros((V, maxlen), np.uint8)
    for i, b in enumerate(tok_bytes):
        if b:
            mat[i, :len(b)] = np.frombuffer(b, np.uint8)
    # dead state = S (absorbing)
    tr = np.full((S + 1, 256), S, np.int32)
    tr[:S] = np.where(trans >= 0, trans, S)
    cur = np.broadcast_to(np.arange(S, dtype=np.int32)[:, None], (S, V)).copy()
    for j in range(maxlen):
        active = lens > j
        idx = np.nonzero(active)[0]
        if idx.size == 0:
            break
        cur[:, idx] = tr[cur[:, idx], mat[idx, j][None, :]]
    allowed = (cur != S) & (lens > 0)[None, :]
    for sid in stop_ids[:1]:
        allowed[:, sid] = accept
    words = (V + 31) // 32
    padded = np.zeros((S, words * 32), bool)
    padded[:, :V] = allowed
    bits = padded.reshape(S, words, 32).astype(np.uint32)
    packed = (bits << np.arange(32, dtype=np.uint32)[None, None, :]).sum(axis=2, dtype=np.uint64)
    return packed.astype(np.uint32).view(np.int32)


_CACHE: Dict[Tuple[str, int, str], CompiledGrammar] = {}


def compile_schema(schema: dict, tokenizer) -> CompiledGrammar:
    key = json.dumps(schema, sort_keys=False)
    ck = (key, tokenizer.vocab_size, tokenizer.family)
    g = _CACHE.get(ck)
    if g is None:
        trans, accept = ast_to_dfa(schema_to_ast(schema))
        masks = token_masks(trans, accept, grammar_token_bytes(tokenizer), tokenizer.vocab_size,
                            tokenizer.stop_ids)
        g = CompiledGrammar(key, trans, accept, masks, trans.shape[0])
        _CACHE[ck] = g
    return g
