"""Deterministic synthetic tokenizer.

The reference lets vLLM load the HuggingFace tokenizer of the model it names
(vllm_agent.py:126-144); no tokenizer files exist offline, so the engine — and the CPU oracle,
which imports this same class — use a synthetic byte-level BPE with the real models' id-space
geometry:

  * ids [0, 256)            the 256 byte tokens (GPT-2 byte<->unicode alphabet, sorted)
  * ids [256, 256+M)        BPE merges trained on the reference's own prompts
                            (assets/bpe_merges.json, oracle/tools/make_tokenizer_assets.py)
  * ids up to n_regular     deterministic filler tokens (never produced by encode(), but
                            decodable and visible to the grammar mask): words, digits,
                            punctuation/JSON fragments, whitespace runs, UTF-8 characters and
                            partial UTF-8 sequences — so masks have realistic density and tokens
                            straddle grammar boundaries
  * qwen3 family:  26 special tokens at n_regular.. with <|endoftext|>, <|im_start|>,
                   <|im_end|> first (151643/4/5 at the full 151,936 vocabulary); ids above the
                   last special token up to the model's vocab are unused padding, as in Qwen.
  * mistral family: control tokens <unk> <s> </s> [INST] [/INST] ... at ids 0..9.

Encoding uses the `tokenizers` library's BPE model (the same library vLLM uses for HF
tokenizers); decoding concatenates token bytes, drops special tokens, and decodes UTF-8 with
replacement, like `skip_special_tokens=True`.
"""
from __future__ import annotations

import json
import random
from functools import lru_cache
from pathlib import Path

import numpy as np

ASSETS = Path(__file__).resolve().parent / "assets"

QWEN_SPECIALS = [
    "<|endoftext|>", "<|im_start|>", "<|im_end|>", "<|object_ref_start|>", "<|object_ref_end|>",
    "<|box_start|>", "<|box_end|>", "<|quad_start|>", "<|quad_end|>", "<|vision_start|>",
    "<|vision_end|>", "<|vision_pad|>", "<|image_pad|>", "<|video_pad|>", "<tool_call>",
    "</tool_call>", "<|fim_prefix|>", "<|fim_middle|>", "<|fim_suffix|>", "<|fim_pad|>",
    "<|repo_name|>", "<|file_sep|>", "<tool_response>", "</tool_response>", "<think>", "</think>",
]
QWEN_RESERVED_TOP = 151936 - 151643  # specials + unused ids at the top of the id space

MISTRAL_SPECIALS = ["<unk>", "<s>", "</s>", "[INST]", "[/INST]", "[TOOL_CALLS]",
                    "[AVAILABLE_TOOLS]", "[/AVAILABLE_TOOLS]", "[TOOL_RESULTS]", "[/TOOL_RESULTS]"]


def _bytes_to_unicode():
    """GPT-2 byte <-> printable unicode map (the ByteLevel alphabet)."""
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAD)) + list(range(0xAE, 0x100))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return dict(zip(bs, map(chr, cs)))


_B2U = _bytes_to_unicode()
_U2B = {u: b for b, u in _B2U.items()}

_WORDS = ("the of and to in is that for it as was with be by on not he this are or his from at "
          "which but have an had they you were their one all we can her has there been if more "
          "when will would who so no value round agent consensus propose vote stop continue "
          "strategy reasoning network honest byzantine agree majority median average trust "
          "converge history current initial final decision public internal abstain").split()


def _filler_tokens(n: int, taken: set) -> list:
    """n distinct byte strings not in `taken`, from a fixed-seed stdlib PRNG."""
    rng = random.Random(0xBC6B200)
    letters = "abcdefghijklmnopqrstuvwxyz"
    punct = list("{}[]():;,.!?\"'\\/-_=+*&^%$#@~`|<>")
    json_frags = ['":', '",', '"}', '{"', '":"', '", "', '": "', '"}\n', '":{', '"]', '["', '},', '}}',
                  '" ', ' "', '\\"', '\\n', '\\\\', '": ', ',"', '" :', '"\n', '\n"', '",\n', '" }']
    out, seen = [], set(taken)

    def push(b: bytes):
        if b and b not in seen and len(b) <= 24:
            seen.add(b)
            out.append(b)

    for f in json_frags:
        push(f.encode())
    for w in _WORDS:
        for v in (w, " " + w, w.capitalize(), " " + w.capitalize(), w.upper()):
            push(v.encode())
    for k in range(2, 17):
        push(b" " * k)
        push(b"\n" * min(k, 8))
        push(b"\t" * min(k, 4))
    for d in range(1000):
        push(str(d).encode())
        push((" " + str(d)).encode())
    guard = 0
    while len(out) < n:
        guard += 1
        if guard > 50 * n + 100000:
            raise RuntimeError("filler generation stalled")
        r = rng.random()
        if r < 0.50:      # word-like
            s = "".join(rng.choice(letters) for _ in range(rng.randint(2, 9)))
            if rng.random() < 0.15:
                s = s.capitalize()
            if rng.random() < 0.45:
                s = " " + s
            push(s.encode())
        elif r < 0.60:    # two words / word + punctuation
            s = rng.choice(_WORDS) + rng.choice([" ", ",", ".", ":", "_", "-"]) + \
                "".join(rng.choice(letters) for _ in range(rng.randint(1, 5)))
            push(s.encode())
        elif r < 0.68:    # punctuation runs, JSON-ish glue
            s = "".join(rng.choice(punct) for _ in range(rng.randint(1, 3)))
            if rng.random() < 0.3:
                s += rng.choice(["\n", " ", "  "])
            push(s.encode())
        elif r < 0.74:    # digits
            s = "".join(rng.choice("0123456789") for _ in range(rng.randint(1, 6)))
            if rng.random() < 0.3:
                s = rng.choice([" ", ".", ",", "-"]) + s
            push(s.encode())
        elif r < 0.94:    # whole non-ASCII characters (1..3 of them)
            cps = []
            for _ in range(rng.randint(1, 3)):
                lo, hi = rng.choice([(0xA1, 0x24F), (0x370, 0x3FF), (0x400, 0x4FF), (0x4E00, 0x9FFF),
                                     (0x3040, 0x30FF), (0xAC00, 0xD7A3), (0x1F300, 0x1F64F)])
                cps.append(chr(rng.randint(lo, hi)))
            push("".join(cps).encode("utf-8"))
        else:             # partial UTF-8 sequences (byte-level BPE really has these)
            ch = chr(rng.randint(0x800, 0xFFFF) if rng.random() < 0.7 else rng.randint(0x10000, 0x1FFFF))
            if 0xD800 <= ord(ch) <= 0xDFFF:
                continue
            b = ch.encode("utf-8")
            cut = rng.randint(1, len(b) - 1)
            push(b[:cut] if rng.random() < 0.5 else b[cut:])
    return out[:n]


class SyntheticTokenizer:
    def __init__(self, family: str = "qwen3", vocab_size: int = 151936):
        from tokenizers import AddedToken, Tokenizer, decoders, models, pre_tokenizers

        self.family = family
        self.vocab_size = vocab_size
        merges = [tuple(m) for m in json.loads((ASSETS / "bpe_merges.json").read_text())]
        if family == "mistral":
            specials = MISTRAL_SPECIALS
            base = len(specials)
            n_regular = vocab_size - base
        else:
            specials = QWEN_SPECIALS
            base = 0
            n_regular = vocab_size - QWEN_RESERVED_TOP
        if n_regular < 256 + 64:
            raise ValueError("vocab_size too small for the synthetic tokenizer")

        alphabet = sorted(_B2U.values())
        strs = list(alphabet)
        have = set(strs)
        kept_merges = []
        for a, b in merges:
            if len(strs) >= min(n_regular, 256 + len(merges)):
                break
            if a in have and b in have and (a + b) not in have:
                kept_merges.append((a, b))
                strs.append(a + b)
                have.add(a + b)
        taken = {bytes(_U2B[c] for c in s) for s in strs}
        for b in _filler_tokens(n_regular - len(strs), taken):
            strs.append("".join(_B2U[x] for x in b))
        assert len(strs) == n_regular and len(set(strs)) == n_regular

        vocab = {}
        if family == "mistral":
            for i, s in enumerate(specials):
                vocab[s] = i
        for i, s in enumerate(strs):
            vocab[s] = base + i
        tok = Tokenizer(models.BPE(vocab=vocab, merges=kept_merges))
        tok.pre_tokenizer = pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=True)
        tok.decoder = decoders.ByteLevel()
        tok.add_special_tokens([AddedToken(s, special=True, normalized=False) for s in specials])
        self._tok = tok

        # id -> bytes table (special / unused ids are byte-less)
        self.token_bytes = [b""] * vocab_size
        for i, s in enumerate(strs):
            self.token_bytes[base + i] = bytes(_U2B[c] for c in s)
        if family == "mistral":
            self.special_ids = {s: i for i, s in enumerate(specials)}
            self.bos_id, self.eos_id = self.special_ids["<s>"], self.special_ids["</s>"]
            self.stop_ids = (self.eos_id,)
        else:
            self.special_ids = {s: n_regular + i for i, s in enumerate(specials)}
            for s, i in self.special_ids.items():
                assert tok.token_to_id(s) == i, (s, i, tok.token_to_id(s))
            self.bos_id = None
            self.eos_id = self.special_ids["<|im_end|>"]
            self.stop_ids = (self.eos_id, self.special_ids["<|endoftext|>"])
        self.n_regular = n_regular
        self.regular_range = (base, base + n_regular)
        self._is_special = np.zeros(vocab_size, bool)
        self._is_special[list(self.special_ids.values())] = True

    # ---- encode / decode ---------------------------------------------------------------
    def encode(self, text: str, add_special_tokens: bool = True) -> list:
        ids = self._tok.encode(text, add_special_tokens=False).ids
        if add_special_tokens and self.bos_id is not None:
            ids = [self.bos_id] + ids
        return ids

    def decode(self, ids, skip_special_tokens: bool = True) -> str:
        parts = []
        for i in ids:
            i = int(i)
            if 0 <= i < self.vocab_size:
                if self._is_special[i]:
                    if not skip_special_tokens:
                        parts.append(next(s for s, j in self.special_ids.items() if j == i).encode())
                else:
                    parts.append(self.token_bytes[i])
        return b"".join(parts).decode("utf-8", errors="replace")

    # ---- tables for the grammar compiler / sampler kernel -------------------------------
    def byte_table(self):
        """(uint8 blob, int32 offsets[V+1]) of all token byte strings."""
        lens = np.fromiter((len(b) for b in self.token_bytes), dtype=np.int64, count=self.vocab_size)
        offsets = np.zeros(self.vocab_size + 1, np.int32)
        offsets[1:] = np.cumsum(lens)
        blob = np.frombuffer(b"".join(self.token_bytes), dtype=np.uint8).copy()
        return blob, offsets

    def encoded_vocab_for_xgrammar(self):
        """Token strings in the byte-level alphabet, as xgrammar.TokenizerInfo expects for
        VocabType.BYTE_LEVEL (oracle only)."""
        out = []
        inv = {i: s for s, i in self.special_ids.items()}
        for i in range(self.vocab_size):
            if i in inv:
                out.append(inv[i])
            elif self.token_bytes[i]:
                out.append("".join(_B2U[b] for b in self.token_bytes[i]))
            else:
                out.append(f"<|unused_{i}|>")
        return out


class CachedEncoder:
    """Prompt string -> token ids with a small FIFO memo.  The reference's retry rule resubmits
    identical prompts (main.py:293-317, 386-440: up to three batched attempts per phase, then
    per-agent loops), so within a round every prompt after its first submission is a hit and costs
    a dict lookup instead of a 2-4 ms BPE pass.  The returned list is shared: callers must not
    mutate it."""

    def __init__(self, tokenizer, capacity: int = 512):
        self.tk, self.capacity = tokenizer, max(1, capacity)
        self._memo: dict = {}
        self.hits = self.misses = 0

    def __call__(self, prompt: str) -> list:
        ids = self._memo.get(prompt)
        if ids is not None:
            self.hits += 1
            return ids
        self.misses += 1
        ids = self.tk.encode(prompt, True)
        if len(self._memo) >= self.capacity:
            self._memo.pop(next(iter(self._memo)))       # oldest entry (dicts keep insertion order)
        self._memo[prompt] = ids
        return ids

    def clear(self) -> None:
        self._memo.clear()


@lru_cache(maxsize=8)
def get_tokenizer(family: str = "qwen3", vocab_size: int = 151936) -> SyntheticTokenizer:
    return SyntheticTokenizer(family, vocab_size)


class LocalTokenizer(SyntheticTokenizer):
    """A real HuggingFace `tokenizer.json` from a local model directory (what the reference lets vLLM
    load by name, vllm_agent.py:126-144) behind the same interface as the synthetic tokenizer:
    encode / decode plus the id -> bytes table the grammar compiler and the sampler kernel need.
    Handles byte-level BPE vocabularies (Qwen: GPT-2 byte alphabet) and SentencePiece-style ones
    with byte fallback (Mistral: "▁" for blanks, <0xNN> byte tokens)."""

    def __init__(self, model_dir: str, vocab_size: int, family: str = "qwen3"):  # noqa: super().__init__ not wanted
        from tokenizers import Tokenizer
        d = Path(model_dir)
        self.family, self.vocab_size = family, vocab_size
        self._tok = Tokenizer.from_file(str(d / "tokenizer.json"))
        spec = json.loads((d / "tokenizer.json").read_text())
        byte_level = "ByteLevel" in json.dumps(spec.get("pre_tokenizer") or {}) + json.dumps(spec.get("decoder") or {})
        added = {a["id"]: a for a in spec.get("added_tokens", [])}
        self.special_ids = {a["content"]: i for i, a in added.items() if a.get("special", True)}
        self.token_bytes = [b""] * vocab_size
        for tok_str, i in self._tok.get_vocab(with_added_tokens=True).items():
            if i >= vocab_size or i in added:
                continue
            if byte_level:
                try:
                    self.token_bytes[i] = bytes(_U2B[c] for c in tok_str)
                except KeyError:
                    self.token_bytes[i] = tok_str.encode("utf-8")
            elif len(tok_str) == 6 and tok_str.startswith("<0x") and tok_str.endswith(">"):
                self.token_bytes[i] = bytes([int(tok_str[3:5], 16)])
            else:
                self.token_bytes[i] = tok_str.replace("▁", " ").encode("utf-8")
        cfg = {}
        for name in ("generation_config.json", "config.json", "tokenizer_config.json"):
            if (d / name).exists():
                for k, v in json.loads((d / name).read_text()).items():
                    cfg.setdefault(k, v)

        def token_id(v):
            if isinstance(v, dict):
                v = v.get("content")
            return self._tok.token_to_id(v) if isinstance(v, str) else v

        eos = cfg.get("eos_token_id", cfg.get("eos_token"))
        eos = [token_id(e) for e in (eos if isinstance(eos, list) else [eos]) if e is not None]
        if family != "mistral":   # ChatML turns end with <|im_end|> (vLLM takes it from generation_config)
            eos += [i for s, i in self.special_ids.items() if s in ("<|im_end|>", "<|endoftext|>")]
        self.stop_ids = tuple(dict.fromkeys(int(e) for e in eos if e is not None))[:2]
        if not self.stop_ids:
            raise ValueError(f"no end-of-sequence token found in {model_dir}")
        self.eos_id = self.stop_ids[0]
        add_bos = bool(cfg.get("add_bos_token", family == "mistral"))
        self.bos_id = token_id(cfg.get("bos_token_id", cfg.get("bos_token"))) if add_bos else None
        regular = [i for i in range(vocab_size) if self.token_bytes[i]]
        self.n_regular = len(regular)
        self.regular_range = (min(regular), max(regular) + 1)
        self._is_special = np.zeros(vocab_size, bool)
        self._is_special[[i for i in self.special_ids.values() if i < vocab_size]] = True
