"""The grammar-constrained sampler kernel must reproduce oracle/sampler.py bit-for-bit:
same token ids, same DFA states, same stop bookkeeping (integer work -> exact equality)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from byzantine_consensus_llm_agents_b200 import lib as L  # noqa: E402
from oracle import sampler as O  # noqa: E402

DEV = "cuda"


class Harness:
    """Owns the device arrays of bcg_sampler_args for B rows."""

    def __init__(self, B, V, n_states=5, max_steps=8, page=16, max_blocks=4, seed=0, n_grammars=2):
        rng = np.random.default_rng(seed)
        self.B, self.V, self.max_steps, self.page, self.max_blocks = B, V, max_steps, page, max_blocks
        self.words = (V + 31) // 32
        # token bytes: token i has 0..4 random bytes (a few byte-less "special" tokens)
        lens = rng.integers(1, 5, size=V)
        lens[-4:] = 0
        self.offsets = np.zeros(V + 1, np.int32)
        self.offsets[1:] = np.cumsum(lens)
        self.bytes = rng.integers(0, 256, size=int(self.offsets[-1]), dtype=np.uint8)
        self.masks, self.trans = [], []
        for _ in range(n_grammars):
            m = rng.integers(0, 2 ** 32, size=(n_states, self.words), dtype=np.uint64).astype(np.uint32)
            m &= rng.integers(0, 2 ** 32, size=(n_states, self.words), dtype=np.uint64).astype(np.uint32)
            if V % 32:
                m[:, -1] &= np.uint32((1 << (V % 32)) - 1)
            m[n_states - 1] = 0          # a state where nothing is allowed
            self.masks.append(m.view(np.int32))
            t = rng.integers(-1, n_states - 1, size=(n_states, 256)).astype(np.int16)
            self.trans.append(t)
        self.d_masks = [torch.from_numpy(m).to(DEV) for m in self.masks]
        self.d_trans = [torch.from_numpy(t).to(DEV) for t in self.trans]
        self.mask_ptrs = torch.tensor([t.data_ptr() for t in self.d_masks], dtype=torch.int64, device=DEV)
        self.trans_ptrs = torch.tensor([t.data_ptr() for t in self.d_trans], dtype=torch.int64, device=DEV)
        self.d_bytes = torch.from_numpy(self.bytes).to(DEV)
        self.d_offsets = torch.from_numpy(self.offsets).to(DEV)
        i32 = lambda a: torch.tensor(a, dtype=torch.int32, device=DEV)
        self.grammar_id = i32(rng.integers(-1, n_grammars, size=B))
        self.state = i32(rng.integers(0, n_states - 1, size=B))
        self.temperature = torch.tensor(rng.choice([0.0, 0.3, 0.5, 1.0], size=B), dtype=torch.float32, device=DEV)
        self.seed = torch.from_numpy(rng.integers(0, 2 ** 32, size=(B, 2), dtype=np.uint64).astype(np.uint32).view(np.int32)).to(DEV)
        self.max_tokens = i32(rng.integers(2, max_steps + 1, size=B))
        self.n_out = i32(np.zeros(B))
        self.finished = i32(np.zeros(B))
        self.out_tokens = torch.full((B, max_steps), -7, dtype=torch.int32, device=DEV)
        self.chosen = torch.full((B, max_steps), -7, dtype=torch.int32, device=DEV)
        self.next_ids = i32(np.zeros(B))
        self.positions = i32(np.zeros(B))
        self.seq_lens = i32(rng.integers(1, page * max_blocks - max_steps, size=B))
        self.slot_map = i32(np.zeros(B))
        self.block_table = i32(rng.integers(0, 1000, size=(B, max_blocks)))
        self.n_active = i32([0])
        self.eos0, self.eos1 = V - 1, V - 3
        self.forced = None

    def args(self, logits):
        a = L.SamplerArgs()
        a.logits, a.ld_logits = logits.data_ptr(), logits.stride(0)
        a.B, a.V, a.mask_words, a.max_steps = self.B, self.V, self.words, self.max_steps
        a.eos_id0, a.eos_id1, a.page_size, a.max_blocks_per_seq = self.eos0, self.eos1, self.page, self.max_blocks
        a.grammar_id = self.grammar_id.data_ptr()
        a.mask_tables, a.trans_tables = self.mask_ptrs.data_ptr(), self.trans_ptrs.data_ptr()
        a.tok_bytes, a.tok_offsets = self.d_bytes.data_ptr(), self.d_offsets.data_ptr()
        a.temperature, a.seed, a.max_tokens = self.temperature.data_ptr(), self.seed.data_ptr(), self.max_tokens.data_ptr()
        a.forced = None if self.forced is None else self.forced.data_ptr()
        a.state, a.n_out, a.finished = self.state.data_ptr(), self.n_out.data_ptr(), self.finished.data_ptr()
        a.out_tokens, a.chosen = self.out_tokens.data_ptr(), self.chosen.data_ptr()
        a.next_ids, a.positions, a.seq_lens, a.slot_map = (self.next_ids.data_ptr(), self.positions.data_ptr(),
                                                            self.seq_lens.data_ptr(), self.slot_map.data_ptr())
        a.block_table, a.n_active = self.block_table.data_ptr(), self.n_active.data_ptr()
        return a

    def snapshot(self):
        names = ["state", "n_out", "finished", "out_tokens", "chosen", "next_ids", "positions", "seq_lens", "slot_map"]
        return {n: getattr(self, n).cpu().numpy().copy() for n in names}


def oracle_step(h: Harness, s: dict, logits: np.ndarray, forced=None):
    """Pure-python restatement of the kernel's bookkeeping on top of oracle.sampler.sample."""
    gid = h.grammar_id.cpu().numpy()
    temp = h.temperature.cpu().numpy()
    seed = h.seed.cpu().numpy().view(np.uint32)
    max_tokens = h.max_tokens.cpu().numpy()
    bt = h.block_table.cpu().numpy()
    for b in range(h.B):
        if s["finished"][b]:
            continue
        step = int(s["n_out"][b])
        mask = None if gid[b] < 0 else h.masks[gid[b]][s["state"][b]]
        tok = O.sample(logits[b], mask, float(temp[b]), step, int(seed[b, 0]), int(seed[b, 1]))
        fin = 0
        if tok < 0:
            tok, fin = h.eos0, 3
        chosen = tok
        if forced is not None and forced[b, step] >= 0:
            tok = int(forced[b, step])
        s["chosen"][b, step] = chosen
        s["out_tokens"][b, step] = tok
        if tok in (h.eos0, h.eos1):
            fin = fin or 1
        elif gid[b] >= 0:
            st = int(s["state"][b])
            bs = h.bytes[h.offsets[tok]:h.offsets[tok + 1]]
            if len(bs) == 0:
                st = -1
            for by in bs:
                if st < 0:
                    break
                st = int(h.trans[gid[b]][st, by])
            if st < 0:
                fin, st = 3, 0
            s["state"][b] = st
        n = step + 1
        s["n_out"][b] = n
        if fin == 0 and (n >= max_tokens[b] or n >= h.max_steps):
            fin = 2
        s["next_ids"][b] = tok
        if fin == 0:
            pos = int(s["seq_lens"][b])
            s["positions"][b] = pos
            s["seq_lens"][b] = pos + 1
            s["slot_map"][b] = bt[b, pos // h.page] * h.page + pos % h.page
        else:
            s["finished"][b] = fin
            s["slot_map"][b] = -1
    return s


@pytest.mark.parametrize("B,V", [(1, 128), (10, 1024), (7, 4100), (10, 151936), (3, 32768)])
def test_sampler_matches_oracle_over_steps(B, V):
    lib = L.load()
    h = Harness(B, V, seed=B * 31 + V)
    if B >= 7:  # one row sits in a state whose mask row is empty -> grammar dead end (finished = 3)
        h.grammar_id[1] = 0
        h.state[1] = 4
    ref = h.snapshot()
    rng = np.random.default_rng(99)
    for step in range(h.max_steps):
        logits_np = (rng.standard_normal((B, V)) * 0.64).astype(np.float32)
        logits = torch.from_numpy(logits_np).to(DEV)
        L.check(lib.bcg_mask_sample(C.byref(h.args(logits)), L.stream_ptr()))
        torch.cuda.synchronize()
        ref = oracle_step(h, ref, logits_np)
        got = h.snapshot()
        for k in ref:
            assert np.array_equal(ref[k], got[k]), f"step {step}: {k} differs\n{ref[k]}\n{got[k]}"
        assert int(h.n_active.item()) == int((ref["finished"] == 0).sum())
    assert (ref["finished"] != 0).all()  # max_tokens <= max_steps: everything stopped


def test_sampler_forced_tokens_and_chosen():
    lib = L.load()
    B, V = 6, 2048
    h = Harness(B, V, seed=5)
    rng = np.random.default_rng(3)
    forced = rng.integers(0, V - 8, size=(B, h.max_steps)).astype(np.int32)
    forced[::2] = -1
    h.forced = torch.from_numpy(forced).to(DEV)
    ref = h.snapshot()
    for step in range(4):
        logits_np = rng.standard_normal((B, V)).astype(np.float32)
        logits = torch.from_numpy(logits_np).to(DEV)
        L.check(lib.bcg_mask_sample(C.byref(h.args(logits)), L.stream_ptr()))
        torch.cuda.synchronize()
        ref = oracle_step(h, ref, logits_np, forced)
        got = h.snapshot()
        for k in ref:
            assert np.array_equal(ref[k], got[k]), f"step {step}: {k}"


def test_sampler_greedy_ties_pick_lowest_index():
    lib = L.load()
    B, V = 2, 1024
    h = Harness(B, V, seed=8)
    h.grammar_id.fill_(-1)
    h.temperature.fill_(0.0)
    logits = torch.zeros(B, V, device=DEV)
    logits[0, [900, 37, 512]] = 3.0
    logits[1, [1000, 999]] = 1.0
    L.check(lib.bcg_mask_sample(C.byref(h.args(logits)), L.stream_ptr()))
    torch.cuda.synchronize()
    assert h.out_tokens[:, 0].tolist() == [37, 999]


def test_sampler_rejects_bad_vocab():
    lib = L.load()
    h = Harness(2, 128)
    logits = torch.zeros(2, 128, device=DEV)
    a = h.args(logits)
    a.V = 127
    assert lib.bcg_mask_sample(C.byref(a), L.stream_ptr()) != 0
    assert b"bcg_mask_sample" in lib.bcg_last_error()
