"""The inference engine behind VLLMAgent: what `vllm.LLM.generate` is to the reference
(vllm_agent.py:144,195,331,430,502), rebuilt for one B200.

Host side (this file, plain Python): tokenise, look prompts up in the block-hash prefix cache,
lay out KV pages, build the ragged prefill batch, launch.  Device side (libbcg_b200.so): the
whole model forward, paged attention, grammar-masked sampling and ALL per-token bookkeeping —
after prefill the decode loop is a CUDA graph replayed once per token with no host round trip;
the host only polls a single counter every few steps to stop early.

Differences from the engine the reference reaches (deliberate):
  * every row of a call is decoded together, each with its own grammar (the reference falls
    back to one-prompt-at-a-time when schemas differ, vllm_agent.py:417-455) and without the
    `max_num_seqs=4` cap (config.py:38) — per-row results are batch-invariant, so this changes
    scheduling, not outputs;
  * KV pages of a prompt's full 16-token blocks stay cached ACROSS calls and rounds, keyed by a
    hash chain over token ids (same idea as vLLM's enable_prefix_caching): per-agent system
    prompts and identical retry prompts (main.py:293-317) are not prefilled twice.
"""
from __future__ import annotations

import ctypes as C
import hashlib
import json
import math
import struct
import time
from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import grammar as G
from . import lib as L
from .model import ModelConfig, ModelHandle, Weights
from .tokenizer import CachedEncoder

PAGE = 16
SPLIT_TOKENS = 256
MAX_GRAMMARS = 64
POLL_EVERY = 8      # tensor-parallel engines: steps between two synchronised looks at the active-row count
IN_FLIGHT = 2       # single-GPU engines: decode steps queued ahead of the one whose result the host has seen
# query rows per prefill-attention CTA (bcg_prefill_attn_paged)
PREFILL_TILE = 128


def row_seed(base_seed: int, call_index: int, row: int):
    """Per-row Philox key (splitmix-style mix); restated in oracle/engine.py."""
    x = (base_seed * 0x9E3779B97F4A7C15 + call_index * 0xBF58476D1CE4E5B9 + row * 0x94D049BB133111EB
         + 0x2545F4914F6CDD1D) & 0xFFFFFFFFFFFFFFFF
    x ^= x >> 30
    x = (x * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
    x ^= x >> 27
    x = (x * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
    x ^= x >> 31
    return x & 0xFFFFFFFF, x >> 32


@dataclass
class Request:
    prompt: str
    schema: Optional[dict] = None
    temperature: float = 0.0
    max_tokens: int = 256
    forced: Optional[Sequence[int]] = None     # teacher-forced ids (parity tests)
    prompt_ids: Optional[List[int]] = None     # pre-tokenised prompt (tests / bench)
    seed: Optional[Sequence[int]] = None       # explicit Philox key (lo, hi); default: row_seed()


@dataclass
class Result:
    text: str
    token_ids: List[int]
    chosen: List[int]
    finish: str
    prompt_tokens: int
    cached_tokens: int


_FINISH = {0: "length", 1: "stop", 2: "length", 3: "dead"}


class PagePool:
    """Free list + block-hash prefix cache over the physical KV pages."""

    def __init__(self, n_pages: int):
        self.n_pages = n_pages
        self.free = list(range(n_pages - 1, -1, -1))
        self.cached: "OrderedDict[bytes, int]" = OrderedDict()  # hash -> page (LRU order)
        self.page_hash: Dict[int, bytes] = {}
        self.ref: Dict[int, int] = {}
        self.hits = 0
        self.misses = 0

    @staticmethod
    def chain(prev: bytes, block: Sequence[int]) -> bytes:
        return hashlib.blake2b(prev + struct.pack(f"<{len(block)}i", *block), digest_size=16).digest()

    def lookup(self, h: bytes) -> Optional[int]:
        p = self.cached.get(h)
        if p is not None:
            self.cached.move_to_end(h)
        return p

    def acquire(self, page: int) -> None:
        self.ref[page] = self.ref.get(page, 0) + 1

    def alloc(self) -> int:
        if not self.free:
            # evict least-recently-used cached pages nobody holds
            for h, p in list(self.cached.items()):
                if self.ref.get(p, 0) == 0:
                    del self.cached[h]
                    del self.page_hash[p]
                    self.free.append(p)
                    if len(self.free) >= 64:
                        break
        if not self.free:
            raise MemoryError("KV page pool exhausted")
        p = self.free.pop()
        self.ref[p] = 1
        return p

    def publish(self, page: int, h: bytes) -> None:
        """Make a fully written prompt page reusable by later requests."""
        if h not in self.cached and page not in self.page_hash:
            self.cached[h] = page
            self.page_hash[page] = h

    def release(self, page: int) -> None:
        r = self.ref.get(page, 0) - 1
        if r > 0:
            self.ref[page] = r
            return
        self.ref.pop(page, None)
        if page not in self.page_hash:
            self.free.append(page)

    def reset_cache(self) -> None:
        for h, p in list(self.cached.items()):
            if self.ref.get(p, 0) == 0:
                del self.cached[h]
                del self.page_hash[p]
                self.free.append(p)


class Engine:
    def __init__(self, cfg: ModelConfig, weights: Weights, tokenizer, *, max_model_len: int = 8192,
                 max_rows: int = 64, prefill_tokens: int = 16384, kv_pages: Optional[int] = None,
                 gpu_memory_utilization: float = 0.9, base_seed: int = 0, device: str = "cuda:0",
                 enable_prefix_cache: bool = True, use_cuda_graph: bool = True, pdl: bool = True,
                 fused_norm: bool = False, deferred_norm: Optional[bool] = None,
                 fused_rope: bool = False,
                 tp_rank: int = 0, tp_size: int = 1, tp_fused_allreduce: bool = True, tp_group=None):
        """`cfg` is the FULL architecture; with tp_size > 1 `weights` must be this rank's shard
        (Weights.shard / random_init_sharded) and bcg_tp_init must have been called (collective)."""
        if not torch.cuda.is_available():
            raise L.BcgError("the engine needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = L.load()
        self.cfg, self.w, self.tk = cfg, weights, tokenizer
        self.tp_rank, self.tp_size = tp_rank, tp_size
        full_cfg = cfg
        lcfg = cfg.shard(tp_size)          # this rank's heads / d_ff / vocab
        self.lcfg = lcfg
        if tp_size > 1 and weights.cfg.n_heads != lcfg.n_heads:
            raise L.BcgError("tensor-parallel engine needs sharded weights (Weights.shard)")
        self.dev = torch.device(device)
        torch.cuda.set_device(self.dev)
        self.max_model_len = min(max_model_len, cfg.max_pos)
        self.max_rows = min(max_rows, 128)
        self.T_max = max(prefill_tokens, self.max_model_len)
        self.base_seed = base_seed
        self.calls = 0
        self.enable_prefix_cache = enable_prefix_cache
        self.use_cuda_graph = use_cuda_graph
        # How decode runs the residual add + RMSNorm between projections (include/bcg_b200.h,
        # bcg_model_decode).  Default on one GPU: statistics kernel + 1/RMS in the next GEMM's epilogue
        # (1.5 % faster per step than the complete norm kernel).  TP engines with the fused all-reduce
        # do the same inside bcg_tp_add_norm_stats; with NCCL all-reduces they run the complete norm.
        self.tp_group = tp_group
        if deferred_norm is None:
            deferred_norm = tp_size == 1 and not fused_norm
        self.flags = ((L.GEMM_PDL if pdl else 0) | (L.STEP_FUSED_NORM if fused_norm else 0)
                      | (L.STEP_DEFERRED_NORM if deferred_norm else 0)
                      | (L.STEP_FUSED_ROPE if fused_rope else 0))
        self.max_blocks = (self.max_model_len + PAGE - 1) // PAGE
        self.max_steps = self.max_model_len

        d, B, T = cfg.d_model, self.max_rows, self.T_max
        bf, f32, i32 = torch.bfloat16, torch.float32, torch.int32
        z = lambda *s, dt=bf: torch.zeros(*s, dtype=dt, device=self.dev)
        # ---- workspace (include/bcg_b200.h bcg_workspace)
        self.max_splits_cap = (self.max_model_len + SPLIT_TOKENS - 1) // SPLIT_TOKENS
        lc = lcfg
        splits = [L.splits_for(1, n, k) for n, k in ((lc.qkv_dim, d), (d, lc.n_heads * 128),
                                                      (2 * lc.d_ff, d), (d, lc.d_ff))]
        pf = max(splits[0] * B * lc.qkv_dim, splits[1] * B * d,
                 (splits[2] * B * 2 * lc.d_ff) if splits[2] > 1 else 0, splits[3] * B * d)
        self.buf = dict(
            residual=z(T, d, dt=f32), xn=z(T, d), qkv=z(T, lc.qkv_dim, dt=f32), q=z(T, lc.n_heads * 128),
            attn=z(T, lc.n_heads * 128), act=z(T, lc.d_ff), delta=z(T, d, dt=f32), partials=z(pf, dt=f32),
            so=z(B * lc.n_heads * self.max_splits_cap * 128, dt=f32),
            sml=z(B * lc.n_heads * self.max_splits_cap * 2, dt=f32),
            attn_counters=z(B * lc.n_kv_heads, dt=i32),
            row_sumsq=z(2 * max((d + 127) // 128, 8) * 128, dt=f32), tile_counters=z((d + 127) // 128, dt=i32),
            # under tensor parallelism the logits stay sharded: [B, vocab / tp] per rank
            logits=z(B, lc.vocab, dt=f32), xn_last=z(B, d))
        ws = L.Workspace()
        b = self.buf
        ws.residual, ws.xn, ws.qkv, ws.q = b["residual"].data_ptr(), b["xn"].data_ptr(), b["qkv"].data_ptr(), b["q"].data_ptr()
        ws.attn, ws.act, ws.delta = b["attn"].data_ptr(), b["act"].data_ptr(), b["delta"].data_ptr()
        ws.partials, ws.partials_floats = b["partials"].data_ptr(), pf
        ws.attn_scratch_o, ws.attn_scratch_ml = b["so"].data_ptr(), b["sml"].data_ptr()
        ws.attn_counters = b["attn_counters"].data_ptr()
        ws.max_splits = self.max_splits_cap
        ws.logits, ws.xn_last, ws.T_max, ws.B_max = b["logits"].data_ptr(), b["xn_last"].data_ptr(), T, B
        ws.logits_local = None
        # tables for the deferred / fused norm (row sums of squares; tile counters for fused_norm=True)
        ws.row_sumsq = b["row_sumsq"].data_ptr()
        ws.tile_counters = b["tile_counters"].data_ptr() if (tp_size == 1 and d % 128 == 0) else None
        ws.tp_region_floats = 0
        if tp_size > 1:
            # The symmetric buffer (cudaIpc over NVLink) carries the sampler's candidate exchange
            # and, with tp_fused_allreduce, the decode step's all-reduces (bcg_tp_add_norm_stats).
            from . import tp as _tp
            region = B * d if tp_fused_allreduce else 0
            _tp.setup_symmetric(int(self.lib.bcg_tp_symm_bytes_for(region, tp_size)), tp_rank, tp_size, tp_group)
            ws.tp_region_floats = region
        self.ws = ws
        # ---- KV pages
        page_bytes = 2 * lc.n_kv_heads * PAGE * 128 * 2 * cfg.n_layers
        if kv_pages is None:
            free_b, total_b = torch.cuda.mem_get_info(self.dev)
            budget = int(total_b * gpu_memory_utilization) - (total_b - free_b)
            want = self.max_rows * self.max_blocks * 2
            kv_pages = max(self.max_blocks + 8, min(want, budget // page_bytes))
            if tp_size > 1:
                # every rank must run the SAME page pool (eviction, cache hits and prefill sizes are
                # decided per rank and have to agree): take the smallest budget of the replica
                import torch.distributed as dist
                t = torch.tensor([kv_pages], dtype=torch.int64)
                dist.all_reduce(t, op=dist.ReduceOp.MIN, group=tp_group)
                kv_pages = int(t.item())
        self.kv_layers = [z(kv_pages, 2, lc.n_kv_heads, PAGE, 128) for _ in range(cfg.n_layers)]
        self.pool = PagePool(kv_pages)
        self.model = ModelHandle(weights, self.kv_layers, PAGE, tp_size)
        # ---- per-row device state (bcg_sampler_args).  Everything the host writes at the start of a
        # call lives in ONE int32 block with a pinned host mirror, so a call costs one H2D copy
        # instead of one per field.
        fields = [("grammar_id", 1), ("state", 1), ("temperature", 1), ("seed", 2), ("max_tokens", 1),
                  ("n_out", 1), ("finished", 1), ("seq_lens", 1), ("block_table", self.max_blocks)]
        words = sum(w for _, w in fields) * B
        self._state_dev = z(words, dt=i32)
        self._state_host = torch.zeros(words, dtype=i32).pin_memory()
        self.row, self.row_host, off = {}, {}, 0
        for name, w in fields:
            dv, hv = self._state_dev[off:off + w * B], self._state_host[off:off + w * B].numpy()
            if w > 1:
                dv, hv = dv.view(B, w), hv.reshape(B, w)
            if name == "temperature":
                dv, hv = dv.view(f32), hv.view(np.float32)
            self.row[name], self.row_host[name] = dv, hv
            off += w * B
        self.row.update(
            out_tokens=z(B, self.max_steps, dt=i32), chosen=z(B, self.max_steps, dt=i32),
            forced=torch.full((B, self.max_steps), -1, dtype=i32, device=self.dev),
            next_ids=z(B, dt=i32), positions=z(B, dt=i32), slot_map=z(B, dt=i32))
        # Number of unfinished rows: the sampler's bookkeeping kernel writes it straight into pinned
        # HOST memory (mapped into the device's address space), so the host reads it without a copy.
        self.n_active_host = torch.zeros(1, dtype=i32).pin_memory()
        self.row["n_active"] = self.n_active_host
        # ---- prefill staging: ids | positions | slots | last rows | tiles | block table, one copy per group
        self._pf_cap = (3 * self.T_max, B, 4 * (self.T_max // PREFILL_TILE + B), B * self.max_blocks, 32)
        self._pf_dev = z(sum(self._pf_cap), dt=i32)
        self._pf_host = torch.zeros(sum(self._pf_cap), dtype=i32).pin_memory()
        # ---- tokenizer / grammar tables on the device
        tok_bytes = G.grammar_token_bytes(tokenizer)
        lens = np.fromiter((len(x) for x in tok_bytes), np.int64, count=cfg.vocab)
        offs = np.zeros(cfg.vocab + 1, np.int32)
        offs[1:] = np.cumsum(lens)
        blob = np.frombuffer(b"".join(tok_bytes) or b"\0", dtype=np.uint8).copy()
        self.tok_bytes = torch.from_numpy(blob).to(self.dev)
        self.tok_offsets = torch.from_numpy(offs).to(self.dev)
        self.mask_words = (cfg.vocab + 31) // 32
        self.grammars: Dict[str, int] = {}
        self.grammar_objs: List[G.CompiledGrammar] = []
        self._gram_dev = []
        self.mask_ptrs = torch.zeros(MAX_GRAMMARS, dtype=torch.int64, device=self.dev)
        self.trans_ptrs = torch.zeros(MAX_GRAMMARS, dtype=torch.int64, device=self.dev)
        self._graphs: Dict[tuple, torch.cuda.CUDAGraph] = {}
        self.stats = dict(calls=0, prompts=0, prompt_tokens=0, cached_tokens=0, prefill_tokens=0,
                          decode_steps=0, generated_tokens=0, prefill_s=0.0, decode_s=0.0, host_s=0.0,
                          gpu_launches=0, h2d_bytes=0, d2h_bytes=0)
        self._capture_stream = torch.cuda.Stream(self.dev)
        self._encode = CachedEncoder(tokenizer)    # retries resubmit identical prompts
        # parity tests: when set to a list, the logits of every step are appended (forces the
        # eager, graph-less decode loop)
        self.debug_logits: Optional[list] = None

    # ------------------------------------------------------------------ grammars
    def grammar_index(self, schema: Optional[dict]) -> int:
        if schema is None:
            return -1
        key = json.dumps(schema, sort_keys=False)
        gi = self.grammars.get(key)
        if gi is None:
            if len(self.grammar_objs) >= MAX_GRAMMARS:
                raise L.BcgError("too many distinct schemas in one engine")
            g = G.compile_schema(schema, self.tk)
            masks = torch.from_numpy(g.masks).to(self.dev).contiguous()
            trans = torch.from_numpy(g.trans).to(self.dev).contiguous()
            self._gram_dev.append((masks, trans))
            gi = len(self.grammar_objs)
            self.grammar_objs.append(g)
            self.grammars[key] = gi
            self.mask_ptrs[gi] = masks.data_ptr()
            self.trans_ptrs[gi] = trans.data_ptr()
        return gi

    # ------------------------------------------------------------------ sampler / steps
    def _sampler_args(self, B: int, use_forced: bool) -> L.SamplerArgs:
        r, a = self.row, L.SamplerArgs()
        a.logits, a.ld_logits = self.buf["logits"].data_ptr(), self.lcfg.vocab
        a.B, a.V, a.mask_words, a.max_steps = B, self.lcfg.vocab, self.mask_words, self.max_steps
        stops = list(self.tk.stop_ids) + [-1]
        a.eos_id0, a.eos_id1 = stops[0], stops[1]
        a.page_size, a.max_blocks_per_seq = PAGE, self.max_blocks
        a.grammar_id = r["grammar_id"].data_ptr()
        a.mask_tables, a.trans_tables = self.mask_ptrs.data_ptr(), self.trans_ptrs.data_ptr()
        a.tok_bytes, a.tok_offsets = self.tok_bytes.data_ptr(), self.tok_offsets.data_ptr()
        a.temperature, a.seed, a.max_tokens = r["temperature"].data_ptr(), r["seed"].data_ptr(), r["max_tokens"].data_ptr()
        a.forced = r["forced"].data_ptr() if use_forced else None
        a.state, a.n_out, a.finished = r["state"].data_ptr(), r["n_out"].data_ptr(), r["finished"].data_ptr()
        a.out_tokens, a.chosen = r["out_tokens"].data_ptr(), r["chosen"].data_ptr()
        a.next_ids, a.positions = r["next_ids"].data_ptr(), r["positions"].data_ptr()
        a.seq_lens, a.slot_map = r["seq_lens"].data_ptr(), r["slot_map"].data_ptr()
        a.block_table, a.n_active = r["block_table"].data_ptr(), r["n_active"].data_ptr()
        if self.tp_size > 1:   # vocab-parallel sampling: local shard + one candidate per row exchanged
            a.tp_rank, a.tp_size, a.vocab_base = self.tp_rank, self.tp_size, self.tp_rank * self.lcfg.vocab
        return a

    def _sample(self, B: int, use_forced: bool) -> None:
        a = self._sampler_args(B, use_forced)
        L.check(self.lib.bcg_mask_sample(C.byref(a), L.stream_ptr()), "bcg_mask_sample")

    def _decode_step(self, B: int, max_splits: int, use_forced: bool) -> None:
        r = self.row
        self.ws.max_splits = max_splits
        L.check(self.lib.bcg_model_decode(C.byref(self.model.struct), C.byref(self.ws),
                                          r["next_ids"].data_ptr(), r["positions"].data_ptr(),
                                          r["slot_map"].data_ptr(), r["seq_lens"].data_ptr(),
                                          r["block_table"].data_ptr(), self.max_blocks, B, self.flags,
                                          L.stream_ptr()), "bcg_model_decode")
        self._sample(B, use_forced)

    def launches_per_decode_step(self, B: int) -> int:
        cfg = self.cfg
        s_gu = L.splits_for(B, 2 * cfg.d_ff, cfg.d_model)
        per_layer = 8 + (1 if s_gu > 1 else 0)   # norm, qkv, rope, attn, o, norm, gate_up(+reduce), down
        return cfg.n_layers * per_layer + 2 + 2  # final norm + lm_head + sampler + active count

    def _graph_for(self, B: int, max_splits: int, use_forced: bool):
        key = (B, max_splits, use_forced)
        g = self._graphs.get(key)
        if g is None:
            # The graph replays real state transitions, so capture must not disturb a live batch:
            # warm-up happens in _warmup() on scratch state; here we only record.
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=self._capture_stream):
                self._decode_step(B, max_splits, use_forced)
            self._graphs[key] = g
        return g

    # ------------------------------------------------------------------ generation
    def generate(self, requests: List[Request]) -> List[Result]:
        out: List[Result] = []
        for i in range(0, len(requests), self.max_rows):
            out.extend(self._generate_wave(requests[i:i + self.max_rows], i))
        self.calls += 1
        self.stats["calls"] += 1
        return out

    def _generate_wave(self, reqs: List[Request], row0: int) -> List[Result]:
        t_host0 = time.perf_counter()
        cfg, B = self.cfg, len(reqs)
        r = self.row
        ids_list, cached, pages_of, hashes_of = [], [], [], []
        use_forced = any(q.forced is not None for q in reqs)
        for q in reqs:
            ids = list(q.prompt_ids) if q.prompt_ids is not None else self._encode(q.prompt)
            if len(ids) + q.max_tokens > self.max_model_len:
                raise L.BcgError(f"prompt ({len(ids)} tokens) + max_tokens ({q.max_tokens}) exceeds "
                                 f"max_model_len ({self.max_model_len})")
            if not ids:
                raise L.BcgError("empty prompt")
            ids_list.append(ids)
        # ---- page layout + prefix-cache lookup
        try:
            for ids, q in zip(ids_list, reqs):
                n_full = len(ids) // PAGE
                hs, h = [], b"\0" * 16
                for j in range(n_full):
                    h = PagePool.chain(h, ids[j * PAGE:(j + 1) * PAGE])
                    hs.append(h)
                pages, hit = [], 0
                if self.enable_prefix_cache:
                    for j in range(n_full):
                        if (j + 1) * PAGE >= len(ids):
                            break  # always compute at least the last prompt token
                        p = self.pool.lookup(hs[j])
                        if p is None:
                            break
                        self.pool.acquire(p)
                        pages.append(p)
                        hit += 1
                total_pages = (len(ids) + q.max_tokens + PAGE - 1) // PAGE
                while len(pages) < total_pages:
                    pages.append(self.pool.alloc())
                pages_of.append(pages)
                hashes_of.append(hs)
                cached.append(hit * PAGE)
        except Exception:
            for pages in pages_of:
                for p in pages:
                    self.pool.release(p)
            raise

        try:
            results = self._run(reqs, ids_list, cached, pages_of, hashes_of, use_forced, row0, t_host0)
        finally:
            for pages in pages_of:
                for p in pages:
                    self.pool.release(p)
        return results

    def _run(self, reqs, ids_list, cached, pages_of, hashes_of, use_forced, row0, t_host0):
        cfg, B, r, dev = self.cfg, len(reqs), self.row, self.dev
        i32 = torch.int32
        bt = np.zeros((B, self.max_blocks), np.int32)
        for b, pages in enumerate(pages_of):
            bt[b, :len(pages)] = pages
        gids = [self.grammar_index(q.schema) for q in reqs]
        seeds = np.array([tuple(q.seed) if q.seed is not None else row_seed(self.base_seed, self.calls, row0 + b)
                          for b, q in enumerate(reqs)], np.uint32)
        # the previous call's copy out of the pinned mirror has long completed (every call ends with
        # a device synchronisation), so the mirror may be rewritten in place
        h = self.row_host
        h["grammar_id"][:B] = gids
        h["state"][:B] = 0
        h["temperature"][:B] = [q.temperature for q in reqs]
        h["max_tokens"][:B] = [q.max_tokens for q in reqs]
        h["n_out"][:B] = 0
        h["finished"][:B] = 0
        h["seq_lens"][:B] = [len(x) for x in ids_list]
        h["seed"][:B] = seeds.view(np.int32).reshape(B, 2)
        h["block_table"][:B] = bt
        self._state_dev.copy_(self._state_host, non_blocking=True)     # ONE copy for all per-row state
        self.stats["h2d_bytes"] += self._state_host.numel() * 4
        if use_forced:
            f = np.full((B, self.max_steps), -1, np.int32)
            for b, q in enumerate(reqs):
                if q.forced is not None:
                    f[b, :len(q.forced)] = np.asarray(q.forced, np.int32)
            r["forced"][:B].copy_(torch.from_numpy(f))
            self.stats["h2d_bytes"] += f.nbytes

        # ---- prefill in groups of <= T_max new tokens
        ev0, ev1, ev2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        ev0.record()
        b0 = 0
        while b0 < B:
            b1, T = b0, 0
            while b1 < B and T + (len(ids_list[b1]) - cached[b1]) <= self.T_max:
                T += len(ids_list[b1]) - cached[b1]
                b1 += 1
            self._prefill_group(b0, b1, ids_list, cached, bt)
            b0 = b1
        # publish freshly written full prompt pages for later calls
        if self.enable_prefix_cache:
            for b in range(B):
                for j, h in enumerate(hashes_of[b]):
                    self.pool.publish(pages_of[b][j], h)
        if self.debug_logits is not None:
            self.debug_logits.append(self._logits_rows(B))
        self._sample(B, use_forced)          # first token of every row
        self.stats["gpu_launches"] += 2
        ev1.record()
        self.stats["host_s"] += time.perf_counter() - t_host0

        # ---- decode loop: one graph replay per token, no host sync except the poll
        max_new = max(q.max_tokens for q in reqs)
        max_len = max(len(x) + q.max_tokens for x, q in zip(ids_list, reqs))
        max_splits = max(1, (max_len + SPLIT_TOKENS - 1) // SPLIT_TOKENS)
        steps = 0
        if max_new > 1:
            per_step = self.launches_per_decode_step(B)
            key = (B, max_splits, use_forced)
            graph = self._graphs.get(key)
            done = False
            use_graph = self.use_cuda_graph and self.debug_logits is None
            if not use_graph:
                graph = None
            if use_graph and graph is None:
                # first use of this shape: one eager step (it is a real step and also performs the
                # kernels' one-time attribute set-up), then record — recording executes nothing.
                self._decode_step(B, max_splits, use_forced)
                steps += 1
                self.stats["gpu_launches"] += per_step
                graph = self._graph_for(B, max_splits, use_forced)
            def one_step():
                if graph is not None:
                    graph.replay()
                else:
                    self._decode_step(B, max_splits, use_forced)
                    if self.debug_logits is not None:
                        self.debug_logits.append(self._logits_rows(B))

            if self.tp_size == 1:
                # The bookkeeping kernel of every step leaves the number of unfinished rows in pinned
                # host memory.  The host keeps at most IN_FLIGHT steps queued: before launching step
                # k it waits for step k - IN_FLIGHT and looks at the counter, so a call whose rows
                # have all finished runs at most IN_FLIGHT dead steps (it used to poll every 16
                # steps: ~8 dead steps of 5.5 ms on average for every call that ends early) and the
                # GPU never waits for the host (one step is queued behind the running one).
                ring = [torch.cuda.Event() for _ in range(IN_FLIGHT + 1)]
                launched = 0
                while steps < max_new - 1:
                    one_step()
                    ring[launched % len(ring)].record()
                    launched += 1
                    steps += 1
                    if launched >= IN_FLIGHT:
                        ring[(launched - IN_FLIGHT) % len(ring)].synchronize()
                        if int(self.n_active_host[0]) == 0:
                            break
                self.stats["gpu_launches"] += launched * per_step
            else:
                # Tensor parallelism: every rank must stop after the SAME step (a rank that ran one
                # step more would wait for peers that never come), so the ranks decide at fixed step
                # counts on a fully synchronised counter, which is identical everywhere.
                while steps < max_new - 1 and not done:
                    n = min(POLL_EVERY, max_new - 1 - steps)
                    for _ in range(n):
                        one_step()
                    steps += n
                    self.stats["gpu_launches"] += n * per_step
                    torch.cuda.current_stream().synchronize()
                    done = int(self.n_active_host[0]) == 0
        ev2.record()
        torch.cuda.synchronize()
        t_host1 = time.perf_counter()
        n_out = r["n_out"][:B].cpu().numpy()
        fin = r["finished"][:B].cpu().numpy()
        width = int(n_out.max()) if B else 0
        toks = r["out_tokens"][:B, :width].cpu().numpy()
        chosen = r["chosen"][:B, :width].cpu().numpy()
        self.stats["d2h_bytes"] += n_out.nbytes + fin.nbytes + toks.nbytes + chosen.nbytes
        results = []
        stop = set(self.tk.stop_ids)
        for b in range(B):
            t = toks[b, :n_out[b]].tolist()
            text = self.tk.decode([x for x in t if x not in stop])
            results.append(Result(text, t, chosen[b, :n_out[b]].tolist(), _FINISH.get(int(fin[b]), "length"),
                                  len(ids_list[b]), cached[b]))
        st = self.stats
        st["prompts"] += B
        st["prompt_tokens"] += sum(len(x) for x in ids_list)
        st["cached_tokens"] += sum(cached)
        st["prefill_tokens"] += sum(len(x) - c for x, c in zip(ids_list, cached))
        st["decode_steps"] += steps
        st["generated_tokens"] += int(n_out.sum())
        st["prefill_s"] += ev0.elapsed_time(ev1) / 1e3
        st["decode_s"] += ev1.elapsed_time(ev2) / 1e3
        st["host_s"] += time.perf_counter() - t_host1
        return results

    def _logits_rows(self, B: int) -> torch.Tensor:
        """[B, V] logits of the current step (parity tests only; under TP the vocabulary shards are
        gathered over the control group — the product path never moves logits between ranks)."""
        if self.tp_size == 1:
            return self.buf["logits"][:B].clone()
        import torch.distributed as dist
        mine = self.buf["logits"][:B].cpu()
        parts = [torch.empty_like(mine) for _ in range(self.tp_size)]
        dist.all_gather(parts, mine, group=self.tp_group)
        return torch.cat(parts, dim=1).to(self.dev)

    def _prefill_group(self, b0, b1, ids_list, cached, bt) -> None:
        cfg, dev = self.cfg, self.dev
        ids, pos, slots, tiles, last = [], [], [], [], []
        row = 0
        for b in range(b0, b1):
            c, n = cached[b], len(ids_list[b]) - cached[b]
            ids.extend(ids_list[b][c:])
            p = np.arange(c, c + n)
            pos.append(p)
            slots.append(bt[b, p // PAGE] * PAGE + p % PAGE)
            for q0 in range(0, n, PREFILL_TILE):
                tiles.append((b - b0, row + q0, c + q0, min(PREFILL_TILE, n - q0)))
            row += n
            last.append(row - 1)
        T, n_seqs = row, b1 - b0
        # one pinned staging block, one H2D copy per group: ids | positions | slots | last | tiles | bt.
        # A group's copy must have left the pinned block before the next group refills it.
        if getattr(self, "_pf_event", None) is not None:
            self._pf_event.synchronize()
        hs, o0 = self._pf_host.numpy(), 0
        views = []
        for arr in (ids, np.concatenate(pos), np.concatenate(slots), last, np.array(tiles).reshape(-1),
                    bt[b0:b1].reshape(-1)):
            a = np.asarray(arr, dtype=np.int32)
            hs[o0:o0 + a.size] = a
            views.append(self._pf_dev[o0:o0 + a.size])
            o0 += (a.size + 3) // 4 * 4        # keep every array 16-byte aligned
        self._pf_dev[:o0].copy_(self._pf_host[:o0], non_blocking=True)
        self._pf_event = torch.cuda.Event()
        self._pf_event.record()
        d_ids, d_pos, d_slots, d_last, d_tiles, d_bt = views
        self.stats["h2d_bytes"] += o0 * 4
        # this group's last-token logits go straight to rows b0..b1 of the row-indexed buffer
        base = self.buf["logits"].data_ptr()
        self.ws.logits = base + b0 * self.lcfg.vocab * 4
        try:
            L.check(self.lib.bcg_model_prefill(C.byref(self.model.struct), C.byref(self.ws),
                                               d_ids.data_ptr(), d_pos.data_ptr(), d_slots.data_ptr(),
                                               d_bt.data_ptr(), self.max_blocks, d_tiles.data_ptr(),
                                               len(tiles), d_last.data_ptr(), n_seqs, T, L.stream_ptr()),
                    "bcg_model_prefill")
        finally:
            self.ws.logits = base
        self.stats["gpu_launches"] += cfg.n_layers * 8 + 2

    # ------------------------------------------------------------------ misc
    def reset_prefix_cache(self) -> None:
        self.pool.reset_cache()

    def close(self) -> None:
        self._graphs.clear()
        self.kv_layers = None
        self.buf = None
