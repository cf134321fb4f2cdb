#!/usr/bin/env python
"""Device-side timeline of one decode step (dev tool; there is no nsys on the box).

Builds/loads the -DBCG_TIMELINE variant of the library (kernels stamp %globaltimer at entry,
after griddepcontrol.wait, when their main loop ends and at exit), replays one decode step from
a CUDA graph and prints, averaged over the middle layers, when each kernel of a layer starts,
gets past its dependency wait and finishes — i.e. where the hand-offs between the layer's seven
kernels lose time.  Numbers include ~0.1 us of stamping overhead per mark.
    python tools/timeline.py [--model qwen3-14b] [--rows 10] [--ctx 1600]
"""
import argparse
import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from byzantine_consensus_llm_agents_b200 import build as B  # noqa: E402
from byzantine_consensus_llm_agents_b200 import lib as L  # noqa: E402

L.LIB_PATH = B.LIB_TIMELINE  # must happen before anything loads the library
from byzantine_consensus_llm_agents_b200.engine import Engine  # noqa: E402
from byzantine_consensus_llm_agents_b200.model import PRESETS, Weights  # noqa: E402
from byzantine_consensus_llm_agents_b200.tokenizer import get_tokenizer  # noqa: E402

PH = {0: "entry", 1: "past wait", 2: "loop done", 3: "exit", 5: "merge exit"}


def dump(lib, name):
    fn = getattr(lib, name)
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_int, C.c_int]
    cap = 1 << 18
    buf = np.zeros(cap, dtype=np.uint64)
    n = fn(buf.ctypes.data, cap, 1)
    assert n >= 0, name
    v = buf[:n]
    return (v >> np.uint64(20)).astype(np.int64), ((v >> np.uint64(16)) & np.uint64(15)).astype(int), \
        ((v >> np.uint64(12)) & np.uint64(15)).astype(int), (v & np.uint64(4095)).astype(int)


def instances_by_gap(t, gap_ns):
    """Split sorted stamps into kernel instances wherever consecutive stamps are > gap apart."""
    t = np.sort(t)
    cuts = np.where(np.diff(t) > gap_ns)[0] + 1
    return np.split(t, cuts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="qwen3-14b")
    ap.add_argument("--rows", type=int, default=10)
    ap.add_argument("--ctx", type=int, default=1600)
    ap.add_argument("--base-flags", type=lambda x: int(x, 0), default=0,
                    help="extra bcg_model_decode flags, e.g. 0x20000 = norm folded into the GEMMs")
    a = ap.parse_args()
    if not B.LIB_TIMELINE.exists():
        B.build_timeline()
    cfg = PRESETS[a.model]
    dev = torch.device("cuda:0")
    eng = Engine(cfg, Weights.random_init(cfg, dev), get_tokenizer(cfg.family, cfg.vocab), max_model_len=4096,
                 max_rows=max(16, a.rows), prefill_tokens=4096, pdl=True)
    lib = eng.lib
    nb, r = a.rows, eng.row
    pages = (a.ctx + 15) // 16
    bt = torch.arange(nb * pages, dtype=torch.int32, device=dev).view(nb, pages)
    r["block_table"][:nb, :pages] = bt
    r["seq_lens"][:nb] = a.ctx
    r["positions"][:nb] = a.ctx - 1
    r["slot_map"][:nb] = (bt[:, -1] * 16 + (a.ctx - 1) % 16)
    r["next_ids"][:nb] = 17
    eng.buf["attn_counters"].zero_()
    eng.ws.max_splits = (a.ctx + 255) // 256

    def step():
        L.check(lib.bcg_model_decode(C.byref(eng.model.struct), C.byref(eng.ws), r["next_ids"].data_ptr(),
                                     r["positions"].data_ptr(), r["slot_map"].data_ptr(), r["seq_lens"].data_ptr(),
                                     r["block_table"].data_ptr(), eng.max_blocks, nb, eng.flags | a.base_flags, L.stream_ptr()))

    step()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        step()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    for name in ("bcg_tl_dump_gemm", "bcg_tl_dump_norm", "bcg_tl_dump_attn"):
        dump(lib, name)  # reset
    g.replay()
    torch.cuda.synchronize()
    ev = {}
    for name in ("bcg_tl_dump_gemm", "bcg_tl_dump_norm", "bcg_tl_dump_attn"):
        t, kid, ph, blk = dump(lib, name)
        for k in np.unique(kid):
            for p in np.unique(ph[kid == k]):
                ev[(int(k), int(p))] = t[(kid == k) & (ph == p)]
    nl = cfg.n_layers
    # GEMM: every instance has the same CTA count -> chunk the sorted stamps; others: split on gaps
    n_gemm = 4 * nl + 1
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    slots = 2 * sms if nb <= 32 else sms          # two CTAs per SM for decode batches of <= 32 rows
    d = cfg.d_model

    def ctas(n, k, fixed_splits=None):
        sp = fixed_splits or L.splits_for(nb, n, k)
        return min(slots, ((n + 127) // 128) * sp)

    s_gu = L.splits_for(nb, 2 * cfg.d_ff, d)
    per_layer = [ctas(cfg.qkv_dim, d), ctas(d, cfg.n_heads * 128), ctas(2 * cfg.d_ff, d), ctas(d, cfg.d_ff)]
    counts = per_layer * nl + [ctas(cfg.vocab, d, 1)]
    bounds = np.concatenate([[0], np.cumsum(counts)])
    gem = {}
    for p in (0, 1, 2, 3):
        t = np.sort(ev[(1, p)])
        assert len(t) == bounds[-1], (p, len(t), int(bounds[-1]))
        gem[p] = [t[bounds[i]:bounds[i + 1]] for i in range(n_gemm)]
    have_norm = (2, 0) in ev
    norm = {p: instances_by_gap(ev[(2, p)], 8000) for p in (0, 1, 3, 4, 5, 6, 7) if (2, p) in ev} if have_norm else {0: []}
    have_rope = (3, 0) in ev
    rope = {p: instances_by_gap(ev[(3, p)], 30000) for p in (0, 1)} if have_rope else {}
    att = {p: instances_by_gap(ev[(4, p)], 30000) for p in (0, 1, 2)}
    att_exit = instances_by_gap(np.concatenate([ev[(4, 3)], ev.get((4, 5), np.zeros(0, np.int64))]), 30000)
    print(f"{a.model} rows={nb} ctx={a.ctx}: gemm CTAs/instance {per_layer}, norm instances "
          f"{len(norm[0])}, rope {len(rope[0]) if have_rope else 0}, attention {len(att[0])}")
    step_ns = gem[3][-1].max() - gem[0][0].min()
    print(f"step (first GEMM entry -> lm_head exit): {step_ns / 1e6:.3f} ms")

    rows = []
    layers = range(5, nl - 5)
    for l in layers:
        # add_rmsnorm instances: 2l = after o-proj of layer l, 2l+1 = after down of layer l
        t0 = gem[1][4 * l].min()  # qkv GEMM: first CTA past its wait
        e = {}
        if have_norm:
            e["normA past wait"] = norm[1][2 * l - 1].min()
            e["normA exit (last CTA)"] = norm[3][2 * l - 1].max()
        else:
            e["previous layer's down GEMM epilogue done first/last"] = (gem[3][4 * l - 1].min(), gem[3][4 * l - 1].max())
        for i, nm in enumerate(("qkv", "o", "gate_up", "down")):
            gi = 4 * l + i
            e[f"{nm} GEMM entry first/last CTA"] = (gem[0][gi].min(), gem[0][gi].max())
            e[f"{nm} GEMM past wait first/last"] = (gem[1][gi].min(), gem[1][gi].max())
            e[f"{nm} GEMM last load issued first/last"] = (gem[2][gi].min(), gem[2][gi].max())
            e[f"{nm} GEMM epilogue done first/last"] = (gem[3][gi].min(), gem[3][gi].max())
            if nm == "qkv":
                if have_rope:
                    e["rope entry"] = rope[0][l].min()
                    e["rope past wait"] = rope[1][l].min()
                e["attn entry first/last"] = (att[0][l].min(), att[0][l].max())
                e["attn past wait first/last"] = (att[1][l].min(), att[1][l].max())
                e["attn loop done first/last"] = (att[2][l].min(), att[2][l].max())
                e["attn exit first/last"] = (att_exit[l].min(), att_exit[l].max())
            if nm == "o" and have_norm:
                e["normB past wait first/last"] = (norm[1][2 * l].min(), norm[1][2 * l].max())
                if 4 in norm:  # the complete (cluster) norm stamps its phases; the statistics kernel does not
                    e["normB loads+block sum done first/last"] = (norm[4][2 * l].min(), norm[4][2 * l].max())
                    e["normB cluster.sync passed first/last"] = (norm[5][2 * l].min(), norm[5][2 * l].max())
                    e["normB DSMEM read done first/last"] = (norm[6][2 * l].min(), norm[6][2 * l].max())
                    e["normB stores issued first/last"] = (norm[7][2 * l].min(), norm[7][2 * l].max())
                e["normB exit (last CTA)"] = norm[3][2 * l].max()
        rows.append({k: (np.array(v) - t0) for k, v in e.items()})
    print(f"average over layers {layers.start}..{layers.stop - 1}, us relative to the qkv GEMM passing its wait:")
    for k in rows[0]:
        m = np.mean([rw[k] for rw in rows], axis=0) / 1e3
        if np.ndim(m) == 0:
            print(f"  {k:44s} {m:9.2f}")
        else:
            print(f"  {k:44s} {m[0]:9.2f} {m[1]:9.2f}")


if __name__ == "__main__":
    main()
