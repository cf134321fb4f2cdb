"""GPU parity tests of the individual sm_100a kernels, called through the C ABI (ctypes).

Floating-point kernels are compared with a plain PyTorch fp32 evaluation of the same op on the
same bf16-valued inputs; tolerances are written next to each comparison.  The sampler is integer
work and must match oracle/sampler.py exactly (see test_gpu_sampler.py).
"""
import math
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from byzantine_consensus_llm_agents_b200 import lib as L  # noqa: E402

DEV = "cuda"


def _lib():
    return L.load()


def _rand_bf16(*shape, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(torch.bfloat16).to(DEV)


# ------------------------------------------------------------------ GEMM (tcgen05)

GEMM_SHAPES = [
    # M, N, K
    (1, 128, 64),
    (10, 512, 1024),
    (16, 256, 128),
    (17, 384, 320),
    (33, 1024, 512),
    (64, 640, 192),
    (100, 256, 2048),
    (128, 512, 512),
    (129, 256, 128),
    (300, 512, 256),
    (1000, 1024, 512),
    (4096, 2048, 1024),
    # prefill shapes on CTA pairs (M >= 512, N % 256 == 0): one k-block, ragged K, ragged M inside the
    # pair's second CTA, more tiles than clusters (several tiles per cluster, both accumulator stages)
    (513, 256, 64),
    (700, 768, 320),
    (641, 512, 4096),
    (4096, 7168, 512),
]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_bf16_out(M, N, K):
    x = _rand_bf16(M, K, seed=1)
    w = _rand_bf16(N, K, scale=0.05, seed=2)
    out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    L.gemm(x, w, out, L.EPI_BF16)
    torch.cuda.synchronize()
    ref = x.float() @ w.float().T
    # fp32 accumulation of exact bf16 products, one rounding to bf16 at the end:
    # |err| <= 2^-8 relative (one bf16 ulp) + accumulation-order noise
    err = (out.float() - ref).abs()
    tol = ref.abs() * 2 ** -7 + 1e-3
    assert torch.isfinite(out.float()).all()
    assert (err <= tol).all(), f"max err {err.max().item()} at {M}x{N}x{K}"


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_f32_out(M, N, K):
    x = _rand_bf16(M, K, seed=3)
    w = _rand_bf16(N, K, scale=0.05, seed=4)
    out = torch.full((M, N), float("nan"), dtype=torch.float32, device=DEV)
    L.gemm(x, w, out, L.EPI_F32)
    torch.cuda.synchronize()
    ref = x.double() @ w.double().T
    err = (out.double() - ref).abs().max().item()
    # fp32 accumulate over K terms of magnitude ~0.05: tolerance 1e-4 * sqrt(K)
    assert err <= 1e-4 * math.sqrt(K) + 1e-5, f"max err {err}"


@pytest.mark.parametrize("M,N,K,splits", [(10, 512, 1024, 4), (7, 256, 2048, 8), (64, 1024, 512, 2),
                                           (128, 384, 640, 3), (10, 5120, 5120, 7)])
def test_gemm_split_k_partials(M, N, K, splits):
    x = _rand_bf16(M, K, seed=5)
    w = _rand_bf16(N, K, scale=0.05, seed=6)
    out = torch.full((splits, M, N), float("nan"), dtype=torch.float32, device=DEV)
    L.gemm(x, w, out, L.EPI_F32, splits=splits)
    torch.cuda.synchronize()
    got = out.double().sum(0)
    ref = x.double() @ w.double().T
    err = (got - ref).abs().max().item()
    assert err <= 1e-4 * math.sqrt(K) + 1e-5, f"max err {err}"
    # determinism: a second launch must be bit-identical (fixed reduction order, no atomics)
    out2 = torch.empty_like(out)
    L.gemm(x, w, out2, L.EPI_F32, splits=splits)
    torch.cuda.synchronize()
    assert torch.equal(out, out2)


def _interleave(gate, up):
    from byzantine_consensus_llm_agents_b200.model import interleave_gate_up
    return interleave_gate_up(gate, up)


@pytest.mark.parametrize("M,F,K", [(10, 256, 512), (64, 512, 256), (128, 128, 128), (300, 512, 256),
                                   (1000, 1024, 512)])
def test_gemm_swiglu_epilogue(M, F, K):
    x = _rand_bf16(M, K, seed=7)
    gate = _rand_bf16(F, K, scale=0.05, seed=8)
    up = _rand_bf16(F, K, scale=0.05, seed=9)
    w = _interleave(gate, up).contiguous()
    out = torch.full((M, F), float("nan"), dtype=torch.bfloat16, device=DEV)
    L.gemm(x, w, out, L.EPI_SWIGLU)
    torch.cuda.synchronize()
    g = x.float() @ gate.float().T
    u = x.float() @ up.float().T
    ref = torch.nn.functional.silu(g) * u
    err = (out.float() - ref).abs()
    tol = ref.abs() * 2 ** -7 + 2e-3  # bf16 output ulp + __expf in the epilogue
    assert (err <= tol).all(), f"max err {err.max().item()}"


def test_gemm_batch_invariance():
    """A row's result must not depend on which other rows share the batch."""
    K, N = 1024, 768
    w = _rand_bf16(N, K, scale=0.05, seed=10)
    x = _rand_bf16(40, K, seed=11)
    outs = []
    for M in (1, 10, 16, 17, 40):
        s = L.splits_for(M, N, K)
        o = torch.empty((s, M, N), dtype=torch.float32, device=DEV)
        L.gemm(x[:M].contiguous(), w, o, L.EPI_F32, splits=s)
        outs.append(o.sum(0)[0].clone())
    torch.cuda.synchronize()
    for o in outs[1:]:
        assert torch.equal(outs[0], o)


def test_gemm_rejects_bad_args():
    x = _rand_bf16(4, 64)
    w = _rand_bf16(128, 64)
    out = torch.empty((4, 128), dtype=torch.bfloat16, device=DEV)
    with pytest.raises(L.BcgError):
        L.gemm(x, w, out, L.EPI_BF16, splits=2)  # split-K needs the fp32 epilogue


# ------------------------------------------------------------------ norms / rope / kv write

def _rmsnorm_ref(x, w, eps):
    v = x.pow(2).mean(-1, keepdim=True)
    return x * torch.rsqrt(v + eps) * w


# ------------------------------------------------------------------ GEMM with the norm folded in

def _fusion(**kw):
    f = L.GemmFusion()
    for k, v in kw.items():
        setattr(f, k, v.data_ptr() if torch.is_tensor(v) else v)
    return f


@pytest.mark.parametrize("M,N,K,splits", [(10, 5120, 5120, 7), (10, 5120, 17408, 11), (1, 256, 512, 2),
                                          (16, 1024, 1024, 1), (33, 384, 2048, 16), (128, 256, 256, 3)])
def test_gemm_producer_finishes_residual_and_norm_statistics(M, N, K, splits):
    """bcg_gemm_bf16_tn_fused, producer half: split-K slabs -> residual (same fp32 order as
    bcg_add_rmsnorm), bf16(residual*w), per-tile row sums of squares; counters back at zero."""
    import ctypes as C
    lib = _lib()
    x = _rand_bf16(M, K, seed=31)
    w = _rand_bf16(N, K, scale=0.05, seed=32)
    res0 = torch.randn(M, N, device=DEV)
    res = res0.clone()
    nw = (torch.rand(N, device=DEV) + 0.5).float()
    P = torch.full((splits, M, N), float("nan"), device=DEV)
    xn = torch.zeros(M, N, device=DEV, dtype=torch.bfloat16)
    tiles = N // 128
    sumsq = torch.full((tiles, 128), float("nan"), device=DEV)
    counters = torch.zeros(tiles, device=DEV, dtype=torch.int32)
    f = _fusion(red_residual=res, red_w=nw, red_xn=xn, red_sumsq=sumsq, red_counters=counters)
    for _ in range(2):  # twice: the counters must come back to zero
        res.copy_(res0)
        L.check(lib.bcg_gemm_bf16_tn_fused(x.data_ptr(), w.data_ptr(), P.data_ptr(), M, N, K, K, K, N, L.EPI_F32,
                                           splits, 0, C.byref(f), L.stream_ptr()))
        torch.cuda.synchronize()
        assert int(counters.abs().sum()) == 0
        acc = P[0].clone()
        for sp in range(1, splits):
            acc = acc + P[sp]
        want = res0 + acc
        assert torch.equal(res, want)                                   # same operation order
        assert torch.equal(xn, (want * nw).to(torch.bfloat16))          # r*w rounded once
        ss = want.view(M, tiles, 128).pow(2).sum(-1).T                  # [tiles, M]
        assert torch.allclose(sumsq[:, :M], ss, rtol=1e-5, atol=1e-6)
    # the partial slabs are an ordinary split-K result
    ref = x.float() @ w.float().T
    assert ((acc - ref).abs() <= ref.abs() * 2 ** -9 + 2e-3).all()


@pytest.mark.parametrize("epi", ["f32", "bf16", "swiglu"])
def test_gemm_consumer_applies_row_scale(epi):
    import ctypes as C
    lib = _lib()
    M, N, K, d = 10, 1024, 2048, 2048
    tiles = d // 128
    x = _rand_bf16(M, K, seed=41)
    w = _rand_bf16(N, K, scale=0.05, seed=42)
    sumsq = torch.rand(tiles, 128, device=DEV) * 300 + 1.0
    scale = torch.rsqrt(sumsq[:, :M].sum(0) / d + 1e-6)                  # [M]
    f = _fusion(scale_sumsq=sumsq, scale_tiles=tiles, scale_d=d, scale_eps=1e-6)
    if epi == "f32":
        splits = 3
        plain = torch.zeros(splits, M, N, device=DEV)
        fused = torch.zeros(splits, M, N, device=DEV)
        L.gemm(x, w, plain, L.EPI_F32, splits=splits)
        L.check(lib.bcg_gemm_bf16_tn_fused(x.data_ptr(), w.data_ptr(), fused.data_ptr(), M, N, K, K, K, N,
                                           L.EPI_F32, splits, 0, C.byref(f), L.stream_ptr()))
        torch.cuda.synchronize()
        # one fp32 multiply per element; the tile sums are added in index order
        tot = torch.zeros(M, device=DEV)
        for t in range(tiles):
            tot = tot + sumsq[t, :M]
        exact = torch.rsqrt(tot / d + 1e-6)
        assert torch.allclose(fused, plain * exact[None, :, None], rtol=3e-7, atol=0)
    else:
        ref = (x.float() @ w.float().T) * scale[:, None]
        if epi == "bf16":
            out = torch.zeros(M, N, device=DEV, dtype=torch.bfloat16)
            L.check(lib.bcg_gemm_bf16_tn_fused(x.data_ptr(), w.data_ptr(), out.data_ptr(), M, N, K, K, K, N,
                                               L.EPI_BF16, 1, 0, C.byref(f), L.stream_ptr()))
        else:
            out = torch.zeros(M, N // 2, device=DEV, dtype=torch.bfloat16)
            L.check(lib.bcg_gemm_bf16_tn_fused(x.data_ptr(), w.data_ptr(), out.data_ptr(), M, N, K, K, K, N // 2,
                                               L.EPI_SWIGLU, 1, 0, C.byref(f), L.stream_ptr()))
            r = ref.view(M, N // 32, 2, 16)
            ref = (torch.nn.functional.silu(r[:, :, 0]) * r[:, :, 1]).reshape(M, N // 2)
        torch.cuda.synchronize()
        assert ((out.float() - ref).abs() <= ref.abs() * 2 ** -6 + 2e-3).all()


def test_gemm_fusion_argument_checks():
    import ctypes as C
    lib = _lib()
    M, N, K = 4, 256, 256
    x, w = _rand_bf16(M, K), _rand_bf16(N, K)
    P = torch.zeros(2, M, N, device=DEV)
    res = torch.zeros(M, N, device=DEV)
    nw = torch.ones(N, device=DEV)
    xn = torch.zeros(M, N, device=DEV, dtype=torch.bfloat16)
    sumsq = torch.zeros(2, 128, device=DEV)
    cnt = torch.zeros(2, device=DEV, dtype=torch.int32)
    f = _fusion(red_residual=res, red_w=nw, red_xn=xn, red_sumsq=sumsq, red_counters=cnt)
    call = lambda m, n, epi, ld: lib.bcg_gemm_bf16_tn_fused(x.data_ptr(), w.data_ptr(), P.data_ptr(), m, n, K, K, K,
                                                            ld, epi, 1, 0, C.byref(f), L.stream_ptr())
    assert call(M, N, L.EPI_BF16, N) != 0           # producer needs the fp32 partial epilogue
    assert call(M, 192, L.EPI_F32, 192) != 0        # N % 128
    assert call(M, N, L.EPI_F32, N + 8) != 0        # ld_out == N
    assert b"fusion" in lib.bcg_last_error()
    assert call(M, N, L.EPI_F32, N) == 0
    torch.cuda.synchronize()


@pytest.mark.parametrize("T,d", [(1, 256), (10, 1024), (37, 5120), (5, 8192)])
def test_embed_rmsnorm(T, d):
    lib = _lib()
    V = 1000
    embed = _rand_bf16(V, d, scale=0.02, seed=12)
    w = (torch.rand(d, device=DEV) + 0.5).float()
    ids = torch.randint(0, V, (T,), device=DEV, dtype=torch.int32)
    res = torch.empty(T, d, device=DEV)
    xn = torch.empty(T, d, device=DEV, dtype=torch.bfloat16)
    L.check(lib.bcg_embed_rmsnorm(ids.data_ptr(), embed.data_ptr(), w.data_ptr(), res.data_ptr(),
                                  xn.data_ptr(), T, d, 1e-6, L.stream_ptr()))
    torch.cuda.synchronize()
    ref_res = embed[ids.long()].float()
    assert torch.equal(res, ref_res)
    ref = _rmsnorm_ref(ref_res, w, 1e-6)
    assert ((xn.float() - ref).abs() <= ref.abs() * 2 ** -7 + 1e-5).all()


@pytest.mark.parametrize("T,d,parts", [(10, 1024, 0), (10, 5120, 4), (33, 2048, 1), (3, 8192, 7),
                                       (200, 5120, 11), (129, 1024, 0), (7, 6144, 17), (5, 1000, 3),
                                       (300, 16384, 2), (2, 16384, 5), (1, 256, 1)])
def test_add_rmsnorm(T, d, parts):
    lib = _lib()
    res0 = torch.randn(T, d, device=DEV)
    w = (torch.rand(d, device=DEV) + 0.5).float()
    res = res0.clone()
    xn = torch.empty(T, d, device=DEV, dtype=torch.bfloat16)
    if parts == 0:
        delta = _rand_bf16(T, d, seed=13)
        L.check(lib.bcg_add_rmsnorm(res.data_ptr(), delta.data_ptr(), 0, 1, 0, w.data_ptr(),
                                    xn.data_ptr(), None, T, d, 1e-6, L.stream_ptr()))
        dsum = delta.float()
    else:
        delta = torch.randn(parts, T, d, device=DEV)
        L.check(lib.bcg_add_rmsnorm(res.data_ptr(), delta.data_ptr(), 1, parts, T * d, w.data_ptr(),
                                    xn.data_ptr(), None, T, d, 1e-6, L.stream_ptr()))
        dsum = delta[0].clone()
        for s in range(1, parts):
            dsum = dsum + delta[s]
    torch.cuda.synchronize()
    ref_res = res0 + dsum
    assert torch.equal(res, ref_res)  # same fp32 operation order
    ref = _rmsnorm_ref(ref_res, w, 1e-6)
    assert ((xn.float() - ref).abs() <= ref.abs() * 2 ** -7 + 1e-5).all()


@pytest.mark.parametrize("d,parts", [(5120, 11), (1024, 0), (1000, 3), (12288, 2)])
def test_add_rmsnorm_row_is_independent_of_launch_size(d, parts):
    """Decode-sized launches (<= 128 rows: one 8-CTA cluster per row) and prefill-sized ones (one CTA
    per row) form the sum of squares in the same order: a row's xn is bit-identical either way."""
    lib = _lib()
    few, many = 6, 200
    res_m = torch.randn(many, d, device=DEV)
    w = (torch.rand(d, device=DEV) + 0.5).float()
    if parts == 0:
        delta_m = _rand_bf16(many, d, seed=21)
    else:
        delta_m = torch.randn(parts, many, d, device=DEV)
    pick = [0, 57, 128, 199, 3, 77]
    res_f = res_m[pick].clone()
    delta_f = delta_m[pick].clone() if parts == 0 else delta_m[:, pick].contiguous()
    xn_m = torch.empty(many, d, device=DEV, dtype=torch.bfloat16)
    xn_f = torch.empty(few, d, device=DEV, dtype=torch.bfloat16)
    for res, delta, xn, T in ((res_m, delta_m, xn_m, many), (res_f, delta_f, xn_f, few)):
        L.check(lib.bcg_add_rmsnorm(res.data_ptr(), delta.data_ptr(), 1 if parts else 0, max(parts, 1),
                                    T * d, w.data_ptr(), xn.data_ptr(), None, T, d, 1e-6, L.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(res_m[pick], res_f)
    assert torch.equal(xn_m[pick].view(torch.int16), xn_f.view(torch.int16))


@pytest.mark.parametrize("T,d,parts", [(10, 5120, 11), (1, 256, 1), (128, 1024, 3), (7, 6144, 0), (5, 1000, 2)])
def test_add_norm_stats_plus_consumer_scale_equals_add_rmsnorm(T, d, parts):
    """bcg_add_norm_stats leaves the residual exactly as bcg_add_rmsnorm does, xn = bf16(r*w), and
    segment sums from which the consuming GEMM's epilogue rebuilds 1/RMS."""
    import ctypes as C
    lib = _lib()
    res0 = torch.randn(T, d, device=DEV)
    w = (torch.rand(d, device=DEV) + 0.5).float()
    if parts == 0:
        delta, is_f32, n = _rand_bf16(T, d, seed=51), 0, 1
    else:
        delta, is_f32, n = torch.randn(parts, T, d, device=DEV), 1, parts
    res_a, res_b = res0.clone(), res0.clone()
    xn_a = torch.empty(T, d, device=DEV, dtype=torch.bfloat16)
    xn_b = torch.empty(T, d, device=DEV, dtype=torch.bfloat16)
    sumsq = torch.full((8, 128), float("nan"), device=DEV)
    L.check(lib.bcg_add_rmsnorm(res_a.data_ptr(), delta.data_ptr(), is_f32, n, T * d, w.data_ptr(),
                                xn_a.data_ptr(), None, T, d, 1e-6, L.stream_ptr()))
    L.check(lib.bcg_add_norm_stats(res_b.data_ptr(), delta.data_ptr(), is_f32, n, T * d, w.data_ptr(),
                                   xn_b.data_ptr(), sumsq.data_ptr(), T, d, L.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(res_a, res_b)
    assert torch.equal(xn_b, (res_b * w).to(torch.bfloat16))
    tot = sumsq[:, :T].sum(0)
    assert torch.allclose(tot, res_b.pow(2).sum(-1), rtol=1e-5)
    # a projection fed with xn_b and the table reproduces the projection of the normalised row
    if d % 8 == 0:
        N = 256
        wt = _rand_bf16(N, d, scale=0.05, seed=52)
        f = _fusion(scale_sumsq=sumsq, scale_tiles=8, scale_d=d, scale_eps=1e-6)
        got = torch.zeros(T, N, device=DEV)
        L.check(lib.bcg_gemm_bf16_tn_fused(xn_b.data_ptr(), wt.data_ptr(), got.data_ptr(), T, N, d, d, d, N,
                                           L.EPI_F32, 1, 0, C.byref(f), L.stream_ptr()))
        ref = torch.zeros(T, N, device=DEV)
        L.gemm(xn_a, wt, ref, L.EPI_F32)
        torch.cuda.synchronize()
        assert ((got - ref).abs() <= ref.abs() * 2 ** -6 + 2e-2 * ref.abs().mean()).all()
    assert lib.bcg_add_norm_stats(res_b.data_ptr(), delta.data_ptr(), is_f32, n, T * d, w.data_ptr(),
                                  xn_b.data_ptr(), sumsq.data_ptr(), 129, d, L.stream_ptr()) != 0


def test_add_rmsnorm_gather_rows():
    lib = _lib()
    T, d, n = 50, 1024, 4
    res = torch.randn(T, d, device=DEV)
    keep = res.clone()
    delta = _rand_bf16(T, d, seed=14)
    w = torch.ones(d, device=DEV)
    rows = torch.tensor([49, 0, 17, 33], device=DEV, dtype=torch.int32)
    xn = torch.empty(n, d, device=DEV, dtype=torch.bfloat16)
    L.check(lib.bcg_add_rmsnorm(res.data_ptr(), delta.data_ptr(), 0, 1, 0, w.data_ptr(),
                                xn.data_ptr(), rows.data_ptr(), n, d, 1e-6, L.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.equal(res, keep)
    ref = _rmsnorm_ref((keep + delta.float())[rows.long()], w, 1e-6)
    assert ((xn.float() - ref).abs() <= ref.abs() * 2 ** -7 + 1e-5).all()


def _rope_ref(x, pos, cos_sin):
    # x [T, H, 128] fp32; rotate-half as transformers apply_rotary_pos_emb
    cos = torch.cat([cos_sin[pos, :64], cos_sin[pos, :64]], -1)[:, None, :]
    sin = torch.cat([cos_sin[pos, 64:], cos_sin[pos, 64:]], -1)[:, None, :]
    x1, x2 = x[..., :64], x[..., 64:]
    rot = torch.cat([-x2, x1], -1)
    return x * cos + rot * sin


def _make_cos_sin(max_pos=4096, theta=1e6):
    inv = 1.0 / (theta ** (torch.arange(0, 128, 2, dtype=torch.int64).float() / 128))
    fr = torch.arange(max_pos, dtype=torch.float32)[:, None] * inv[None, :]
    return torch.cat([fr.cos(), fr.sin()], 1).contiguous().to(DEV)


@pytest.mark.parametrize("T,H,HKV,parts,qknorm", [(10, 4, 2, 0, True), (10, 40, 8, 3, True),
                                                   (70, 16, 8, 0, True), (5, 6, 1, 2, False),
                                                   # prefill-sized fp32 slab: the token-per-CTA kernel
                                                   (300, 40, 8, 1, True), (700, 64, 8, 1, True),
                                                   (257, 48, 8, 1, False), (130, 3, 1, 1, True)])
def test_qknorm_rope_kvwrite(T, H, HKV, parts, qknorm):
    lib = _lib()
    page = 16
    n_pages = max(64, (T + 15) // 16 + 8)
    qkv_dim = (H + 2 * HKV) * 128
    cos_sin = _make_cos_sin()
    pos = torch.randint(0, 4000, (T,), device=DEV, dtype=torch.int32)
    slots = torch.randperm(n_pages * page, device=DEV)[:T].to(torch.int32)
    slots[0] = -1  # skipped write
    qw = (torch.rand(128, device=DEV) + 0.5) if qknorm else None
    kw = (torch.rand(128, device=DEV) + 0.5) if qknorm else None
    cache = torch.zeros(n_pages, 2, HKV, page, 128, device=DEV, dtype=torch.bfloat16)
    q_out = torch.empty(T, H * 128, device=DEV, dtype=torch.bfloat16)
    if parts == 0:
        qkv = _rand_bf16(T, qkv_dim, seed=15)
        full = qkv.float()
        L.check(lib.bcg_qknorm_rope_kvwrite(qkv.data_ptr(), 0, 1, 0, qkv_dim, L.ptr(qw), L.ptr(kw),
                                            cos_sin.data_ptr(), pos.data_ptr(), slots.data_ptr(),
                                            q_out.data_ptr(), cache.data_ptr(), T, H, HKV, page, 1e-6,
                                            L.stream_ptr()))
    else:
        qkv = torch.randn(parts, T, qkv_dim, device=DEV)
        full = qkv[0].clone()
        for s in range(1, parts):
            full = full + qkv[s]
        L.check(lib.bcg_qknorm_rope_kvwrite(qkv.data_ptr(), 1, parts, T * qkv_dim, qkv_dim, L.ptr(qw),
                                            L.ptr(kw), cos_sin.data_ptr(), pos.data_ptr(),
                                            slots.data_ptr(), q_out.data_ptr(), cache.data_ptr(), T, H,
                                            HKV, page, 1e-6, L.stream_ptr()))
    torch.cuda.synchronize()
    q = full[:, :H * 128].view(T, H, 128)
    k = full[:, H * 128:(H + HKV) * 128].view(T, HKV, 128)
    v = full[:, (H + HKV) * 128:].view(T, HKV, 128)
    if qknorm:
        q = _rmsnorm_ref(q, qw, 1e-6)
        k = _rmsnorm_ref(k, kw, 1e-6)
    q = _rope_ref(q, pos.long(), cos_sin)
    k = _rope_ref(k, pos.long(), cos_sin)
    tol = lambda r: r.abs() * 2 ** -7 + 1e-4
    assert ((q_out.float().view(T, H, 128) - q).abs() <= tol(q)).all()
    for t in range(T):
        s = int(slots[t])
        if s < 0:
            continue
        p, o = divmod(s, page)
        assert ((cache[p, 0, :, o].float() - k[t]).abs() <= tol(k[t])).all()
        assert ((cache[p, 1, :, o].float() - v[t]).abs() <= tol(v[t])).all()
    # the skipped token wrote nothing anywhere: total non-zero slots == T-1
    written = (cache.float().abs().sum(dim=(1, 2, 4)) > 0).sum().item()
    assert written == T - 1


def test_swiglu_reduce():
    lib = _lib()
    T, F, parts = 10, 256, 3
    gate = torch.randn(parts, T, F, device=DEV)
    up = torch.randn(parts, T, F, device=DEV)
    inter = torch.stack([gate.view(parts, T, F // 16, 16), up.view(parts, T, F // 16, 16)], dim=3)
    inter = inter.reshape(parts, T, 2 * F).contiguous()
    act = torch.empty(T, F, device=DEV, dtype=torch.bfloat16)
    L.check(lib.bcg_swiglu_reduce(inter.data_ptr(), parts, T * 2 * F, act.data_ptr(), T, 2 * F,
                                  L.stream_ptr()))
    torch.cuda.synchronize()
    g = gate[0] + gate[1] + gate[2]
    u = up[0] + up[1] + up[2]
    ref = torch.nn.functional.silu(g) * u
    assert ((act.float() - ref).abs() <= ref.abs() * 2 ** -7 + 2e-3).all()


# ------------------------------------------------------------------ attention

def _build_paged(seqs_kv, HKV, page=16, n_pages=None, seed=0):
    """seqs_kv: list of lengths.  Returns cache, block_table, K list, V list (fp32 per seq)."""
    g = torch.Generator().manual_seed(seed)
    total_pages = sum((s + page - 1) // page for s in seqs_kv)
    n_pages = n_pages or total_pages + 3
    perm = torch.randperm(n_pages, generator=g).tolist()
    cache = torch.zeros(n_pages, 2, HKV, page, 128, dtype=torch.bfloat16)
    max_blocks = max((s + page - 1) // page for s in seqs_kv)
    bt = torch.zeros(len(seqs_kv), max_blocks, dtype=torch.int32)
    Ks, Vs = [], []
    nxt = 0
    for b, s in enumerate(seqs_kv):
        k = torch.randn(s, HKV, 128, generator=g).to(torch.bfloat16)
        v = torch.randn(s, HKV, 128, generator=g).to(torch.bfloat16)
        Ks.append(k.float().to(DEV))
        Vs.append(v.float().to(DEV))
        for j in range((s + page - 1) // page):
            p = perm[nxt]
            nxt += 1
            bt[b, j] = p
            n = min(page, s - j * page)
            cache[p, 0, :, :n] = k[j * page:j * page + n].transpose(0, 1)
            cache[p, 1, :, :n] = v[j * page:j * page + n].transpose(0, 1)
    return cache.to(DEV), bt.to(DEV), Ks, Vs


def _attn_ref(q, K, V, G, causal_pos=None):
    # q [Tq, H, 128] fp32, K,V [S, HKV, 128] fp32
    Kx = K.repeat_interleave(G, dim=1)
    Vx = V.repeat_interleave(G, dim=1)
    s = torch.einsum("thd,shd->hts", q, Kx) / math.sqrt(128)
    if causal_pos is not None:
        S = K.shape[0]
        mask = torch.arange(S, device=q.device)[None, :] > causal_pos[:, None]
        s = s.masked_fill(mask[None], float("-inf"))
    p = torch.softmax(s, dim=-1)
    return torch.einsum("hts,shd->thd", p, Vx)


@pytest.mark.parametrize("H,HKV,lens", [(4, 2, [1, 16, 17, 300]), (40, 8, [2000, 1700, 5, 256, 257]),
                                        (16, 8, [1023, 1024, 1025]), (32, 8, [700, 3]),
                                        (48, 8, [513]), (64, 8, [100, 900]),
                                        # more than 8 chunks per row: the merge's long-row branch
                                        (8, 2, [4100, 2049, 300, 2048]), (40, 8, [8000])])
def test_decode_attn_paged(H, HKV, lens):
    lib = _lib()
    B = len(lens)
    G = H // HKV
    cache, bt, Ks, Vs = _build_paged(lens, HKV, seed=20)
    q = _rand_bf16(B, H * 128, seed=21)
    seq = torch.tensor(lens, device=DEV, dtype=torch.int32)
    max_splits = (max(lens) + 255) // 256 + 1
    so = torch.empty(B * H * max_splits * 128, device=DEV)
    sml = torch.empty(B * H * max_splits * 2, device=DEV)
    cnt = torch.zeros(B * HKV, device=DEV, dtype=torch.int32)
    out = torch.empty(B, H * 128, device=DEV, dtype=torch.bfloat16)
    for _ in range(2):  # twice: the counters must come back to zero
        L.check(lib.bcg_decode_attn_paged(q.data_ptr(), cache.data_ptr(), bt.data_ptr(), bt.shape[1],
                                          seq.data_ptr(), out.data_ptr(), so.data_ptr(), sml.data_ptr(),
                                          cnt.data_ptr(), B, H, HKV, 16, max_splits, 1.0 / math.sqrt(128),
                                          L.stream_ptr()))
    assert int(cnt.abs().sum()) == 0
    torch.cuda.synchronize()
    for b in range(B):
        ref = _attn_ref(q[b].float().view(1, H, 128), Ks[b], Vs[b], G)[0]
        got = out[b].float().view(H, 128)
        err = (got - ref).abs().max().item()
        assert err <= 2e-2, f"row {b} len {lens[b]} err {err}"  # bf16 output of O(1) values


@pytest.mark.parametrize("H,HKV,lens,parts,qknorm", [
    (40, 8, [1600, 257, 1, 16, 17, 512, 300], 5, True),      # 14B shape: chunk / page boundaries, 1-token row
    (8, 8, [33, 700], 1, True),                               # G = 1
    (48, 8, [1025, 64], 3, False),                            # no QK-norm (Mistral), G = 6
    (64, 8, [2049, 900, 4100], 2, True),                      # G = 8, rows with more than 8 chunks
    (4, 2, [5, 256], 11, True)])
def test_decode_attn_from_qkv_slabs_equals_rope_kernel_plus_attention(H, HKV, lens, parts, qknorm):
    """bcg_decode_attn_paged_qkv == bcg_qknorm_rope_kvwrite followed by bcg_decode_attn_paged: the KV
    cache ends up bit-identical (the new token's K/V use the rope kernel's arithmetic), the outputs
    agree to bf16 rounding of the queries (their RMS is summed in a different lane order)."""
    lib = _lib()
    B, G = len(lens), H // HKV
    qkv_dim = (H + 2 * HKV) * 128
    cache0, bt, _, _ = _build_paged(lens, HKV, seed=60)
    # garbage where the new token goes: neither path may read it
    cos_sin = _make_cos_sin(max_pos=8192)
    pos = torch.tensor([n - 1 for n in lens], device=DEV, dtype=torch.int32)
    slots = torch.tensor([int(bt[b, (n - 1) // 16]) * 16 + (n - 1) % 16 for b, n in enumerate(lens)],
                         device=DEV, dtype=torch.int32)
    for b, n in enumerate(lens):
        cache0[int(bt[b, (n - 1) // 16]), :, :, (n - 1) % 16] = float("nan")
    slots_skip = slots.clone()
    if B > 2:
        slots_skip[2] = -1   # a row that must not write the cache (its attention still sees the new token)
    seq = torch.tensor(lens, device=DEV, dtype=torch.int32)
    P = torch.randn(parts, B, qkv_dim, device=DEV) * 0.7
    qw = (torch.rand(128, device=DEV) + 0.5) if qknorm else None
    kw = (torch.rand(128, device=DEV) + 0.5) if qknorm else None
    max_splits = (max(lens) + 255) // 256
    so = torch.empty(B * H * max_splits * 128, device=DEV)
    sml = torch.empty(B * H * max_splits * 2, device=DEV)
    cnt = torch.zeros(B * HKV, device=DEV, dtype=torch.int32)
    scale = 1.0 / math.sqrt(128)
    # reference path: two launches
    cache_a = cache0.clone()
    q = torch.empty(B, H * 128, device=DEV, dtype=torch.bfloat16)
    out_a = torch.empty(B, H * 128, device=DEV, dtype=torch.bfloat16)
    L.check(lib.bcg_qknorm_rope_kvwrite(P.data_ptr(), 1, parts, B * qkv_dim, qkv_dim, L.ptr(qw), L.ptr(kw),
                                        cos_sin.data_ptr(), pos.data_ptr(), slots.data_ptr(), q.data_ptr(),
                                        cache_a.data_ptr(), B, H, HKV, 16, 1e-6, L.stream_ptr()))
    L.check(lib.bcg_decode_attn_paged(q.data_ptr(), cache_a.data_ptr(), bt.data_ptr(), bt.shape[1], seq.data_ptr(),
                                      out_a.data_ptr(), so.data_ptr(), sml.data_ptr(), cnt.data_ptr(), B, H, HKV,
                                      16, max_splits, scale, L.stream_ptr()))
    torch.cuda.synchronize()
    for trial_slots in (slots, slots_skip):
        cache_b = cache0.clone()
        out_b = torch.full((B, H * 128), float("nan"), device=DEV, dtype=torch.bfloat16)
        for _ in range(2):  # second call: the row is already in the cache, the result must not change
            L.check(lib.bcg_decode_attn_paged_qkv(P.data_ptr(), parts, B * qkv_dim, qkv_dim, L.ptr(qw), L.ptr(kw),
                                                  cos_sin.data_ptr(), pos.data_ptr(), trial_slots.data_ptr(), 1e-6,
                                                  cache_b.data_ptr(), bt.data_ptr(), bt.shape[1], seq.data_ptr(),
                                                  out_b.data_ptr(), so.data_ptr(), sml.data_ptr(), cnt.data_ptr(),
                                                  B, H, HKV, 16, max_splits, scale, L.stream_ptr()))
            torch.cuda.synchronize()
            assert int(cnt.abs().sum()) == 0
            assert torch.isfinite(out_b.float()).all()
            assert (out_b.float() - out_a.float()).abs().max().item() <= 1.5e-2
        written = [b for b in range(B) if int(trial_slots[b]) >= 0]
        for b in range(B):
            pg, off = int(bt[b, (lens[b] - 1) // 16]), (lens[b] - 1) % 16
            if b in written:
                assert torch.equal(cache_b[pg, :, :, off].view(torch.int16), cache_a[pg, :, :, off].view(torch.int16))
            else:
                assert torch.isnan(cache_b[pg, :, :, off].float()).all()   # untouched


def test_decode_attn_batch_invariance():
    lib = _lib()
    H, HKV = 40, 8
    lens = [1500, 37, 800]
    cache, bt, Ks, Vs = _build_paged(lens, HKV, seed=22)
    q = _rand_bf16(3, H * 128, seed=23)

    def run(rows, max_splits):
        B = len(rows)
        qq = q[rows].contiguous()
        btt = bt[rows].contiguous()
        seq = torch.tensor([lens[r] for r in rows], device=DEV, dtype=torch.int32)
        so = torch.empty(B * H * max_splits * 128, device=DEV)
        sml = torch.empty(B * H * max_splits * 2, device=DEV)
        cnt = torch.zeros(B * HKV, device=DEV, dtype=torch.int32)
        out = torch.empty(B, H * 128, device=DEV, dtype=torch.bfloat16)
        L.check(lib.bcg_decode_attn_paged(qq.data_ptr(), cache.data_ptr(), btt.data_ptr(), btt.shape[1],
                                          seq.data_ptr(), out.data_ptr(), so.data_ptr(), sml.data_ptr(),
                                          cnt.data_ptr(), B, H, HKV, 16, max_splits, 1.0 / math.sqrt(128),
                                          L.stream_ptr()))
        torch.cuda.synchronize()
        return out

    full = run([0, 1, 2], 8)
    solo = run([2], 4)
    assert torch.equal(full[2], solo[0])


@pytest.mark.parametrize("H,HKV,spec", [
    (4, 2, [(0, 1)]),                       # single token, no prefix
    (4, 2, [(0, 64), (0, 65), (0, 17)]),    # ragged rows inside one tile
    (4, 2, [(0, 128), (0, 129), (0, 127), (0, 257)]),   # query-tile / KV-block boundaries
    (16, 8, [(0, 200), (48, 130), (160, 1)]),  # cached prefixes (multiples of the page size)
    (40, 8, [(416, 1300), (0, 700)]),
    (8, 1, [(112, 16), (1008, 40), (2032, 300)]),       # prefix ends 16 short of a KV block; G = 8
    (48, 8, [(0, 1700)]),                               # Mistral head layout, G = 6, 14 KV blocks
])
def test_prefill_attn_paged(H, HKV, spec):
    """spec: list of (cached_prefix_len, new_tokens)."""
    _check_prefill_attn(H, HKV, spec, 1.0)


@pytest.mark.parametrize("H,HKV,spec", [
    (48, 8, [(1984, 3), (368, 1641), (2000, 9)]),       # Mistral head layout, long cached prefixes, few new tokens
    (40, 8, [(0, 700), (512, 130)]),
    (4, 2, [(0, 300)]),
])
def test_prefill_attn_paged_peaky_scores(H, HKV, spec):
    """Queries 3x larger: score maxima differ by far more than 8 (log2) between KV blocks, so the
    kernel's lazy rescale of the TMEM-resident output runs in most blocks instead of almost never
    (models without QK-norm — Mistral — produce such scores with random-init weights)."""
    _check_prefill_attn(H, HKV, spec, 3.0)


def _check_prefill_attn(H, HKV, spec, q_scale):
    TILE = 128
    lib = _lib()
    G = H // HKV
    lens = [c + n for c, n in spec]
    cache, bt, Ks, Vs = _build_paged(lens, HKV, seed=24)
    T = sum(n for _, n in spec)
    q = _rand_bf16(T, H * 128, scale=q_scale, seed=25)
    tiles = []
    row = 0
    for b, (c, n) in enumerate(spec):
        for r0 in range(0, n, TILE):
            tiles.append([b, row + r0, c + r0, min(TILE, n - r0)])
        row += n
    ti = torch.tensor(tiles, device=DEV, dtype=torch.int32)
    out = torch.full((T, H * 128), float("nan"), device=DEV, dtype=torch.bfloat16)
    L.check(lib.bcg_prefill_attn_paged(q.data_ptr(), cache.data_ptr(), bt.data_ptr(), bt.shape[1],
                                       ti.data_ptr(), len(tiles), out.data_ptr(), T, H, HKV, 16,
                                       1.0 / math.sqrt(128), L.stream_ptr()))
    torch.cuda.synchronize()
    row = 0
    for b, (c, n) in enumerate(spec):
        qq = q[row:row + n].float().view(n, H, 128)
        pos = torch.arange(c, c + n, device=DEV)
        ref = _attn_ref(qq, Ks[b], Vs[b], G, causal_pos=pos)
        got = out[row:row + n].float().view(n, H, 128)
        assert torch.isfinite(got).all(), f"seq {b}: non-finite output"
        err = (got - ref).abs().max().item()
        assert err <= 3e-2, f"seq {b} err {err}"  # P rounded to bf16 before PV + bf16 output
        row += n
