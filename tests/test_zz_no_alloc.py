"""bcg_decode_attn_paged_static: decode attention for callers that cannot allocate device memory.

Who needs it: a host that sized weights + KV pages to fill its device-memory budget at start-up
and so has nothing left to cudaMalloc when attention runs, or a decode loop that must never call
cudaMalloc/cudaFree because they stall the whole device.  The library keeps the chunk scratch in
a fixed 16 MiB device array of its own instead.

The test calls the entry point through the C ABI twice in a row — the second time on a different,
shorter batch, so anything stale left in the shared array by the first call would show — and
checks each result two ways: bit for bit against the existing caller-scratch entry point (both run
the same kernels in the same order, so nothing short of equality is acceptable), and — so that the
check does not rest on a sibling alone — against an fp32 softmax(QK^T)V computed by torch over the
gathered pages, within the 2e-2 tolerance of the other decode-attention tests.  Collected last on
purpose (file name): a run that stops at the first failure has then already run every other test.
It needs no state from other tests and passes in a fresh process as well.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from byzantine_consensus_llm_agents_b200 import lib as L  # noqa: E402

DEV = "cuda"
TOL = 2e-2


def _paged_case(lens, H, HKV, seed):
    g = torch.Generator().manual_seed(seed)
    page = 16
    blocks = [(n + page - 1) // page for n in lens]
    n_pages = sum(blocks) + 2
    perm = torch.randperm(n_pages, generator=g)
    cache = torch.randn(n_pages, 2, HKV, page, 128, generator=g).to(torch.bfloat16)
    bt = torch.zeros(len(lens), max(blocks), dtype=torch.int32)
    nxt = 0
    for b, nb in enumerate(blocks):
        bt[b, :nb] = perm[nxt:nxt + nb].to(torch.int32)
        nxt += nb
    q = torch.randn(len(lens), H * 128, generator=g).to(torch.bfloat16)
    return q.to(DEV), cache.to(DEV), bt.to(DEV), torch.tensor(lens, dtype=torch.int32, device=DEV)


def _run_both(lib, lens, H, HKV, seed):
    q, cache, bt, seq = _paged_case(lens, H, HKV, seed)
    B = len(lens)
    max_splits = (max(lens) + 255) // 256
    scale = 1.0 / math.sqrt(128)
    # existing entry point: caller-owned scratch
    so = torch.empty(B * H * max_splits * 128, device=DEV)
    sml = torch.empty(B * H * max_splits * 2, device=DEV)
    cnt = torch.zeros(B * HKV, device=DEV, dtype=torch.int32)
    want = torch.empty(B, H * 128, device=DEV, dtype=torch.bfloat16)
    L.check(lib.bcg_decode_attn_paged(q.data_ptr(), cache.data_ptr(), bt.data_ptr(), bt.shape[1],
                                      seq.data_ptr(), want.data_ptr(), so.data_ptr(), sml.data_ptr(),
                                      cnt.data_ptr(), B, H, HKV, 16, max_splits, scale, L.stream_ptr()))
    # new entry point: inputs and outputs only
    got = torch.full((B, H * 128), float("nan"), device=DEV, dtype=torch.bfloat16)
    L.check(lib.bcg_decode_attn_paged_static(q.data_ptr(), cache.data_ptr(), bt.data_ptr(), bt.shape[1],
                                             seq.data_ptr(), got.data_ptr(), B, H, HKV, 16, max_splits,
                                             scale, L.stream_ptr()))
    torch.cuda.synchronize()
    return want.float(), got.float(), _fp32_reference(q, cache, bt, lens, H, HKV)


def _fp32_reference(q, cache, bt, lens, H, HKV):
    """softmax(q K^T / sqrt(128)) V per row in fp32, K/V gathered from the pages by the block table."""
    G = H // HKV
    out = torch.empty(len(lens), H * 128, device=q.device)
    for b, n in enumerate(lens):
        pages = bt[b, :(n + 15) // 16].long()
        kv = cache[pages].float()                                   # [P, 2, HKV, 16, 128]
        K = kv[:, 0].permute(1, 0, 2, 3).reshape(HKV, -1, 128)[:, :n]
        V = kv[:, 1].permute(1, 0, 2, 3).reshape(HKV, -1, 128)[:, :n]
        qb = q[b].float().view(HKV, G, 128)
        p = torch.softmax(torch.einsum("kgd,ksd->kgs", qb, K) / math.sqrt(128), dim=-1)
        out[b] = torch.einsum("kgs,ksd->kgd", p, V).reshape(-1)
    return out


def test_static_scratch_matches_caller_scratch_on_two_calls_in_a_row():
    lib = L.load()
    # call 1: long rows fill many chunks of the shared array
    want, got, ref = _run_both(lib, [2000, 1300, 257, 40], 40, 8, seed=1)
    assert torch.isfinite(got).all()
    assert torch.equal(want, got)
    assert (got - ref).abs().max().item() <= TOL
    # call 2: fewer, shorter rows and another head layout reuse the same (now dirty) array
    want, got, ref = _run_both(lib, [17, 300], 16, 8, seed=2)
    assert torch.isfinite(got).all()
    assert torch.equal(want, got)
    assert (got - ref).abs().max().item() <= TOL


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_static_scratch_on_a_second_device():
    """The array's address and the kernel's shared-memory limit are per device: a process that
    drives two GPUs must be able to make its first call on cuda:1 after having used cuda:0."""
    global DEV
    lib = L.load()
    _run_both(lib, [300, 40], 16, 8, seed=4)                 # cuda:0 first
    DEV = "cuda:1"
    try:
        with torch.cuda.device(1):
            want, got, ref = _run_both(lib, [700, 17, 256], 40, 8, seed=5)
            assert torch.equal(want, got)
            assert (got - ref).abs().max().item() <= TOL
    finally:
        DEV = "cuda"


def test_static_scratch_rejects_oversized_problems():
    lib = L.load()
    q, cache, bt, seq = _paged_case([16, 16], 40, 8, seed=3)
    out = torch.empty(2, 40 * 128, device=DEV, dtype=torch.bfloat16)
    # 128 rows x 64 heads x 32 chunks x 130 floats = 34 M floats > the 4 Mi-float array
    rc = lib.bcg_decode_attn_paged_static(q.data_ptr(), cache.data_ptr(), bt.data_ptr(), bt.shape[1],
                                          seq.data_ptr(), out.data_ptr(), 128, 64, 8, 16, 32,
                                          1.0 / math.sqrt(128), L.stream_ptr())
    assert rc == 4  # BCG_STATUS_TOO_LARGE
    msg = lib.bcg_last_error().decode()
    assert "16 MiB" in msg and "bcg_decode_attn_paged" in msg
