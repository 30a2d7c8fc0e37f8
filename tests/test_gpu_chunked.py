"""GPU parity of the context path: query rows that are only the tail of
their sequence, keys read back from the paged KV cache through the block table -- chunked
prefill, prefix-cache hits and decode steps, as the reference backend serves them with
flash_attn_varlen_func(q, key_cache, value_cache, cu_seqlens_q, seqused_k, block_table)
(vllm/v1/attention/backends/flash_attn.py:518-548).

Oracle: the reference op's arithmetic on the chosen rows of the FULL sequence
(oracle.attention_rows_fp32: mask rows from the closed form with N = the sequence length so far)."""
import pytest
import torch

from oracle import trianglemix_oracle as orc

pytestmark = pytest.mark.gpu
ATOL, RTOL = 1.5e-2, 1e-2
DEV = "cuda:0"
D = 128


@pytest.fixture(scope="module")
def tmx():
    import vllm_trianglemix_b200 as m
    return m


@pytest.fixture(scope="module")
def ops():
    from vllm_trianglemix_b200 import ops as o
    return o


def _seqs(seq_lens, hq, hkv, seed):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(n, hq, D, generator=g).bfloat16(), torch.randn(n, hkv, D, generator=g).bfloat16(),
             torch.randn(n, hkv, D, generator=g).bfloat16()) for n in seq_lens]


def _cu(lens):
    cu = [0]
    for n in lens:
        cu.append(cu[-1] + n)
    return cu


def _oracle(seqs, q_lens, scale, si, sl, last, mode, rows_cap=192):
    """Per sequence: (row indices inside the chunk, oracle output for them)."""
    res = []
    for (q, k, v), ql in zip(seqs, q_lens):
        n = q.shape[0]
        r0 = n - ql
        if ql <= rows_cap:
            res.append((torch.arange(ql), orc.attention_rows_fp32(q, k, v, scale, n, si, sl, last, mode, r0, n)))
        else:   # head, a stretch around the first dense row, tail
            lrow = max(r0, min(n, n - last + 1))
            picks = sorted(set(list(range(r0, r0 + 48)) + list(range(max(r0, lrow - 40), min(n, lrow + 40))) +
                               list(range(n - 48, n))))
            outs = [orc.attention_rows_fp32(q, k, v, scale, n, si, sl, last, mode, i, i + 1) for i in picks]
            res.append((torch.tensor(picks) - r0, torch.cat(outs, dim=0)))
    return res


def _compare(out, q_lens, oracle, what):
    cu = _cu(q_lens)
    for b, (rows, ref) in enumerate(oracle):
        got = out[cu[b]:cu[b + 1]].cpu()[rows]
        rep = orc.error_report(got, ref, ATOL, RTOL)
        assert rep["violations"] == 0, (what, b, rep)


def _paged_cache(seqs, page, seed, extra_pages=3):
    """Scatter every sequence's K/V into a zero-initialised [2, num_blocks, page, Hkv, D] cache through a
    shuffled block table (page 0 is left unused so that a wrong table entry is visible)."""
    hkv = seqs[0][1].shape[1]
    need = [(s[0].shape[0] + page - 1) // page for s in seqs]
    nb = sum(need) + extra_pages + 1
    g = torch.Generator().manual_seed(seed)
    perm = (torch.randperm(nb - 1, generator=g) + 1).tolist()
    cache = torch.zeros(2, nb, page, hkv, D, dtype=torch.bfloat16)
    bt = torch.zeros(len(seqs), max(need) + 2, dtype=torch.int32)
    for b, (_, k, v) in enumerate(seqs):
        n = k.shape[0]
        for j in range(need[b]):
            pg = perm.pop()
            bt[b, j] = pg
            rows = slice(j * page, min(n, (j + 1) * page))
            cache[0, pg, :rows.stop - rows.start] = k[rows]
            cache[1, pg, :rows.stop - rows.start] = v[rows]
    return cache.to(DEV), bt.to(DEV)


CASES = [
    # (seq_lens, q_lens, hq, hkv, si, sl, last)
    ([700, 300, 2500], [200, 300, 1], 8, 2, 4, 32, 64),          # chunk, whole sequence, decode step
    ([4100], [1000], 4, 2, 4, 32, 64),                           # chunk holds the dense tail rows
    ([4100], [130], 4, 1, 4, 32, 200),                           # chunk entirely inside the dense tail
    ([5000, 2304], [1, 1], 8, 2, 4, 32, 64),                     # two decode steps
    ([3333, 129, 1], [777, 128, 1], 4, 2, 0, 0, 1),
    ([3000], [1500], 2, 2, 7, 100, 300),                         # MHA: mma.sync kernel under "auto"
    ([6000], [2000], 16, 2, 20, 150, 700),                       # sink / window beyond the streaming kernel's
]


@pytest.mark.parametrize("kernel", ["auto", "mma", "tcgen05", "pair"])
@pytest.mark.parametrize("mode", [0, 1])
@pytest.mark.parametrize("case", range(len(CASES)))
def test_context_rows_contiguous_kv(ops, case, mode, kernel):
    seq_lens, q_lens, hq, hkv, si, sl, last = CASES[case]
    if kernel == "tcgen05" and (hq // hkv) % 2:
        pytest.skip("tcgen05 kernel pairs two query heads of one KV head")
    if kernel == "pair" and (hq // hkv) % 4:
        pytest.skip("the CTA-pair kernel takes four query heads of a KV head per cluster")
    seqs = _seqs(seq_lens, hq, hkv, seed=case)
    scale = D ** -0.5
    q = torch.cat([s[0][-ql:] for s, ql in zip(seqs, q_lens)]).to(DEV)
    k = torch.cat([s[1] for s in seqs]).to(DEV)
    v = torch.cat([s[2] for s in seqs]).to(DEV)
    out = ops.chunked_prefill_attn(q, torch.tensor(_cu(q_lens), dtype=torch.int32, device=DEV),
                                   torch.tensor(seq_lens, dtype=torch.int32, device=DEV), max(q_lens), scale,
                                   si, sl, last, mode, k=k, v=v,
                                   cu_seqlens_k=torch.tensor(_cu(seq_lens), dtype=torch.int32, device=DEV),
                                   kernel=kernel)
    _compare(out, q_lens, _oracle(seqs, q_lens, scale, si, sl, last, mode), (case, mode, kernel))


@pytest.mark.parametrize("kernel", ["mma", "tcgen05"])
@pytest.mark.parametrize("page", [16, 32, 64, 128, 256])
@pytest.mark.parametrize("mode", [0, 1])
def test_context_rows_paged_cache(ops, mode, page, kernel):
    seq_lens, q_lens, hq, hkv, si, sl, last = [1000, 300, 4500, 77], [200, 300, 1, 76], 8, 2, 4, 32, 64
    seqs = _seqs(seq_lens, hq, hkv, seed=page)
    cache, bt = _paged_cache(seqs, page, seed=page)
    scale = D ** -0.5
    q = torch.cat([s[0][-ql:] for s, ql in zip(seqs, q_lens)]).to(DEV)
    cu_q = torch.tensor(_cu(q_lens), dtype=torch.int32, device=DEV)
    lens = torch.tensor(seq_lens, dtype=torch.int32, device=DEV)
    out, lse = ops.chunked_prefill_attn(q, cu_q, lens, max(q_lens), scale, si, sl, last, mode, key_cache=cache[0],
                                        value_cache=cache[1], block_table=bt, kernel=kernel, return_lse=True)
    _compare(out, q_lens, _oracle(seqs, q_lens, scale, si, sl, last, mode), ("paged", mode, page, kernel))
    # the same kernel over a contiguous copy of the keys gives the same bits (only the addressing differs)
    k = torch.cat([s[1] for s in seqs]).to(DEV)
    v = torch.cat([s[2] for s in seqs]).to(DEV)
    ref, lse2 = ops.chunked_prefill_attn(q, cu_q, lens, max(q_lens), scale, si, sl, last, mode, k=k, v=v,
                                         cu_seqlens_k=torch.tensor(_cu(seq_lens), dtype=torch.int32, device=DEV),
                                         kernel=kernel, return_lse=True)
    assert torch.equal(out, ref) and torch.equal(lse, lse2)


@pytest.mark.parametrize("kernel", ["tcgen05", "auto"])
@pytest.mark.parametrize("layer_idx", [0, 1])
def test_backend_three_chunk_prefill_then_decode(tmx, layer_idx, kernel):
    """TriangleMixImpl.forward with a real paged kv_cache: a 20,000-token prompt in three chunks, then one
    decode step.  Dense layer, same kernel on both paths (kernel="tcgen05"; under "auto" the one-shot prefill runs
    on CTA pairs and the paged chunks on the single-CTA kernel): every chunk is bit-identical to the one-shot
    prefill (chunks start on 128-row tile boundaries, so the tile walk is the same).  Triangle layer: the last chunk (whose N is the
    prompt length) matches the one-shot prefill within tolerance; earlier chunks are checked against the
    oracle with N = the length so far (the mask's N is seq_lens[b], as in the reference generator)."""
    n, hq, hkv, page = 20000, 8, 2, 16
    si, sl, last = 4, 32, 64
    cfg = tmx.TriangleMixConfig(si, sl, last, triangle_layer_indices=[1])
    scale = D ** -0.5
    impl = tmx.TriangleMixImpl(hq, D, scale, num_kv_heads=hkv, trianglemix_config=cfg, layer_idx=layer_idx,
                               num_layers=2, kernel=kernel)
    mode = orc.TRIANGLE if layer_idx == 1 else orc.DENSE_CAUSAL
    (q, k, v), = _seqs([n + 1], hq, hkv, seed=20)
    qd, kd, vd = q.to(DEV).view(n + 1, -1), k.to(DEV).view(n + 1, -1), v.to(DEV).view(n + 1, -1)
    one = torch.zeros(n, hq * D, dtype=torch.bfloat16, device=DEV)
    impl.forward(None, qd[:n], kd[:n], vd[:n], None, tmx.TriangleMixMetadataBuilder.from_seq_lens([n], DEV),
                 output=one)
    nb = (n + 1 + page - 1) // page
    g = torch.Generator().manual_seed(1)
    bt = torch.randperm(nb, generator=g).to(torch.int32).view(1, nb).to(DEV)
    kv_cache = torch.zeros(tmx.TriangleMixBackend.get_kv_cache_shape(nb, page, hkv, D), dtype=torch.bfloat16, device=DEV)
    pos = torch.arange(n + 1, device=DEV)
    slots_all = bt[0].to(torch.int64)[pos // page] * page + pos % page
    chunked = torch.zeros(n + 1, hq * D, dtype=torch.bfloat16, device=DEV)
    bounds = [(0, 6144), (6144, 14080), (14080, n), (n, n + 1)]          # three prefill chunks + a decode step
    for c0, c1 in bounds:
        md = tmx.TriangleMixMetadataBuilder.from_seq_lens([c1], DEV, query_lens=[c1 - c0], block_table=bt,
                                                          slot_mapping=slots_all[c0:c1])
        assert md.context_present() == (c0 > 0)
        impl.forward(None, qd[c0:c1], kd[c0:c1], vd[c0:c1], kv_cache, md, output=chunked[c0:c1])
    torch.cuda.synchronize()
    # the cache holds every K/V row at its slot
    got_k = kv_cache[0].view(-1, hkv * D)[slots_all]
    assert torch.equal(got_k, kd) and torch.equal(kv_cache[1].view(-1, hkv * D)[slots_all], vd)
    if layer_idx == 0 and kernel == "tcgen05":
        assert torch.equal(chunked[:n], one)
    else:
        lo = 0 if layer_idx == 0 else 14080
        rep = orc.error_report(chunked[lo:n].cpu(), one[lo:n].cpu().float(), ATOL, RTOL)
        assert rep["violations"] == 0, rep
    for c0, c1 in bounds:
        seq = [(q[:c1], k[:c1], v[:c1])]
        _compare(chunked[c0:c1].view(-1, hq, D), [c1 - c0], _oracle(seq, [c1 - c0], scale, si, sl, last, mode),
                 ("backend chunk", layer_idx, c0, c1))


def test_backend_context_without_cache_raises(tmx):
    impl = tmx.TriangleMixImpl(4, D, 0.1, num_kv_heads=2)
    md = tmx.TriangleMixMetadataBuilder.from_seq_lens([300], DEV, query_lens=[1])
    x = torch.zeros(1, 4 * D, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(NotImplementedError, match="context keys"):
        impl.forward(None, x, x[:, :2 * D], x[:, :2 * D], None, md, output=torch.empty_like(x))


def test_backend_short_sequences_are_dense_in_triangle_layers(tmx):
    """The reference helper returns no mask at seq_len <= 2048 (trianglemix_attention.py:330-332): such a
    sequence is plain causal even in a triangle layer.  Mixed varlen batch: short, long, short."""
    lens, hq, hkv = [700, 2600, 2048], 4, 2
    si, sl, last = 4, 32, 64
    cfg = tmx.TriangleMixConfig(si, sl, last, triangle_layer_indices=[0])
    impl = tmx.TriangleMixImpl(hq, D, D ** -0.5, num_kv_heads=hkv, trianglemix_config=cfg, layer_idx=0, num_layers=1)
    seqs = _seqs(lens, hq, hkv, seed=31)
    q = torch.cat([s[0] for s in seqs]).to(DEV)
    k = torch.cat([s[1] for s in seqs]).to(DEV)
    v = torch.cat([s[2] for s in seqs]).to(DEV)
    t = sum(lens)
    out = torch.zeros(t, hq * D, dtype=torch.bfloat16, device=DEV)
    impl.forward(None, q.view(t, -1), k.view(t, -1), v.view(t, -1), None,
                 tmx.TriangleMixMetadataBuilder.from_seq_lens(lens, DEV), output=out)
    cu = _cu(lens)
    for b, (n, (qq, kk, vv)) in enumerate(zip(lens, seqs)):
        mode = orc.TRIANGLE if n > 2048 else orc.DENSE_CAUSAL
        ref = orc.attention_varlen_fp32(qq, kk, vv, [0, n], D ** -0.5, si, sl, last, mode)
        rep = orc.error_report(out[cu[b]:cu[b + 1]].cpu().view(n, hq, D), ref, ATOL, RTOL)
        assert rep["violations"] == 0, (b, n, rep)
