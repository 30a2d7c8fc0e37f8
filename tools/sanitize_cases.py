"""Small shapes of every kernel family, for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool racecheck python tools/sanitize_cases.py
Each case also checks the result against the CPU oracle, so a sanitizer-clean run is a correct run.
Shapes are tiny because the sanitizer slows kernels down by 10-100x."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import vllm_trianglemix_b200 as tmx
from vllm_trianglemix_b200 import ops
from oracle import trianglemix_oracle as orc

DEV = "cuda:0"


def rand(t, hq, hkv, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(t, hq, 128, generator=g).bfloat16(), torch.randn(t, hkv, 128, generator=g).bfloat16(),
            torch.randn(t, hkv, 128, generator=g).bfloat16())


def check(out, ref, what):
    rep = orc.error_report(out, ref)
    assert rep["violations"] == 0, (what, rep)
    print("ok", what, "max_abs %.2e" % rep["max_abs"], flush=True)


def main():
    scale = 128 ** -0.5
    # varlen batch: boxes of one sequence run into the next one / past the end of the buffer
    lens = [300, 1, 129, 700]
    cu = [0]
    for n in lens:
        cu.append(cu[-1] + n)
    cud = torch.tensor(cu, dtype=torch.int32, device=DEV)
    for hq, hkv in ((8, 2), (4, 2), (3, 1)):
        q, k, v = rand(cu[-1], hq, hkv, hq)
        for mode, (si, sl, last) in ((ops.TRIANGLE, (4, 32, 64)), (ops.DENSE_CAUSAL, (0, 0, 0)), (ops.TRIANGLE, (20, 150, 40)),
                                     (ops.NONE, (0, 0, 0))):
            for kernel in ("auto", "mma"):
                out = ops.prefill_attn(q.to(DEV), k.to(DEV), v.to(DEV), cud, max(lens), scale, si, sl, last, mode,
                                       kernel=kernel)
                if mode == ops.NONE:
                    ref = torch.cat([orc.attention_fp32(q[a:b][None], k[a:b][None], v[a:b][None], None, scale)[0]
                                     for a, b in zip(cu[:-1], cu[1:])])
                else:
                    ref = orc.attention_varlen_fp32(q, k, v, cu, scale, si, sl, last, mode)
                check(out.cpu(), ref, ("attn", hq, hkv, mode, si, sl, last, kernel))
    # context rows over a paged cache (one TMA box per 16-row page)
    n, ql, hq, hkv, page = 900, 200, 4, 2, 16
    q, k, v = rand(n, hq, hkv, 77)
    nb = (n + page - 1) // page
    bt = torch.randperm(nb).to(torch.int32).view(1, nb)
    cache = torch.zeros(2, nb, page, hkv, 128, dtype=torch.bfloat16)
    for j in range(nb):
        rows = slice(j * page, min(n, (j + 1) * page))
        cache[0, bt[0, j], :rows.stop - rows.start] = k[rows]
        cache[1, bt[0, j], :rows.stop - rows.start] = v[rows]
    cache = cache.to(DEV)
    for kernel in ("tcgen05", "mma"):
        for mode in (ops.TRIANGLE, ops.DENSE_CAUSAL):
            out = ops.chunked_prefill_attn(q[-ql:].to(DEV), torch.tensor([0, ql], dtype=torch.int32, device=DEV),
                                           torch.tensor([n], dtype=torch.int32, device=DEV), ql, scale, 4, 32, 64, mode,
                                           key_cache=cache[0], value_cache=cache[1], block_table=bt.to(DEV), kernel=kernel)
            check(out.cpu(), orc.attention_rows_fp32(q, k, v, scale, n, 4, 32, 64, mode, n - ql, n), ("paged", kernel, mode))
    # masks
    bits = ops.mask_bits(333, 4, 32, 64, "triangle", DEV)
    add = ops.mask_additive(333, 4, 32, 64, ops.TRIANGLE, DEV, torch.bfloat16)
    b2, bad = ops.bits_from_additive(add)
    assert torch.equal(bits, b2) and int(bad.item()) == 0
    # prologue (norm + rope + cache write)
    t, hq, hkv = 100, 4, 2
    qkv = torch.randn(t, (hq + 2 * hkv) * 128).bfloat16().to(DEV)
    blk = tmx.Qwen3TriangleMixAttention(hq, hkv, 128, max_position=256, device=DEV)
    md = tmx.TriangleMixMetadataBuilder.from_seq_lens([t], DEV)
    md.slot_mapping = torch.arange(t, dtype=torch.int64, device=DEV)
    kvc = torch.zeros(tmx.TriangleMixBackend.get_kv_cache_shape(8, 16, hkv, 128), dtype=torch.bfloat16, device=DEV)
    blk(torch.arange(t, device=DEV), qkv, kvc, md)
    torch.cuda.synchronize()
    print("cases ok")


if __name__ == "__main__":
    main()
