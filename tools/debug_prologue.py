"""Developer probe for one prologue fuzz seed: where does the GPU differ from the oracle?"""
import os, sys, random
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import qknorm_rope_oracle as orc
from vllm_trianglemix_b200 import ops
seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1385
r = random.Random(seed)
t = r.choice([1, 2, 7, 33, 64, 100, 257, 1000, r.randint(1, 3000)])
hq, hkv = r.choice([(1, 1), (2, 1), (4, 1), (4, 2), (8, 1), (8, 2), (16, 8), (32, 8), (5, 5), (7, 1), (64, 8)])
max_pos = r.choice([16, 4096, 40960]); bs = r.choice([16, 32, 48, 64]); nb = (t + bs - 1) // bs + r.randint(0, 3)
inplace, with_cache, hnd = r.random() < 0.5, r.random() < 0.7, r.random() < 0.3
d = 128
g = torch.Generator().manual_seed(500 + seed)
qkv = torch.randn(t, (hq + 2 * hkv) * d, generator=g); qkv[:, :hq * d] *= 3.0; qkv = qkv.bfloat16()
qw = (1.0 + 0.1 * torch.randn(d, generator=g)).bfloat16(); kw = (1.0 + 0.1 * torch.randn(d, generator=g)).bfloat16()
pos = torch.randint(0, max_pos, (t,), generator=g, dtype=torch.int64)
base = r.choice([10000.0, 1000000.0]); eps = r.choice([1e-6, 1e-5])
cache = orc.cos_sin_cache(d, max_pos, base)
q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
wq, wk = orc.qk_norm_rope(q, k, pos, qw, kw, cache, eps, d)
qd, kd, vd = qkv.cuda().split([hq * d, hkv * d, hkv * d], dim=-1)
for trial in range(3):
    gq, gk = ops.qk_norm_rope(pos.cuda(), qd, kd, qw.cuda(), kw.cuda(), cache.cuda(), eps, d)
    torch.cuda.synchronize()
    bad = (gk.cpu().view(torch.int16) != wk.view(torch.int16)).nonzero()
    print("trial", trial, "t hq hkv", t, hq, hkv, "pos", pos.tolist()[:4], "k mismatches", bad.shape[0], bad[:12].tolist())
    badq = (gq.cpu().view(torch.int16) != wq.view(torch.int16)).nonzero()
    print("   q mismatches", badq.shape[0], badq[:8].tolist())
# hypotheses
wq2, wk_qw = orc.qk_norm_rope(q, k, pos, qw, qw, cache, eps, d)
print("k with q weights matches GPU:", (gk.cpu().view(torch.int16) == wk_qw.view(torch.int16)).float().mean().item())
for i, j in bad[:6].tolist():
    print(i, j, "gpu", float(gk[i, j]), "want", float(wk[i, j]), "kw", float(kw[j % d]), "qw", float(qw[j % d]))
