#!/usr/bin/env python
"""Headline benchmark: Qwen3-8B-shape 128K-token TriangleMix prefill attention.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one synthetic 128K-token prompt: all 36
layers' masked GQA prefill attention (BASELINE.json configs[2]: dense shallow layers +
triangle deep layers), each layer choosing its mask through the unchanged
should_use_triangle() switch and running through the vLLM-shaped backend call
TriangleMixImpl.forward(layer, q, k, v, kv_cache, metadata, output=...).
With N > 1 the query/KV heads are sharded across ranks (strong scaling: the prompt is
fixed) and every layer's output shard is all-gathered over NCCL.

Prints ONE JSON line (rank 0).  `value` = tokens/s with q/k/v resident in HBM;
`e2e` = the same through host buffers (pinned H2D of q/k/v and D2H of the last
layer's output inside the timed region).  `--impl reference` times the CPU port of the
reference's own algorithm (oracle/) on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SI, SL, LAST = 4, 32, 64                       # reference defaults (trianglemix_attention.py:16-18)
D = 128
UNIT = "tokens/s"
# BASELINE.json configs (0-based).  2 is the metric's configuration (the driver's line); 3 and 4 are the
# multi-GPU configurations, run with --config 3|4 (profiles/r02_scale_config{3,4}.json).
CONFIGS = {
    2: dict(name="BASELINE configs[2]: Qwen3-8B 36-layer prefill attention, 128K tokens, 28 dense-causal shallow "
                 "layers + 8 triangle deep layers (si=4, sl=32, last=64)",
            metric="qwen3_8b_128k_trianglemix_prefill_attn_tokens_per_s",
            seq_lens=[131072], layers=36, triangle=list(range(28, 36)), hq=32, hkv=8),   # for_qwen3("8B") -> 8
    3: dict(name="BASELINE configs[3]: Qwen3-32B (64 q / 8 kv heads) 64-layer prefill attention, 128K tokens, "
                 "52 dense-causal shallow layers + 12 triangle deep layers (for_qwen3('32B')), head-sharded",
            metric="qwen3_32b_128k_trianglemix_prefill_attn_tokens_per_s",
            seq_lens=[131072], layers=64, triangle=list(range(52, 64)), hq=64, hkv=8),
    4: dict(name="BASELINE configs[4]: batched varlen prefill, 8 sequences 8K-64K (249,856 tokens), Qwen3-8B shape, "
                 "36 layers (28 dense-causal + 8 triangle deep), head-sharded",
            metric="qwen3_8b_varlen8_trianglemix_prefill_attn_tokens_per_s",
            seq_lens=[8192, 12288, 16384, 24576, 32768, 40960, 49152, 65536], layers=36,
            triangle=list(range(28, 36)), hq=32, hkv=8),
}
CFG = dict(CONFIGS[2])
SEQ_LEN, NUM_LAYERS, HQ, HKV = sum(CFG["seq_lens"]), CFG["layers"], CFG["hq"], CFG["hkv"]
TRIANGLE_LAYERS = CFG["triangle"]
METRIC = CFG["metric"]


def select_config(idx, seq_len=None):
    global CFG, SEQ_LEN, NUM_LAYERS, HQ, HKV, TRIANGLE_LAYERS, METRIC
    CFG = dict(CONFIGS[idx])
    if seq_len is not None and len(CFG["seq_lens"]) == 1:
        CFG["seq_lens"] = [seq_len]
    SEQ_LEN, NUM_LAYERS, HQ, HKV = sum(CFG["seq_lens"]), CFG["layers"], CFG["hq"], CFG["hkv"]
    TRIANGLE_LAYERS, METRIC = CFG["triangle"], CFG["metric"]


def workload_config(n_gpus):
    qkv_gb = 2.0 * SEQ_LEN * D * (HQ + 2 * HKV) / 1e9
    return {"workload": CFG["name"], "seq_len": SEQ_LEN, "seq_lens": CFG["seq_lens"], "layers": NUM_LAYERS,
            "triangle_layers": TRIANGLE_LAYERS, "q_heads": HQ, "kv_heads": HKV, "head_dim": D,
            "parallelism": "heads/%d" % n_gpus if n_gpus > 1 else "single",
            "l2_policy": "inputs (%.1f GB q/k/v over all ranks) larger than the 126 MB L2; no explicit flush" % qkv_gb}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": p["hbm_gbs"], "tflops": p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                "source": "measured (MEASURED_PEAKS.json, sustained bf16)"}
    return {"hbm_gbs": 6650.0, "tflops": 1590.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# --------------------------------------------------------------------------- CPU arm
def cpu_reference_sample(rows_per_layer_type, threads):
    """Time the oracle port of the reference's algorithm on a bounded sample of the 128K
    workload: for `rows` query rows spread over the prompt, (a) the mask rows by the
    reference's scalar Python loops (triangle layers; dense layers use tril, vectorised),
    (b) the fp32 scores/softmax/PV of those rows against ALL 131072 keys for all 32 heads,
    exactly as the reference's dense eager op would.  Per-row cost is position-independent
    in the reference (no tile skipping), so tokens/s = 1 / sum_over_layers(seconds per row)."""
    import numpy as np
    import torch
    from oracle import trianglemix_oracle as orc
    torch.set_num_threads(threads)
    n = SEQ_LEN
    g = torch.Generator().manual_seed(0)
    q = torch.randn(rows_per_layer_type, HQ, D, generator=g).bfloat16()
    k = torch.randn(n, HKV, D, generator=g).bfloat16()
    v = torch.randn(n, HKV, D, generator=g).bfloat16()
    rows = [int(x) for x in np.linspace(0, n - 1, rows_per_layer_type)]
    scale = D ** -0.5
    kf = k.float().repeat_interleave(HQ // HKV, dim=1).transpose(0, 1).contiguous()
    vf = v.float().repeat_interleave(HQ // HKV, dim=1).transpose(0, 1).contiguous()
    out = {}
    for name, mode in (("triangle", orc.TRIANGLE), ("dense", orc.DENSE_CAUSAL)):
        t0 = time.perf_counter()
        if mode == orc.TRIANGLE:
            m01 = orc.triangle01_loops(n, SI, SL, LAST, rows)          # reference :82-89,126-129
        else:
            m01 = np.stack([(np.arange(n) <= r) for r in rows]).astype(np.float32)   # tril rows
        add = torch.from_numpy(orc.additive(m01))
        t1 = time.perf_counter()
        qf = q.float().transpose(0, 1)
        scores = torch.matmul(qf, kf.transpose(-2, -1)) * scale + add
        w = torch.softmax(scores, dim=-1)
        o = torch.matmul(w, vf)
        t2 = time.perf_counter()
        out[name] = {"mask_s_per_row": (t1 - t0) / len(rows), "attn_s_per_row": (t2 - t1) / len(rows),
                     "checksum": float(o.sum())}
    n_tri = len(TRIANGLE_LAYERS)
    per_tok = (n_tri * (out["triangle"]["mask_s_per_row"] + out["triangle"]["attn_s_per_row"]) +
               (NUM_LAYERS - n_tri) * (out["dense"]["mask_s_per_row"] + out["dense"]["attn_s_per_row"]))
    return 1.0 / per_tok, out


REF_DIR = os.path.join(ROOT, "baseline", "_ref")     # `pip install --target baseline/_ref` of the unmodified reference


def reference_measured(threads, sizes=(2048, 4096)):
    """The UNMODIFIED reference, loaded by path from baseline/_ref (the offline pip install of /root/reference;
    nothing here reads /root/reference): BASELINE configs[0] -- Qwen3-0.6B attention shape (16 q / 8 kv heads
    repeated to 16, head_dim 128), batch 1, seq 2048 -- mask generation by
    TriangleMixMaskGenerator.get_attention_mask(use_triangle=True) (vllm/attention/trianglemix_attention.py:221-261,
    its own Python loops) + AscendTriangleMixAttention.forward with that mask (vllm/attention/ascend_trianglemix.py:
    52-108), fp32, all host threads; and the same at 4096 tokens.  Real loops, real op, measured end to end.
    Returns None when the install is absent."""
    base = os.path.join(REF_DIR, "vllm", "attention")
    if not os.path.exists(os.path.join(base, "trianglemix_attention.py")):
        return None
    import importlib.util
    import types
    import torch
    saved = {k: sys.modules.get(k) for k in ("vllm", "vllm.attention")}
    try:
        for name in ("vllm", "vllm.attention"):
            m = types.ModuleType(name)
            m.__path__ = []
            sys.modules[name] = m
        mods = {}
        for short in ("trianglemix_attention", "ascend_trianglemix"):
            full = "vllm.attention." + short
            spec = importlib.util.spec_from_file_location(full, os.path.join(base, short + ".py"))
            mod = importlib.util.module_from_spec(spec)
            sys.modules[full] = mod
            spec.loader.exec_module(mod)
            mods[short] = mod
        tma, asc = mods["trianglemix_attention"], mods["ascend_trianglemix"]
        torch.set_num_threads(threads)
        out = []
        for n in sizes:
            g = torch.Generator().manual_seed(0)
            hq, hkv = 16, 8
            q = torch.randn(1, n, hq, D, generator=g).bfloat16().float()
            k = torch.randn(1, n, hkv, D, generator=g).bfloat16().float().repeat_interleave(hq // hkv, dim=2)
            v = torch.randn(1, n, hkv, D, generator=g).bfloat16().float().repeat_interleave(hq // hkv, dim=2)
            cfg = tma.TriangleMixConfig(num_sink_tokens=SI, sliding_window_size=SL, num_last_tokens=LAST)
            gen = tma.TriangleMixMaskGenerator(cfg)
            t0 = time.perf_counter()
            mask = gen.get_attention_mask(n, 0, torch.device("cpu"), torch.float32, use_triangle=True)
            t1 = time.perf_counter()
            op = asc.AscendTriangleMixAttention(num_heads=hq, head_dim=D, scale=D ** -0.5)
            o = op.forward(q, k, v, attn_mask=mask)
            t2 = time.perf_counter()
            out.append({"seq_len": n, "mask_s": t1 - t0, "attention_s": t2 - t1, "layer_s": t2 - t0,
                        "tokens_per_s_one_layer": n / (t2 - t0), "checksum": float(o.double().sum())})
        return {"kind": "reference", "source": "baseline/_ref (pip install --target of the unmodified reference)",
                "cores": threads, "shape": "B1, 16 q heads (8 kv heads repeated), head_dim 128, fp32, triangle mask "
                "si=4 sl=32 last=64; ONE layer: get_attention_mask + AscendTriangleMixAttention.forward",
                "points": out}
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
        for short in ("trianglemix_attention", "ascend_trianglemix"):
            sys.modules.pop("vllm.attention." + short, None)


def run_reference_arm(args, rank):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    rows = 8
    vals = []
    for _ in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        tok_s, detail = cpu_reference_sample(rows, threads)
        vals.append((tok_s, time.perf_counter() - t0))
        if sum(x[1] for x in vals) > 240:      # keep the whole arm within a few minutes
            break
    timed = vals[args.warmup:] or vals[-1:]
    tok_s = sum(x[0] for x in timed) / len(timed)
    measured = reference_measured(threads)
    sample = ("%d query rows spread over the 128K prompt x all 131072 keys x 32 heads, one triangle layer "
              "(mask rows by the reference's scalar Python loops) + one dense layer (tril rows), fp32; "
              "scaled to 36 layers (8 triangle + 28 dense)" % rows)
    line = {"impl": "reference", "metric": METRIC, "value": tok_s, "unit": UNIT, "n_gpus": args.gpus,
            "steps": len(timed), "warmup": args.warmup, "ms_per_step": SEQ_LEN / tok_s * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args.gpus),
            "cpu_baseline": {"value": tok_s, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                             "measured": measured},
            "e2e": {"value": tok_s, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "detail": detail}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kernel", default="auto", choices=["auto", "mma", "tcgen05", "pair"])
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS),
                    help="BASELINE.json configs index: 2 = the metric (default), 3 = Qwen3-32B heads, 4 = 8-sequence varlen")
    ap.add_argument("--seq-len", type=int, default=None, help="developer override (invalidates the metric)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="output all-gather at N > 1: NVLink peer-memory push (default; falls back to NCCL) or ncclAllGather")
    args = ap.parse_args()
    select_config(args.config, args.seq_len)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    import torch
    import torch.distributed as dist
    import vllm_trianglemix_b200 as tmx
    from vllm_trianglemix_b200 import ops

    assert torch.cuda.is_available(), "bench.py needs a B200; there is no CPU path"
    assert world == args.gpus, "launch with torchrun --nproc-per-node %d (WORLD_SIZE=%d)" % (args.gpus, world)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's "NCCL version ..." banner (NCCL_DEBUG=VERSION) goes to stderr
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    n = SEQ_LEN
    assert HKV % world == 0, "head sharding needs kv_heads % n_gpus == 0"
    hq, hkv = HQ // world, HKV // world

    # synthetic global tensors from one seed, sliced by head (rank r owns kv heads [r*hkv, (r+1)*hkv))
    g = torch.Generator().manual_seed(0)
    chunk = 8192
    q_host = torch.empty((n, hq, D), dtype=torch.bfloat16).pin_memory()
    k_host = torch.empty((n, hkv, D), dtype=torch.bfloat16).pin_memory()
    v_host = torch.empty((n, hkv, D), dtype=torch.bfloat16).pin_memory()
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        q_host[s:e] = torch.randn(e - s, HQ, D, generator=g).bfloat16()[:, rank * hq:(rank + 1) * hq]
        k_host[s:e] = torch.randn(e - s, HKV, D, generator=g).bfloat16()[:, rank * hkv:(rank + 1) * hkv]
        v_host[s:e] = torch.randn(e - s, HKV, D, generator=g).bfloat16()[:, rank * hkv:(rank + 1) * hkv]
    out_host = torch.empty((n, hq * D), dtype=torch.bfloat16).pin_memory()
    q = q_host.to(dev).view(n, hq * D)
    k = k_host.to(dev).view(n, hkv * D)
    v = v_host.to(dev).view(n, hkv * D)
    # two output / gather buffers: the all-gather of layer L runs on NCCL's stream while layer L+1
    # computes into the other buffer ("overlap with the next layer's attention")
    from vllm_trianglemix_b200 import head_shard
    og, gather_kind = head_shard.make_gather(args.gather, (n, hq * D), torch.bfloat16, dev, depth=2) \
        if world > 1 else (None, None)
    out = torch.empty((n, hq * D), dtype=torch.bfloat16, device=dev) if world == 1 else None

    cfg = tmx.TriangleMixConfig(SI, SL, LAST, triangle_layer_indices=list(TRIANGLE_LAYERS))
    scale = D ** -0.5
    layers = [tmx.TriangleMixImpl(hq, D, scale, num_kv_heads=hkv, trianglemix_config=cfg, layer_idx=i,
                                  num_layers=NUM_LAYERS, kernel=args.kernel) for i in range(NUM_LAYERS)]
    meta = tmx.TriangleMixMetadataBuilder.from_seq_lens(CFG["seq_lens"], dev)
    launches = [0]

    def step(qd, kd, vd, events=None, gather=True):
        for i, impl in enumerate(layers):
            o = out if og is None else og.buffer(i)
            if events is not None:
                events[i][0].record()
            impl.forward(None, qd, kd, vd, None, meta, output=o)
            launches[0] += ops.last_launch_count()
            if events is not None:
                events[i][1].record()
            if og is not None and gather:     # output all-gather of this layer's [T, Hq/P, D] shard, asynchronous
                og.launch(i)
        if og is not None and gather:
            og.finish()
        return o

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(q, k, v)
    barrier()
    ev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in layers]
          for _ in range(args.steps)]
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches[0] = 0
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for s in range(args.steps):
        step(q, k, v, ev[s])
    t1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms.item()) / args.steps
    n_launch = launches[0]

    # per-layer-type kernel time (dominant kernel = the dense-causal attention launch)
    dense_ms = [ev[s][i][0].elapsed_time(ev[s][i][1]) for s in range(args.steps) for i in range(NUM_LAYERS)
                if i not in TRIANGLE_LAYERS]
    tri_ms = [ev[s][i][0].elapsed_time(ev[s][i][1]) for s in range(args.steps) for i in TRIANGLE_LAYERS]
    dense_avg, tri_avg = sum(dense_ms) / len(dense_ms), sum(tri_ms) / len(tri_ms)

    # ---- the same step without the output all-gather (report both) -----------------------
    no_gather_ms = None
    if world > 1:
        barrier()
        t0.record()
        for _ in range(args.steps):
            step(q, k, v, gather=False)
        t1.record()
        barrier()
        ms3 = torch.tensor([t0.elapsed_time(t1)], device=dev)
        dist.all_reduce(ms3, op=dist.ReduceOp.MAX)
        no_gather_ms = float(ms3.item()) / args.steps

    # ---- end to end through host buffers ------------------------------------------------
    def e2e_step():
        qd = q_host.to(dev, non_blocking=True).view(n, hq * D)
        kd = k_host.to(dev, non_blocking=True).view(n, hkv * D)
        vd = v_host.to(dev, non_blocking=True).view(n, hkv * D)
        out_host.copy_(step(qd, kd, vd), non_blocking=True)
    e2e_step()
    barrier()
    t0.record()
    for _ in range(args.steps):
        e2e_step()
    t1.record()
    barrier()
    ms2 = torch.tensor([t0.elapsed_time(t1)], device=dev)
    if world > 1:
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    e2e_ms = float(ms2.item()) / args.steps

    # ---- the step in front of attention, reported beside the metric --------
    # fused q/k RMSNorm + RoPE + KV-cache write, in place on a packed [T, (Hq+2Hkv)*D] qkv buffer of
    # this rank's heads; 1.6 GB/P of inputs, larger than L2 at every P the bench runs.
    prologue = None
    if rank == 0:
        qkv = torch.randn(n, (hq + 2 * hkv) * D, device=dev, dtype=torch.bfloat16)
        pq, pk, pv = qkv.split([hq * D, hkv * D, hkv * D], dim=-1)
        w1 = torch.ones(D, device=dev, dtype=torch.bfloat16)
        cs_cache = tmx.compute_cos_sin_cache(D, n, 1000000.0, device=dev)
        pos = torch.arange(n, device=dev)
        bs = 16
        kvc = torch.zeros(2, (n + bs - 1) // bs, bs, hkv, D, device=dev, dtype=torch.bfloat16)
        slots = torch.arange(n, device=dev)

        def pro():
            ops.qk_norm_rope(pos, pq, pk, w1, w1, cs_cache, 1e-6, D, inplace=True, v=pv, key_cache=kvc[0],
                             value_cache=kvc[1], slot_mapping=slots)
        for _ in range(3):
            pro()
        torch.cuda.synchronize()
        t0.record()
        for _ in range(10):
            pro()
        t1.record()
        torch.cuda.synchronize()
        pms = t0.elapsed_time(t1) / 10
        pbytes = 2 * n * (hq + hkv) * D * 2 + 3 * n * hkv * D * 2 + n * D * 2 + 2 * n * 8
        prologue = {"kernel": "fused q/k RMSNorm + NeoX RoPE + KV-cache write (tmx_qk_norm_rope)", "ms": pms,
                    "algorithmic_bytes": pbytes, "achieved": pbytes / pms / 1e6, "unit": "GB/s", "bound": "hbm"}
        del qkv, kvc

    if rank == 0:
        peaks = load_peaks()
        if prologue is not None and peaks.get("hbm_gbs"):
            prologue["peak"] = peaks["hbm_gbs"]
            prologue["frac"] = prologue["achieved"] / peaks["hbm_gbs"]
        pairs_dense = sum(x * (x + 1) // 2 for x in CFG["seq_lens"])
        flops_dense = 4.0 * D * hq * pairs_dense
        achieved = flops_dense / (dense_avg * 1e-3) / 1e12
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                traffic = json.load(f).get("dense_attn_dram_bytes_per_launch")
        line = {"metric": METRIC, "value": n / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic", "config": workload_config(world),
                "ms_per_layer": {"dense": dense_avg, "triangle": tri_avg},
                "clocks": clocks,
                "e2e": {"value": n / (e2e_ms * 1e-3), "unit": UNIT,
                        "h2d_bytes_per_step": (q_host.numel() + k_host.numel() + v_host.numel()) * 2,
                        "d2h_bytes_per_step": out_host.numel() * 2, "ms_per_step": e2e_ms},
                "gpu_launches": n_launch,
                "all_gather": None if world == 1 else {
                    "ms_per_step_with": ms_per_step, "ms_per_step_without": no_gather_ms,
                    "exposed_ms_per_step": ms_per_step - no_gather_ms,
                    "bytes_per_rank_per_layer": n * hq * D * 2,
                    "kind": gather_kind,
                    "how": ("every rank pushes each layer's [T, Hq/P, D] output shard into the peers' symmetric-memory "
                            "gather buffers with copy-engine peer copies on a side stream + device barrier "
                            "(head_shard.PeerGather)") if gather_kind == "peer" else
                           ("NCCL all_gather_into_tensor of each layer's [T, Hq/P, D] output shard, asynchronous, "
                            "double-buffered (head_shard.OverlappedGather)")},
                "prologue": prologue,
                "roofline": {"bound": "tensor", "kernel": "dense-causal attention layer (per rank)",
                             "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s",
                             "frac": achieved / peaks["tflops"], "traffic": traffic,
                             "peak_source": peaks["source"],
                             "algorithmic_flops_per_launch": flops_dense, "avg_launch_ms": dense_avg}}
        if world == 1 and not args.no_cpu_baseline and args.config == 2:
            threads = os.cpu_count() or 1
            tok_s, detail = cpu_reference_sample(8, threads)
            line["cpu_baseline"] = {
                "measured": reference_measured(threads),
                "value": tok_s, "unit": UNIT, "cores": threads, "kind": "port",
                "sample": "8 query rows x 131072 keys x 32 heads for one triangle + one dense layer "
                          "(reference's Python mask loops + fp32 eager attention), scaled to 36 layers",
                "detail": detail}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
