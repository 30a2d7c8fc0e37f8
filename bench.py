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

SEQ_LEN = 131072
NUM_LAYERS = 36
HQ, HKV, D = 32, 8, 128
SI, SL, LAST = 4, 32, 64                       # reference defaults (trianglemix_attention.py:16-18)
TRIANGLE_LAYERS = list(range(28, 36))          # for_qwen3("8B") -> 8 triangle layers, placed deep
METRIC = "qwen3_8b_128k_trianglemix_prefill_attn_tokens_per_s"
UNIT = "tokens/s"


def workload_config(n_gpus):
    return {"workload": "BASELINE configs[2]: Qwen3-8B 36-layer prefill attention, 128K tokens, "
                        "28 dense-causal shallow layers + 8 triangle deep layers (si=4, sl=32, last=64)",
            "seq_len": SEQ_LEN, "layers": NUM_LAYERS, "triangle_layers": TRIANGLE_LAYERS,
            "q_heads": HQ, "kv_heads": HKV, "head_dim": D,
            "parallelism": "heads/%d" % n_gpus if n_gpus > 1 else "single",
            "l2_policy": "inputs (1.6 GB q/k/v) larger than the 126 MB L2; no explicit flush"}


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
    sample = ("%d query rows spread over the 128K prompt x all 131072 keys x 32 heads, one triangle layer "
              "(mask rows by the reference's scalar Python loops) + one dense layer (tril rows), fp32; "
              "scaled to 36 layers (8 triangle + 28 dense)" % rows)
    line = {"impl": "reference", "metric": METRIC, "value": tok_s, "unit": UNIT, "n_gpus": args.gpus,
            "steps": len(timed), "warmup": args.warmup, "ms_per_step": SEQ_LEN / tok_s * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args.gpus),
            "cpu_baseline": {"value": tok_s, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": tok_s, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "detail": detail}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--kernel", default="auto", choices=["auto", "mma", "tcgen05"])
    ap.add_argument("--seq-len", type=int, default=SEQ_LEN, help="developer override (invalidates the metric)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
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
    n = args.seq_len
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
    og = head_shard.OverlappedGather((n, hq * D), torch.bfloat16, dev, depth=2) if world > 1 else None
    out = torch.empty((n, hq * D), dtype=torch.bfloat16, device=dev) if world == 1 else None

    cfg = tmx.TriangleMixConfig(SI, SL, LAST, triangle_layer_indices=list(TRIANGLE_LAYERS))
    scale = D ** -0.5
    layers = [tmx.TriangleMixImpl(hq, D, scale, num_kv_heads=hkv, trianglemix_config=cfg, layer_idx=i,
                                  num_layers=NUM_LAYERS, kernel=args.kernel) for i in range(NUM_LAYERS)]
    meta = tmx.TriangleMixMetadataBuilder.from_seq_lens([n], dev)
    launches = [0]

    def step(qd, kd, vd, events=None):
        for i, impl in enumerate(layers):
            o = out if og is None else og.buffer(i)
            if events is not None:
                events[i][0].record()
            impl.forward(None, qd, kd, vd, None, meta, output=o)
            launches[0] += ops.last_launch_count()
            if events is not None:
                events[i][1].record()
            if og is not None:     # output all-gather of this layer's [T, Hq/P, D] shard, asynchronous
                og.launch(i)
        if og is not None:
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
        pairs_dense = n * (n + 1) // 2
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
                "prologue": prologue,
                "roofline": {"bound": "tensor", "kernel": "dense-causal attention layer (per rank)",
                             "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s",
                             "frac": achieved / peaks["tflops"], "traffic": traffic,
                             "peak_source": peaks["source"],
                             "algorithmic_flops_per_launch": flops_dense, "avg_launch_ms": dense_avg}}
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            tok_s, detail = cpu_reference_sample(8, threads)
            line["cpu_baseline"] = {
                "value": tok_s, "unit": UNIT, "cores": threads, "kind": "port",
                "sample": "8 query rows x 131072 keys x 32 heads for one triangle + one dense layer "
                          "(reference's Python mask loops + fp32 eager attention), scaled to 36 layers",
                "detail": detail}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
