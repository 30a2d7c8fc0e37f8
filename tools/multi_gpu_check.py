"""Multi-GPU parity of the head-sharded path on real GPUs (NCCL): every rank computes its head
shard with the CUDA kernels, the shards are all-gathered, and rank 0 compares the gathered tensor
with the same layer computed on ONE GPU with all heads.  Heads never interact, so dense layers
must be bit-identical.  In triangle layers the dense-tail rows are split over the keys into as many
chunks as fill the SMs (tmx_api.cu: sm_count / (B * Hkv)), so the per-rank head count changes the fp32
merge order of those last-1 rows: they must agree within the attention tolerance (atol 1.5e-2 /
rtol 1e-2) and every other row must still be bit-identical.  BASELINE configs[3] (Qwen3-32B heads, one 128K... here a
shorter prompt by default) and configs[4] (8 ragged sequences).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node P --master-addr 127.0.0.1 \
        --master-port 29511 tools/multi_gpu_check.py [--n 32768]
"""
import argparse
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import vllm_trianglemix_b200 as tmx
from vllm_trianglemix_b200 import head_shard


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", dest="n", type=int, default=32768)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    d, si, sl, last = 128, 4, 32, 64
    results = []
    cases = [("config4_qwen3_32b_heads", 64, 8, [args.n]),
             ("config5_varlen_8seq", 32, 8, [8192, 12288, 16384, 24576, 32768, 40960, 49152, 65536])]
    for name, hq_tot, hkv_tot, seq_lens in cases:
        if name.startswith("config5") and args.n < 32768:
            seq_lens = [x // 8 for x in seq_lens]
        t = sum(seq_lens)
        part = head_shard.head_partition(rank, world, hq_tot, hkv_tot)
        g = torch.Generator().manual_seed(1234)
        q = torch.randn(t, hq_tot, d, generator=g).bfloat16()
        k = torch.randn(t, hkv_tot, d, generator=g).bfloat16()
        v = torch.randn(t, hkv_tot, d, generator=g).bfloat16()
        cfg = tmx.TriangleMixConfig(si, sl, last, triangle_layer_indices=[1])
        meta = tmx.TriangleMixMetadataBuilder.from_seq_lens(seq_lens, dev)
        for layer_idx in (0, 1):
            impl = tmx.TriangleMixImpl(part.num_heads, d, d ** -0.5, num_kv_heads=part.num_kv_heads,
                                       trianglemix_config=cfg, layer_idx=layer_idx, num_layers=2)
            qs = head_shard.shard_heads(q, part, "q").contiguous().to(dev)
            ks = head_shard.shard_heads(k, part, "kv").contiguous().to(dev)
            vs = head_shard.shard_heads(v, part, "kv").contiguous().to(dev)
            out = torch.empty(t, part.num_heads * d, dtype=torch.bfloat16, device=dev)
            impl.forward(None, qs.view(t, -1), ks.view(t, -1), vs.view(t, -1), None, meta, output=out)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            dist.barrier()
            e0.record()
            gathered, _ = head_shard.all_gather_heads(out.view(t, part.num_heads, d))
            e1.record()
            torch.cuda.synchronize()
            gather_ms = e0.elapsed_time(e1)
            full = head_shard.heads_from_gathered(gathered)
            # the same gather by NVLink peer-memory pushes (head_shard.PeerGather): identical bytes, timed
            peer_ms, peer_same = None, None
            try:
                pg = head_shard.PeerGather((t, part.num_heads * d), torch.bfloat16, dev, depth=1)
                for rep in range(3):
                    pg.buffer(rep).copy_(out)
                    torch.cuda.synchronize()
                    dist.barrier()
                    e0.record()
                    pg.launch(rep)
                    pg.finish()
                    e1.record()
                    torch.cuda.synchronize()
                peer_ms = e0.elapsed_time(e1)
                peer_same = bool(torch.equal(pg.result(0).view(-1), gathered.view(-1)))
                del pg
            except Exception as e:      # noqa: BLE001
                peer_same = "unavailable: %s" % (e,)
            if rank == 0:
                one = tmx.TriangleMixImpl(hq_tot, d, d ** -0.5, num_kv_heads=hkv_tot, trianglemix_config=cfg,
                                          layer_idx=layer_idx, num_layers=2)
                ref = torch.empty(t, hq_tot * d, dtype=torch.bfloat16, device=dev)
                one.forward(None, q.to(dev).view(t, -1), k.to(dev).view(t, -1), v.to(dev).view(t, -1), None, meta,
                            output=ref)
                torch.cuda.synchronize()
                got = full.reshape(t, -1)
                same = torch.equal(got, ref)
                diff = (got.float() - ref.float()).abs()
                viol = int((diff > 1.5e-2 + 1e-2 * ref.float().abs()).sum().item())
                rows_diff = (got != ref).any(dim=1)
                # rows outside the dense tail of their sequence must not differ at all
                tail = torch.zeros(t, dtype=torch.bool, device=dev)
                s0 = 0
                for ln in seq_lens:
                    tail[s0 + max(0, ln - last + 1):s0 + ln] = True
                    s0 += ln
                stray = int((rows_diff & ~tail).sum().item())
                ok = same if layer_idx == 0 else (viol == 0 and stray == 0)
                results.append({"case": name, "layer": "triangle" if layer_idx == 1 else "dense", "world": world,
                                "tokens": t, "heads_per_rank": [part.num_heads, part.num_kv_heads],
                                "ok": bool(ok) and peer_same is not False, "bit_identical_to_1gpu": bool(same),
                                "peer_gather_identical_to_nccl": peer_same, "peer_gather_ms": peer_ms,
                                "peer_gather_GBps_per_rank": None if not peer_ms else
                                (world - 1) * t * part.num_heads * d * 2 / peer_ms / 1e6,
                                "max_abs_diff": float(diff.max().item()), "tolerance_violations": viol,
                                "rows_differing": int(rows_diff.sum().item()),
                                "rows_differing_outside_dense_tail": stray, "all_gather_ms": gather_ms,
                                "all_gather_GBps_per_rank": out.numel() * 2 * (world - 1) / gather_ms / 1e6})
                print(json.dumps(results[-1]), flush=True)
                del ref
            del gathered, full, out
    # ---- o_proj with the reduce-scatter fused into the GEMM epilogue (head_shard.RowParallelOProj) against
    # torch.matmul + NCCL: the reference's RowParallelLinear value (linear.py:1197-1359) ----------------------
    try:
        t, hidden, hq_tot = min(args.n, 65536), 4096, 32
        kshard = hq_tot * d // world
        g = torch.Generator().manual_seed(77)
        x_full = torch.randn(t, hq_tot * d, generator=g).bfloat16()
        w_full = (torch.randn(hidden, hq_tot * d, generator=g) * (hq_tot * d) ** -0.5).bfloat16()
        xs = x_full[:, rank * kshard:(rank + 1) * kshard].contiguous().to(dev)
        ws = w_full[:, rank * kshard:(rank + 1) * kshard].contiguous().to(dev)
        oproj = head_shard.RowParallelOProj(ws, t, dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for rep in range(3):
            torch.cuda.synchronize(); dist.barrier()
            e0.record()
            mine = oproj.forward(xs)
            e1.record()
            torch.cuda.synchronize()
        fused_ms = e0.elapsed_time(e1)
        full = oproj.forward(xs, all_gather=True).clone()
        # NCCL formulation: local matmul, reduce-scatter (and all-reduce for the full value)
        for rep in range(3):
            torch.cuda.synchronize(); dist.barrier()
            e0.record()
            part = torch.matmul(xs, ws.t())
            rs = torch.empty(t // world, hidden, dtype=torch.bfloat16, device=dev)
            dist.reduce_scatter_tensor(rs, part)
            e1.record()
            torch.cuda.synchronize()
        nccl_ms = e0.elapsed_time(e1)
        ref = (x_full.float() @ w_full.float().t())
        ref_mine = ref[rank * (t // world):(rank + 1) * (t // world)].to(dev)
        err = (mine.float() - ref_mine).abs()
        okp = bool((err <= 3e-2 + 2e-2 * ref_mine.abs()).all()) and bool(
            ((full.float() - ref.to(dev)).abs() <= 3e-2 + 2e-2 * ref.to(dev).abs()).all())
        flag = torch.tensor([1 if okp else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            results.append({"case": "o_proj_fused_reduce_scatter", "world": world, "tokens": t, "hidden": hidden,
                            "k_per_rank": kshard, "ok": bool(flag.item()), "max_abs_err_vs_fp32": float(err.max().item()),
                            "fused_gemm_reduce_scatter_ms": fused_ms, "matmul_plus_nccl_reduce_scatter_ms": nccl_ms,
                            "nccl_result_max_abs_diff": float((rs.float() - mine.float()).abs().max().item())})
            print(json.dumps(results[-1]), flush=True)
    except Exception as e:      # noqa: BLE001
        if rank == 0:
            results.append({"case": "o_proj_fused_reduce_scatter", "ok": False, "error": "%s: %s" % (type(e).__name__, e)})
            print(json.dumps(results[-1]), flush=True)
    if rank == 0:
        ok = all(r["ok"] for r in results)
        if args.out:
            with open(args.out, "w") as f:
                json.dump(results, f, indent=1)
        print("multi_gpu_check", "OK" if ok else "MISMATCH", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
