"""Developer tool: clock64 timeline of the streaming triangle kernel (needs the DEV library:
make -C vllm_trianglemix_b200/csrc DEV=1).  Prints mean cycle counts between pipeline events."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_trianglemix_b200 import _lib

_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libtmx_dev.so")
from vllm_trianglemix_b200 import ops  # noqa: E402

EV = ["load_issue", "full_seen", "S_issued", "sfull_seen", "P_stored", "PV_issued", "ofull_seen", "O_staged",
      "staged_seen", "stage_freed"]


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
    hq, hkv = 32, 8
    dev = "cuda:0"
    lib = _lib.load()
    lib.tmx_dev_set_trace.argtypes = [ctypes.c_void_p]
    lib.tmx_dev_set_trace.restype = None
    q = torch.randn(n, hq, 128, device=dev, dtype=torch.bfloat16)
    k = torch.randn(n, hkv, 128, device=dev, dtype=torch.bfloat16)
    v = torch.randn(n, hkv, 128, device=dev, dtype=torch.bfloat16)
    out = torch.empty_like(q)
    cu = torch.tensor([0, n], dtype=torch.int32, device=dev)
    trace = torch.zeros(4, 96, 12, dtype=torch.int64, device=dev)
    run = lambda: ops.prefill_attn(q, k, v, cu, n, 128 ** -0.5, 4, 32, 64, 1, out=out)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    lib.tmx_dev_set_trace(trace.data_ptr())
    run()
    torch.cuda.synchronize()
    lib.tmx_dev_set_trace(None)
    tr = trace.cpu().double()
    for cta in range(2):
        t = tr[cta, 12:90, :10]
        print("CTA %d (tiles 12..89), mean cycles:" % cta)
        print("  tile period            %8.0f" % ((t[-1, 0] - t[0, 0]) / (t.shape[0] - 1)))
        print("  stage cycle (3 tiles)  %8.0f" % (tr[cta, 15:90, 0] - tr[cta, 12:87, 0]).mean())
        for a in range(9):
            print("  %-12s -> %-12s %8.0f  (min %6.0f max %6.0f)" % (EV[a], EV[a + 1], (t[:, a + 1] - t[:, a]).mean(),
                                                                    (t[:, a + 1] - t[:, a]).min(), (t[:, a + 1] - t[:, a]).max()))
        print("  load_issue -> stage_freed %8.0f" % (t[:, 9] - t[:, 0]).mean())
    t = tr[0, 40:46, :10]
    print("raw CTA 0 tiles 40..45 (relative to tile 40 load issue):")
    for i in range(t.shape[0]):
        print("  ", " ".join("%7.0f" % (x - t[0, 0]) for x in t[i]))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        run()
    e1.record()
    torch.cuda.synchronize()
    print("layer %.3f ms (dev library)" % (e0.elapsed_time(e1) / 10))


if __name__ == "__main__":
    main()
