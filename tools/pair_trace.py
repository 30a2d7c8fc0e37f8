"""Developer tool: clock64 timeline of the CTA-pair dense kernel (needs the DEV library: make -C
vllm_trianglemix_b200/csrc DEV=1).  Prints, for cluster 0's leader CTA, the mean cycles between the pipeline
events of the two head groups' softmax warps and of the MMA issuer over KV tiles 16..90."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vllm_trianglemix_b200 import _lib

_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), os.environ.get("TMX_DEV_LIB", "libtmx_dev.so"))
from vllm_trianglemix_b200 import ops  # noqa: E402

EV = ["w0 H0 seen", "w1 H0 seen", "w2 H0 seen", "w3 H0 seen", "w0 P ready", "w1 P ready", "w2 P ready", "w3 P ready",
      "H0_A(t+1) iss", "H1_A(t+1) iss", "H0_B(t+1) iss", "H1_B(t+1) iss"]


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
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
    run = lambda: ops.prefill_attn(q, k, v, cu, n, 128 ** -0.5, 0, 0, 0, 0, out=out, kernel="pair")
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    lib.tmx_dev_set_trace(trace.data_ptr())
    run()
    torch.cuda.synchronize()
    lib.tmx_dev_set_trace(None)
    tr = trace.cpu().double()[0]
    t = tr[16:91]
    period = (t[-1, 0] - t[0, 0]) / (t.shape[0] - 1)
    print("tile period (group A, H0 seen -> H0 seen): %.0f clk  => tensor pipe %.1f %% of the 2048-clk MMA time" %
          (period, 100 * 2048 / period))
    for w in range(4):
        print("warp %d (scheduler %d): pass H0 seen -> P ready %5.0f | P ready -> next H0 seen %5.0f | P ready behind the "
              "first warp's %5.0f" % (w, w, (t[:, 4 + w] - t[:, w]).mean(), (t[1:, w] - t[:-1, 4 + w]).mean(),
                                      (t[:, 4 + w] - t[:, 4:8].min(dim=1).values).mean()))
    last = t[:, 4:8].argmax(dim=1)
    print("last warp to arrive at 'P ready': " + ", ".join("warp %d: %d tiles" % (w, int((last == w).sum())) for w in range(4)))
    print("last P ready -> PV_A + H1_A(t+1) issued %5.0f | H0_A(t+1) issued -> first H0 seen(t+1) %5.0f" %
          ((t[:, 9] - t[:, 4:8].max(dim=1).values).mean(), (t[1:, 0:4].min(dim=1).values - t[:-1, 8]).mean()))
    base = tr[40, 0]
    print("raw tiles 40..45 relative to tile 40's 'A H0 seen':")
    print("   " + " ".join("%13s" % e for e in EV))
    for i in range(40, 46):
        print("   " + " ".join("%13.0f" % (x - base) for x in tr[i]))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("layer %.3f ms = %.1f TFLOP/s (dev library)" % (ms, 4 * 128 * hq * (n * (n + 1) / 2) / ms / 1e9))


if __name__ == "__main__":
    main()
