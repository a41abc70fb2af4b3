"""HBM bandwidth by access mix (write-only, read-only, copy) with torch library kernels: the denominators that bound
the store-dominated GEMM epilogues and the read-dominated reductions."""
import json
import torch

dev = torch.device("cuda", 0)
n = 1 << 30  # 4 GiB of fp32
x = torch.empty(n, dtype=torch.float32, device=dev)
y = torch.empty(n, dtype=torch.float32, device=dev)


def t(fn, it=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / it


ms = t(lambda: x.zero_())
print(json.dumps({"op": "write only (zero_)", "GBs": round(4 * n / ms / 1e6, 1)}))
ms = t(lambda: x.sum())
print(json.dumps({"op": "read only (sum)", "GBs": round(4 * n / ms / 1e6, 1)}))
ms = t(lambda: y.copy_(x))
print(json.dumps({"op": "copy (read + write)", "GBs": round(8 * n / ms / 1e6, 1)}))
