"""K6 alone: hc_transform_to_working on a synthetic balanced-looking float matrix; the specialised kernel against the general
one (HC_TRANSFORM_GENERIC=1), bit for bit, with CUDA-event times.  usage: transform_once.py [n] [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hicembler_b200 import ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(5)
S = torch.ones((n, n), dtype=torch.float32, device=dev)
# a Hi-C looking band: big counts next to the diagonal (outside the series range), ones elsewhere, a few zeros rows apart
idx = torch.arange(n, device=dev)
for d in range(1, 6):
    S[idx[:-d], idx[d:]] = float(4000 // d)
S[idx[: n // 7] * 7, idx[: n // 7] * 3 % n] = 37.0
vmax = float(S.max().item())
r = (torch.rand(n, generator=g, device=dev, dtype=torch.float64) + 0.5) * (vmax / n) ** 0.5
c = (torch.rand(n, generator=g, device=dev, dtype=torch.float64) + 0.5) * (vmax / n) ** 0.5
W = torch.empty((n, n), dtype=torch.float64, device=dev)
def run(generic, square):
    if generic: os.environ["HC_TRANSFORM_GENERIC"] = "1"
    else: os.environ.pop("HC_TRANSFORM_GENERIC", None)
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _, fl = ops.transform_to_working(S, ops.KIND_POWER, 0.5, 1.1, r=r, c=c, vmax=vmax, square=square, out=W)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return ts, int(fl[0].item())
for square in (True, False):
    tg, fg = run(True, square)
    ref = W.clone() if n <= 52000 else None
    tn, fn = run(False, square)
    same = bool(torch.equal(W.view(torch.int64), ref.view(torch.int64))) if ref is not None else None
    gb = (2.0 * n * n + 8.0 * n * n) / 1e9
    print("n=%d square=%s generic ms %s  specialised ms %s  identical=%s flags=%d/%d  specialised GB/s %.0f" % (
        n, square, ["%.3f" % t for t in tg], ["%.3f" % t for t in tn], same, fg, fn, gb / (min(tn) * 1e-3)))
    del ref
