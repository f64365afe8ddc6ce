import torch, sys
sys.path.insert(0, '/root/repo')
from jtsm_b200 import ops, synth
wl = synth.step_case("c2"); cfg = wl["cfg"]
dev = torch.device("cuda:0")
x = wl["feat"].to(dev); rois_all = wl["rois"].to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def t(fn, n=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(n):
        flush.zero_(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
    ts.sort(); return ts[len(ts) // 2]
for R in (8, 64, 500, 1000, 2000):
    rois = rois_all[:R].contiguous()
    print(R, "rois fwd us", t(lambda: ops.roi_pool_forward(x, rois, cfg["scale"], (7, 7), argmax_dtype=torch.int16)))
