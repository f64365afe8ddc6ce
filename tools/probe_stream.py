"""Where does the streamed (e2e) step lose time against its own graph?  Variants of HotPathStep.run_stream's loop at c2:
full, without the H2D copy, without the D2H copy, without the per-step host sync.  Prints ms per step."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import hotpath, synth  # noqa: E402

dev = torch.device("cuda:0")
wl = synth.step_case("c2")
step = hotpath.HotPathStep(wl["cfg"], dev)
host = step.pin_host_inputs(wl)
step.prepare_stream(host, depth=2, overlap="any")
N = 200


dummy = torch.empty_like(step.slots[0]['inputs']['__arena'])


def loop(h2d=True, d2h=True, hostsync=True, to_dummy=False, delay_cycles=0):
    comp = torch.cuda.current_stream()
    ns = len(step.slots)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(N):
        sl = step.slots[i % ns]
        if h2d:
            with torch.cuda.stream(step.copy_stream):
                if i >= ns:
                    step.copy_stream.wait_event(sl["ev_done"])
                if delay_cycles:
                    torch.cuda._sleep(delay_cycles)  # start the upload later inside the running step
                (dummy if to_dummy else sl["inputs"]["__arena"]).copy_(host["__arena"], non_blocking=True)
                sl["ev_in"].record(step.copy_stream)
            comp.wait_event(sl["ev_in"])
        if d2h and i >= ns:
            comp.wait_event(sl["ev_out"])
        sl["graph"].replay()
        sl["ev_done"].record(comp)
        if d2h:
            with torch.cuda.stream(step.d2h_stream):
                step.d2h_stream.wait_event(sl["ev_done"])
                sl["out_buf"].copy_(sl["res"], non_blocking=True)
                sl["ev_out"].record(step.d2h_stream)
        if hostsync and i >= 1:
            prev = step.slots[(i - 1) % ns]
            (prev["ev_out"] if d2h else prev["ev_done"]).synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / N


for name, kw in [("full", {}), ("no H2D", dict(h2d=False)), ("no D2H", dict(d2h=False)), ("no host sync", dict(hostsync=False)),
                 ("graph only", dict(h2d=False, d2h=False, hostsync=False)),
                 ("H2D into a dummy buffer (inputs stay warm)", dict(to_dummy=True)),
                 ("H2D delayed 0.10 ms", dict(delay_cycles=196000)), ("H2D delayed 0.20 ms", dict(delay_cycles=393000)),
                 ("H2D delayed 0.28 ms", dict(delay_cycles=550000))]:
    loop(**kw)
    print("%-14s %.4f ms/step" % (name, loop(**kw)))
