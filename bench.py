#!/usr/bin/env python
"""bench.py -- RoIs/sec fwd+bwd over (pool + MIL + refine) for the JTSM hot path on B200.

Contract (driver): `python bench.py --gpus N --steps K --warmup W [--impl reference]`;
for N>1 launched under torch.distributed.run (one rank per GPU).  Rank 0 prints ONE JSON line.

A "step" = one pass of the hot path over one image batch of the named workload
(SURVEY.md 8d): pool fwd (x objectness) -> [given head logits] WSDDN MIL fwd + BCE ->
Kr x (pseudo-gt top-k -> IoU label assignment -> weighted CE [+ box reg] -> softmax) ->
backward of all of it (d logits) and RoIPool backward down to grad_feature.
fc6/fc7 GEMMs are excluded (they stay in torch).

  value : RoIs/s, whole job, inputs resident in HBM, CUDA-event timed, L2 flushed
          (256 MiB write) between timed steps.
  e2e   : same step driven from pinned HOST buffers through the public ops layer
          (-> C ABI), H2D of feature map / rois / logits and D2H of losses + grad_feature
          inside the timed region.
  roofline : dominant kernel (RoIPool forward): algorithmic bytes / its CUDA-event time.
  cpu_baseline : the CPU oracle port (oracle/jtsm_oracle.c, host threads) on a bounded
          sample of the same workload.  --impl reference times the reference's own CPU
          path (torchvision CPU roi_pool + the eager-PyTorch head composition).
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

METRIC = "rois_per_sec_fwd_bwd_pool_mil_refine"
UNIT = "RoIs/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c2")
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch the step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--cpu-sample-rois", type=int, default=0, help="0 = auto (about 10-30 s)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            d = json.load(fh)
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []
        self.mark_at = 0

    def mark(self):
        """Samples from here on count as 'during the timed region' (falls back to all if none arrive)."""
        self.mark_at = len(self.lines)

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        lines = self.lines[self.mark_at:] or self.lines
        for l in lines:
            f = [t.strip() for t in l.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------
# workload
# ----------------------------------------------------------------------------------------------
def make_workload(name, rank, dtype):
    import torch

    from jtsm_b200 import synth

    w = synth.step_case(name, sweep_index=rank)
    if dtype == "bf16":
        w["feat"] = w["feat"].bfloat16()
    return w


def algorithmic_bytes(cfg, esz):
    """compulsory traffic only."""
    n, c, h, w_, r, p, k, kr = cfg["N"], cfg["C"], cfg["H"], cfg["W"], cfg["N"] * cfg["R"], cfg["P"], cfg["K"], cfg["Kr"]
    pool_fwd = n * c * h * w_ * esz + 20 * r + r * c * p * p * esz
    pool_bwd = r * c * p * p * esz + n * c * h * w_ * esz + 20 * r
    if cfg.get("pooler") == "MOIPool":  # API-visible int32 argmax, oh_labels [R, S], superpixels [N, Hs, Ws]
        extra = r * c * p * p * 4
        pool_fwd += extra + r * cfg["S"] * 4 + n * cfg["img_h"] * cfg["img_w"] * 4
        pool_bwd += extra
    mil = 7 * r * k * 4
    refine = kr * (r * (16 + 12 + 3 * (k + 1) * 4) + (r * (16 + 2 * 4 * k * 4) if cfg.get("reg") else 0))
    return dict(pool_fwd=pool_fwd, pool_bwd=pool_bwd, mil=mil, refine=refine,
                step=pool_fwd + pool_bwd + mil + refine)


def run_b200(a):
    import torch
    import torch.distributed as dist

    from jtsm_b200 import _lib, hotpath

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (impl=b200) needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_bound = bind_to_gpu_numa_node(local) if world > 1 else False
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()  # fail loudly if libjtsm_b200.so is missing

    sampler = ClockSampler(local)  # started early: nvidia-smi needs ~1 s to deliver its first sample
    if rank == 0:
        sampler.start()
    wl = make_workload(a.config, rank, a.dtype)
    cfg = wl["cfg"]
    R = cfg["N"] * cfg["R"]
    esz = 2 if a.dtype == "bf16" else 4
    step = hotpath.HotPathStep(cfg, dev, dtype=torch.bfloat16 if a.dtype == "bf16" else torch.float32)
    host = step.pin_host_inputs(wl)
    resident = step.upload(host)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---------------- device-resident timing (value) ----------------
    use_graph = not a.no_graph
    if use_graph:
        step.capture(resident)
        run_step = step.replay
    else:
        run_step = lambda: step.run(resident)
    for _ in range(max(a.warmup, 3)):
        run_step()
    barrier()
    lc0 = _lib.launch_count()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    barrier()
    sampler.mark()
    t_wall0 = time.perf_counter()
    for i in range(a.steps):
        flush.zero_()
        ev[i][0].record()
        run_step()
        ev[i][1].record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = step.launches_per_step * a.steps if use_graph else _lib.launch_count() - lc0
    dev_ms = sum(e0.elapsed_time(e1) for e0, e1 in ev)
    # dominant op (RoIPool forward = 2 memsets + prep kernel + pooling kernel): its own CUDA graph, CUDA events
    # around each replay, same inputs, L2 flushed (events cannot be recorded inside the step's captured graph)
    step.capture_dominant(resident)
    kev = []
    for i in range(a.steps):
        flush.zero_()
        kev.append(step.time_dominant())
    barrier()
    dom_ms = sum(e0.elapsed_time(e1) for e0, e1 in kev) / len(kev)

    # ---------------- end-to-end from host buffers (e2e) ----------------
    # Every step uploads its inputs from pinned host memory and downloads its results; with the CUDA-graph step
    # the public streaming call (HotPathStep.run_stream) double-buffers, so step i + 1's H2D and step i - 1's D2H
    # overlap step i's kernels.  Cache state: every step's inputs arrive fresh over PCIe and the two buffer sets
    # alternate, each step streaming > 500 MB (pooled tensor, argmax, upstream gradient) through the 126 MB L2
    # before a buffer is touched again, so nothing a step reads can be warm from the previous one; the same loop
    # with an explicit 256 MiB flush before every step is timed as well and reported next to it.
    e2e_flush_ms = None
    h2d_gbps = None
    io_graph_ms = None
    if use_graph:
        step.prepare_stream(host)
        # the link alone: the step's pinned arena copied 10 times back to back
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dst = step.slots[0]["inputs"]["__arena"]
        dst.copy_(host["__arena"], non_blocking=True)
        c0.record()
        for _ in range(10):
            dst.copy_(host["__arena"], non_blocking=True)
        c1.record()
        torch.cuda.synchronize()
        h2d_gbps = 10 * step.h2d_bytes(host) / (c0.elapsed_time(c1) * 1e-3) / 1e9
        step.run_stream([host] * 3)
        barrier()
        # the streamed step's graph alone (derive inputs + step + result packing), inputs resident, back to back
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(a.steps):
            if step.slots[0]["graph"] is not None:
                step.slots[0]["graph"].replay()
            else:
                step.slots[0]["graph_a"].replay()
                step.slots[0]["graph_b"].replay()
        g1.record()
        barrier()
        io_graph_ms = g0.elapsed_time(g1) / a.steps
        seen = []
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step.run_stream([host] * a.steps, consume=lambda i, out: seen.append(float(out[0][0])))
        e1.record()
        barrier()
        assert len(seen) == a.steps
        e2e_ms = e0.elapsed_time(e1)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        step.run_stream([host] * a.steps, before_step=flush.zero_)
        f1.record()
        barrier()
        e2e_flush_ms = f0.elapsed_time(f1)
        e2e_mode = "pipelined: H2D | step | D2H on three streams, %d buffer sets, uploads %s" % (
            len(step.slots), "held back until the running step's pooling forward has finished (two half-arena copies "
            "per step)" if step.overlap == "after_forward" else "as soon as queued")
    else:
        for _ in range(3):
            step.run_from_host(host)
        barrier()
        ev2 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
        for i in range(a.steps):
            flush.zero_()
            ev2[i][0].record()
            step.run_from_host(host)  # ends with a blocking D2H of the results
            ev2[i][1].record()
        barrier()
        e2e_ms = sum(e0.elapsed_time(e1) for e0, e1 in ev2)
        e2e_mode = "sequential: H2D, step, blocking D2H"

    clocks = sampler.stop() if rank == 0 else None
    tm = torch.tensor([dev_ms, e2e_ms, e2e_flush_ms or 0.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, e2e_flush_ms = (float(v) for v in tm.tolist())

    if rank == 0:
        peak, peak_src = peaks()
        ab = algorithmic_bytes(cfg, esz)
        ach = ab["pool_fwd"] / (dom_ms * 1e-3) / 1e9
        out = {
            "metric": METRIC, "value": world * R * a.steps / (dev_ms * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": dev_ms / a.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32" if a.dtype == "fp32" else "bf16",
            "data": "synthetic (seeded MCG-like proposals, relu(randn) conv5 map, random head logits)",
            "config": {"workload": wl["name"], "images_per_gpu": cfg["N"], "rois_per_image": cfg["R"],
                       "feature": [cfg["N"], cfg["C"], cfg["H"], cfg["W"]], "pooled": cfg["P"], "classes": cfg["K"],
                       "refine_branches": cfg["Kr"], "box_reg": bool(cfg.get("reg")), "pooler": cfg.get("pooler", "ROIPool"),
                       "l2": "flushed (256 MiB write) between timed steps (value, roofline); e2e: see e2e.l2", "stages": step.stages(), "schedule": step.schedule(),
                       "launch": "cuda_graph_replay" if use_graph else "eager"},
            "e2e": {"value": world * R * a.steps / (e2e_ms * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": step.h2d_bytes(host), "d2h_bytes_per_step": step.d2h_bytes(),
                    "mode": e2e_mode,
                    "l2": "fresh H2D inputs every step; alternating buffer sets, > 500 MB streamed per step "
                          "(126 MB L2) before a buffer is reused; no explicit flush" if use_graph else
                          "flushed (256 MiB write) before every step",
                    "h2d_link_gbps": h2d_gbps, "cpu_affinity_gpu_local": numa_bound, "step_graph_ms_inputs_resident_l2_warm": io_graph_ms,
                    "value_with_l2_flush": (world * R * a.steps / (e2e_flush_ms * 1e-3)) if e2e_flush_ms else None},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": step.dominant_kernel(), "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak, "peak_source": peak_src, "traffic": step.ncu_traffic(wl["name"] if a.dtype == "fp32" else None),
                         "algorithmic_bytes_per_launch": ab["pool_fwd"], "avg_launch_ms": dom_ms,
                         "step_algorithmic_bytes": ab["step"],
                         "step_frac": ab["step"] / (dev_ms / a.steps * 1e-3) / 1e9 / peak},
            "clocks": clocks,
            "wall_s_timed_region": t_wall,
        }
        if not a.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(wl, a, kind="port")
        emit(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


# ----------------------------------------------------------------------------------------------
# CPU legs (the only places that execute oracle/)
# ----------------------------------------------------------------------------------------------
def cpu_baseline(wl, a, kind):
    from oracle import cpu_step

    cores = os.cpu_count() or 1
    sample = a.cpu_sample_rois or cpu_step.auto_sample_rois(wl, kind)
    t, n_rois, desc = cpu_step.time_step(wl, sample_rois=sample, threads=cores, kind=kind)
    return {"value": n_rois / t, "unit": UNIT, "cores": cores, "kind": kind, "sample": desc, "seconds": t}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    from oracle import cpu_step

    wl = make_workload(a.config, 0, "fp32")
    cfg = wl["cfg"]
    cores = os.cpu_count() or 1
    sample = a.cpu_sample_rois or cpu_step.auto_sample_rois(wl, "reference", budget_s=6.0)
    for _ in range(max(1, min(a.warmup, 1))):
        cpu_step.time_step(wl, sample_rois=sample, threads=cores, kind="reference")
    ts = []
    desc = ""
    for _ in range(a.steps):
        t, n_rois, desc = cpu_step.time_step(wl, sample_rois=sample, threads=cores, kind="reference")
        ts.append(t)
    v = n_rois * len(ts) / sum(ts)
    out = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * sum(ts) / len(ts), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "images_per_gpu": cfg["N"], "rois_per_image": cfg["R"],
                   "feature": [cfg["N"], cfg["C"], cfg["H"], cfg["W"]], "pooled": cfg["P"], "classes": cfg["K"],
                   "refine_branches": cfg["Kr"], "box_reg": bool(cfg.get("reg")), "pooler": cfg.get("pooler", "ROIPool")},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": desc},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(json.dumps(out))


_JSON_OUT = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything libraries print meanwhile (e.g. NCCL's version banner,
    which lands on fd 1) was redirected to stderr by main()."""
    os.write(_JSON_OUT, (line + "\n").encode())


def bind_to_gpu_numa_node(local):
    """Run this rank (and allocate its pinned host buffers) on the CPUs next to its GPU: with 8 ranks streaming
    16 MB per step each, pinned memory on the far socket costs the H2D copies their bandwidth."""
    try:
        import pynvml

        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
        return True
    except Exception:
        return False


def main():
    global _JSON_OUT
    a = parse()
    sys.stdout.flush()
    _JSON_OUT = os.dup(1)
    os.dup2(2, 1)  # stray prints of libraries -> stderr
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)


if __name__ == "__main__":
    main()
