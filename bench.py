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
          (-> C ABI); inside the timed region every step uploads its inputs (feature map in the
          step's dtype, rois, objectness, head logits, labels: one pinned arena) and downloads
          what the host consumes (losses, pseudo-labels, counters and a CHECKSUM of grad_feature:
          the gradient itself stays on the device, where the backbone's backward reads it).  The
          timed loop rotates over distinct host batches.
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
# what the reference arm runs: the pooling op is the reference's own dependency (torchvision CPU roi_pool, or the
# reference's MOIPool kernels built as host code); the head math is oracle/ref_heads.py, a restatement pinned
# bit-equal to the reference's Python (tests/test_ref_parity.py) -- a port, and labelled as one
REF_KIND = "reference(pool)+port(heads)"
SINGLE_PROCESS = "; ONE CPU process on rank 0 whatever --gpus says (at N > 1 the driver's ratio divides N GPUs by it)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c2")
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "bf16"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the DDP training-step leg (train_images_per_sec)")
    ap.add_argument("--no-graph", action="store_true", help="launch the step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--cpu-sample-rois", type=int, default=0, help="0 = auto (about 10-30 s)")
    ap.add_argument("--tune", nargs="*", default=[], help="library tuning knobs key=value (A/B experiments only)")
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


def config_dict(wl):
    """The workload description both arms print verbatim (the driver compares them)."""
    cfg = wl["cfg"]
    d = {"workload": wl["name"], "images_per_gpu": cfg["N"], "rois_per_image": cfg["R"], "pooled": cfg["P"],
         "classes": cfg["K"], "refine_branches": cfg["Kr"], "box_reg": bool(cfg.get("reg")),
         "pooler": cfg.get("pooler", "ROIPool")}
    if "augs" in wl:  # inference chain (c4)
        d.update(mode="inference", augmentations=cfg["A"], feature_channels=cfg["C"], stride=cfg["stride"],
                 min_sizes=list(cfg["min_sizes"]), flip=bool(cfg["flip"]), score_thresh=cfg["score_thresh"],
                 nms_thresh=cfg["nms_thresh"], detections_per_image=cfg["topk"])
    else:
        d.update(mode="train_step", feature=[cfg["N"], cfg["C"], cfg["H"], cfg["W"]])
    return d


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
    for kv in a.tune:  # A/B experiments only; the default run sets nothing
        k, v = kv.split("=")
        _lib.set_tuning(k, int(v))

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (impl=b200) needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_bound = bind_to_gpu_numa_node(local) if world > 1 else False
    if world > 1:
        # the training leg captures a DDP step into a CUDA graph: PyTorch asks for NCCL's async error handling to be off
        os.environ.setdefault("TORCH_NCCL_ASYNC_ERROR_HANDLING", "0")
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
    # a second, different batch of the same shapes: the e2e loop alternates between the two pinned arenas
    host_b = step.pin_host_inputs(make_workload(a.config, rank + 1000, a.dtype))
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
        batches = [host if i % 2 == 0 else host_b for i in range(a.steps)]
        step.run_stream(batches[:3])
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
        step.run_stream(batches, consume=lambda i, out: seen.append(float(out[0][0])))
        e1.record()
        barrier()
        assert len(seen) == a.steps
        e2e_ms = e0.elapsed_time(e1)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        step.run_stream(batches, before_step=flush.zero_)
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

    # what bounds a streamed step: its upload over the host link, or its kernels
    e2e_bound = None
    if h2d_gbps and io_graph_ms:
        up_ms = step.h2d_bytes(host) / (h2d_gbps * 1e9) * 1e3
        e2e_bound = "h2d (%.3f ms upload vs %.3f ms of kernels per step)" % (up_ms, io_graph_ms) if up_ms > io_graph_ms \
            else "kernels (%.3f ms per step vs %.3f ms upload)" % (io_graph_ms, up_ms)
    clocks = sampler.stop() if rank == 0 else None
    train = None
    if not a.no_train and cfg.get("pooler", "ROIPool") == "ROIPool":
        train = train_leg(a, wl, dev, world, rank, barrier)
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
            "config": config_dict(wl),
            "detail": {"l2": "flushed (256 MiB write) between timed steps (value, roofline); e2e: see e2e.l2",
                       "stages": step.stages(), "schedule": step.schedule(),
                       "launch": "cuda_graph_replay" if use_graph else "eager"},
            "e2e": {"value": world * R * a.steps / (e2e_ms * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": step.h2d_bytes(host), "d2h_bytes_per_step": step.d2h_bytes(),
                    "mode": e2e_mode,
                    "uploaded": "feature map (step dtype), rois, objectness, MIL + refinement logits, image labels "
                                "(in the real system the feature map is produced on the device by the backbone)",
                    "downloaded": "losses, pseudo-labels, fg/bg + accuracy counters, checksum of grad_feature "
                                  "(the gradient w.r.t. the feature map stays on the device)",
                    "bound": e2e_bound,
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
        if train is not None:
            out.update(train)
        if not a.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(wl, a, kind="port")
        emit(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def train_leg(a, wl, dev, world, rank, barrier):
    """The data-parallel TRAINING step: RoIPool -> fc6 / fc7 -> MIL + Kr refinement heads (~120 M fc
    parameters) under DistributedDataParallel, loss.backward() through the head autograd Functions, SGD step; conv5
    map resident on the device, per-rank synthetic images.  -> extra keys of the JSON line (identical on every rank
    after the max-reduce; only rank 0 prints)."""
    import torch
    import torch.distributed as dist

    from jtsm_b200 import trainstep

    cfg = wl["cfg"]
    ts = trainstep.TrainStep(cfg, dev, world=world, tf32=True, fused_sgd=True)
    batch = ts.make_batch(wl)
    steps = max(3, min(a.steps, 10))

    def timed(fn):
        for _ in range(3):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        return e0.elapsed_time(e1) / steps

    ms_eager = timed(lambda: ts.step(batch, sync=True))
    ar_alone = None
    if world > 1:
        flat = torch.empty(ts.num_params(), dtype=torch.float32, device=dev)
        ar_alone = timed(lambda: dist.all_reduce(flat))
        del flat
    # the whole step (forward, backward, DDP all-reduce, optimizer) as ONE CUDA graph: the eager step is launch bound
    # (~105 kernels, 3.8 ms of GPU work in 5.6 ms at N = 1); skipped with --no-graph or when the capture fails
    ms_graph, graph_err = None, None
    if not a.no_graph:
        try:
            ts.capture(batch)
        except Exception as e:  # noqa: BLE001 -- reported in the JSON line, the eager number stands
            graph_err = "%s: %s" % (type(e).__name__, str(e).splitlines()[0][:200])
        ok = torch.tensor([0.0 if graph_err else 1.0], device=dev)
        if world > 1:  # replay only if EVERY rank captured (the graph holds the all-reduce)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok) > 0:
            ms_graph = timed(ts.replay)
        elif graph_err is None:
            graph_err = "capture failed on another rank"
    loss = None
    if world == 1 and (ms_graph is not None or graph_err is None):  # after a failed capture the RNG state is unusable
        loss = float(ts.replay() if ms_graph is not None else ts.step(batch))
    ms_local = None
    if world > 1 and graph_err is None:  # no all-reduce: replicas diverge, timing only, run last
        ms_local = timed(lambda: ts.step(batch, sync=False))
    tm = torch.tensor([ms_eager, ms_local or 0.0, ar_alone or 0.0, ms_graph if ms_graph is not None else -1.0],
                      dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    ms_eager, ms_local, ar_alone, ms_graph = (float(v) for v in tm.tolist())
    if ms_graph <= 0:
        ms_graph = None
    ms = ms_graph if ms_graph is not None else ms_eager
    return {
        "train_images_per_sec": world * cfg["N"] * 1e3 / ms,
        "allreduce_ms": (ms_eager - ms_local) if (world > 1 and ms_local) else (0.0 if world == 1 else None),
        "train_step": {
            "ms_per_step": ms, "launch": "cuda_graph_replay" if ms_graph is not None else "eager",
            "ms_per_step_eager": ms_eager, "graph_error": graph_err,
            "ms_per_step_without_allreduce": ms_local if world > 1 else None,
            "allreduce_alone_ms": ar_alone if world > 1 else None, "allreduce_bytes": 4 * ts.num_params(),
            "allreduce_ms_means": "exposed, eager: step time under DDP minus the same step inside no_sync()",
            "params": ts.num_params(), "images_per_gpu": cfg["N"], "steps": steps,
            "what": "RoIPool x objectness -> fc6/fc7 (torch GEMMs, TF32 tensor cores) -> WSDDN MIL + BCE -> Kr x "
                    "(pseudo-gt, label assignment, weighted CE[, box reg]) -> backward (autograd Functions, DDP "
                    "gradient all-reduce over NCCL) -> fused SGD momentum step; conv5 map resident on the device",
            "loss_after": loss,
        },
    }


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
        "config": config_dict(wl),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": REF_KIND, "sample": desc + SINGLE_PROCESS},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(json.dumps(out))


# ----------------------------------------------------------------------------------------------
# c4: the multi-scale + flip INFERENCE chain (BASELINE.json configs[3])
# ----------------------------------------------------------------------------------------------
def nms_pairs(det_classes_counts):
    """Box pairs of the per-class suppression matrices (what torchvision's bitmask kernel evaluates): sum m (m - 1) / 2."""
    return sum(m * (m - 1) // 2 for m in det_classes_counts)


def run_b200_infer(a):
    import torch
    import torch.distributed as dist

    from jtsm_b200 import _lib, inferpath, synth
    for kv in a.tune:  # A/B experiments only; the default run sets nothing
        k, v = kv.split("=")
        _lib.set_tuning(k, int(v))

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (impl=b200) needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.lib()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    wl = synth.infer_case(a.config, sweep_index=rank)
    cfg = wl["cfg"]
    rois_per_step = cfg["A"] * cfg["R"]  # RoIs pooled + scored per image
    step = inferpath.InferenceStep(cfg, dev, dtype=torch.bfloat16 if a.dtype == "bf16" else torch.float32)
    host = step.host_inputs(wl)
    d = step.upload(host)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    step.capture(d)
    for _ in range(max(a.warmup, 3)):
        step.graph.replay()
    barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    sampler.mark()
    t0 = time.perf_counter()
    for i in range(a.steps):
        flush.zero_()
        ev[i][0].record()
        step.graph.replay()
        ev[i][1].record()
    barrier()
    t_wall = time.perf_counter() - t0
    dev_ms = sum(e0.elapsed_time(e1) for e0, e1 in ev)
    # the NMS kernel alone (its own graph: the detection call on the merged tensors), CUDA events, L2 flushed
    from jtsm_b200 import inference
    mb, ms = step.last["mean_boxes"], step.last["mean_scores"]
    g2 = torch.cuda.CUDAGraph()
    s2 = torch.cuda.Stream()
    s2.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s2):
        inference.fast_rcnn_inference_device(mb, ms, step.orig_hw, cfg["score_thresh"], cfg["nms_thresh"], cfg["topk"])
    torch.cuda.current_stream().wait_stream(s2)
    torch.cuda.synchronize()
    with torch.cuda.graph(g2):
        inference.fast_rcnn_inference_device(mb, ms, step.orig_hw, cfg["score_thresh"], cfg["nms_thresh"], cfg["topk"])
    kev = []
    for i in range(a.steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g2.replay()
        e1.record()
        kev.append((e0, e1))
    barrier()
    det_ms = sum(e0.elapsed_time(e1) for e0, e1 in kev) / len(kev)
    # candidates per class (score > thresh) -> pair count of the suppression problem
    sc = ms[:, :-1]
    counts = [int(v) for v in (sc > cfg["score_thresh"]).sum(dim=0).tolist()]

    # e2e: every step uploads ALL its inputs from pinned host memory (the conv5 maps of the 10 augmentations included,
    # which the real system produces on the device) and downloads the packed detections
    out_host = torch.empty(step.graph_packed.shape, dtype=step.graph_packed.dtype).pin_memory()
    hosts = [host, step.host_inputs(synth.infer_case(a.config, sweep_index=rank + 1000))]
    for i in range(2):
        step.upload(hosts[i % 2], into=d)
        step.graph.replay()
        out_host.copy_(step.graph_packed, non_blocking=True)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        step.upload(hosts[i % 2], into=d)
        step.graph.replay()
        out_host.copy_(step.graph_packed, non_blocking=True)
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    tm = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = (float(v) for v in tm.tolist())
    if rank == 0:
        pairs = nms_pairs(counts)
        flops = 16.0 * pairs  # ~16 FP32 ops per box pair
        sm_count = torch.cuda.get_device_properties(dev).multi_processor_count
        peak_tflops = sm_count * 128 * 2 * 1.965e9 / 1e12  # FP32 SIMT peak at the maximum SM clock
        ach = flops / (det_ms * 1e-3) / 1e12
        out = {
            "metric": METRIC, "value": world * rois_per_step * a.steps / (dev_ms * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": dev_ms / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32" if a.dtype == "fp32" else "bf16",
            "data": "synthetic (seeded MCG-like proposals mapped into 10 augmented images, relu(randn) conv5 maps, random head logits)",
            "config": config_dict(wl),
            "detail": {"images_per_sec": world * a.steps / (dev_ms * 1e-3), "l2": "flushed (256 MiB write) between timed steps",
                       "stages": step.stages(), "launch": "cuda_graph_replay",
                       "value_counts": "RoIs pooled and scored = augmentations x proposals per image"},
            "e2e": {"value": world * rois_per_step * a.steps / (e2e_ms * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": step.h2d_bytes(host),
                    "d2h_bytes_per_step": out_host.numel() * out_host.element_size(),
                    "mode": "sequential per image: H2D of all inputs, graph replay, D2H of the packed detections",
                    "uploaded": "conv5 maps of the 10 augmentations (produced on the device in the real system), RoIs, "
                                "objectness, refinement logits / deltas", "downloaded": "top-100 detections (boxes, scores, classes, proposal ids, count)",
                    "bound": "h2d"},
            "gpu_launches": int(step.launches_per_step * a.steps),
            "roofline": {"bound": "alu", "kernel": "nms_segment_kernel (inside jtsm_fast_rcnn_inference: 9 launches)",
                         "achieved": ach, "peak": peak_tflops, "unit": "TFLOP/s", "frac": ach / peak_tflops,
                         "peak_source": "FP32 SIMT: %d SMs x 128 lanes x 2 x 1.965 GHz" % sm_count, "traffic": None,
                         "algorithmic_pairs": pairs, "flops_per_pair": 16, "candidates_per_class": counts,
                         "avg_launch_ms": det_ms,
                         "note": "algorithmic work = the full per-class suppression matrices (what a bitmask NMS evaluates); "
                                 "greedy NMS has a sequential chain of ceil(m / 64) dependent block steps per class and "
                                 "20 classes occupy 20 of the SMs, so the fraction of the chip-wide ALU peak is small by "
                                 "construction; avg_launch_ms covers the whole detection call"},
            "clocks": clocks, "wall_s_timed_region": t_wall,
        }
        if not a.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline_infer(wl, a)
        emit(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline_infer(wl, a):
    from oracle import cpu_infer

    cores = os.cpu_count() or 1
    sample = a.cpu_sample_rois or cpu_infer.auto_sample_rois(wl)
    t, n_rois, desc, parts = cpu_infer.time_step(wl, sample, cores)
    return {"value": n_rois / t, "unit": UNIT, "cores": cores, "kind": REF_KIND, "sample": desc, "seconds_per_image_scaled": t,
            "measured_seconds": parts}


def run_reference_infer(a):
    from jtsm_b200 import synth
    from oracle import cpu_infer

    if int(os.environ.get("RANK", "0")) != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = synth.infer_case(a.config, sweep_index=0)
    cores = os.cpu_count() or 1
    sample = a.cpu_sample_rois or cpu_infer.auto_sample_rois(wl, budget_s=4.0)
    ts, desc, n_rois = [], "", 0
    for i in range(max(1, min(a.warmup, 1)) + a.steps):
        t, n_rois, desc, _ = cpu_infer.time_step(wl, sample, cores)
        if i >= max(1, min(a.warmup, 1)):
            ts.append(t)
    v = n_rois * len(ts) / sum(ts)
    emit(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": world, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * sum(ts) / len(ts), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(wl),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": REF_KIND, "sample": desc + SINGLE_PROCESS},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}))


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
    from jtsm_b200 import synth

    infer = a.config in synth.INFER_CONFIGS
    if a.impl == "reference":
        (run_reference_infer if infer else run_reference)(a)
    else:
        (run_b200_infer if infer else run_b200)(a)


if __name__ == "__main__":
    main()
