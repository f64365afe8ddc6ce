"""pool -> fc6: the fused tcgen05 kernel vs (RoIPool kernel + cuBLAS bf16 GEMM) at config 2 (C = 512, 75 x 100, 2000
RoIs, 7 x 7, fc6 25088 -> 4096).  CUDA events, L2 flushed, median of --iters."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from jtsm_b200 import ops, synth  # noqa: E402


def timeit(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    dev = torch.device("cuda:0")
    cfg, feat, boxes, rois, obj = synth.pool_case("c2")
    nout = 4096
    g = synth.generator(99)
    x = feat.to(dev).to(torch.bfloat16)
    x_nhwc = x.permute(0, 2, 3, 1).contiguous()
    rd, sd = rois.to(dev), (obj + 1.0).to(dev)
    w = (torch.randn(nout, cfg["C"] * 49, generator=g) * 0.01).to(dev)
    b = torch.randn(nout, generator=g).to(dev)
    wb16 = w.to(torch.bfloat16)
    wbin = ops.fc6_weight_binmajor(w, cfg["C"], 7)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    P = (7, 7)

    def unfused():
        pooled, _ = ops.roi_pool_forward(x, rd, cfg["scale"], P, row_scale=sd, argmax_dtype=torch.int16)
        return torch.relu(torch.addmm(b.to(torch.bfloat16), pooled.flatten(1), wb16.t()))

    def gemm_only(pooled):
        return torch.relu(torch.addmm(b.to(torch.bfloat16), pooled.flatten(1), wb16.t()))

    fused = lambda: ops.pool_fc6_forward(x_nhwc, rd, cfg["scale"], P, wbin, b, row_scale=sd, relu=True)
    got, want = fused(), unfused().float()
    err = float((got - want).abs().max() / want.abs().max())
    pooled, _ = ops.roi_pool_forward(x, rd, cfg["scale"], P, row_scale=sd, argmax_dtype=torch.int16)
    flops = 2.0 * rois.shape[0] * cfg["C"] * 49 * nout
    res = dict(config="c2", rois=rois.shape[0], K=cfg["C"] * 49, N=nout, rel_err_vs_unfused_bf16_out=err,
               fused_us=timeit(fused, iters, flush), unfused_pool_plus_cublas_us=timeit(unfused, iters, flush),
               roi_pool_bf16_us=timeit(lambda: ops.roi_pool_forward(x, rd, cfg["scale"], P, row_scale=sd, argmax_dtype=torch.int16), iters, flush),
               cublas_bf16_gemm_us=timeit(lambda: gemm_only(pooled), iters, flush),
               nhwc_staging_copy_us=timeit(lambda: x.permute(0, 2, 3, 1).contiguous(), iters, flush), gemm_flop=flops)
    res["fused_tflops"] = flops / res["fused_us"] / 1e6
    res["cublas_tflops"] = flops / res["cublas_bf16_gemm_us"] / 1e6
    print(json.dumps(res))


if __name__ == "__main__":
    main()
