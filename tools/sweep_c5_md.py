"""Render a tools/sweep_c5.py record file as a markdown table, next to an earlier round's records when given.
Usage: python tools/sweep_c5_md.py new.jsonl [old.jsonl] > table.md"""
import json
import sys


def load(p):
    out = {}
    for l in open(p):
        l = l.strip()
        if l.startswith("{"):
            d = json.loads(l)
            out[(d["stride"], d["P"], d["R"], d["dtype"])] = d
    return out


new = load(sys.argv[1])
old = load(sys.argv[2]) if len(sys.argv) > 2 else {}
print("| stride | map | P | R | dtype | RoIPool fwd | bwd | (earlier fwd / bwd) | tv fwd | tv bwd | RoIAlign fwd | bwd | tv fwd | tv bwd "
      "| fwd+bwd speed-up pool / align |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|")
gains = []
for k in sorted(new, key=lambda t: (t[0], t[1], t[2], t[3] != "float32")):
    d = new[k]
    o = old.get(k)
    f = lambda v: "%.0f" % v if v is not None else "-"
    tv = "tv_roi_pool_fwd_us" in d
    sp = "-"
    if tv:
        sp = "%.1fx / %.1fx" % ((d["tv_roi_pool_fwd_us"] + d["tv_roi_pool_bwd_us"]) / (d["roi_pool_fwd_us"] + d["roi_pool_bwd_us"]),
                                (d["tv_roi_align_fwd_us"] + d["tv_roi_align_bwd_us"]) / (d["roi_align_fwd_us"] + d["roi_align_bwd_us"]))
    if o:
        gains.append((k, (o["roi_pool_fwd_us"] + o["roi_pool_bwd_us"]) / (d["roi_pool_fwd_us"] + d["roi_pool_bwd_us"])))
    print("| %d | %dx%d | %d | %d | %s | %s | %s | %s | %s | %s | %s | %s | %s | %s | %s |" % (
        d["stride"], d["map"][0], d["map"][1], d["P"], d["R"], d["dtype"], f(d["roi_pool_fwd_us"]), f(d["roi_pool_bwd_us"]),
        ("%s / %s" % (f(o["roi_pool_fwd_us"]), f(o["roi_pool_bwd_us"]))) if o else "-",
        f(d.get("tv_roi_pool_fwd_us")), f(d.get("tv_roi_pool_bwd_us")), f(d["roi_align_fwd_us"]), f(d["roi_align_bwd_us"]),
        f(d.get("tv_roi_align_fwd_us")), f(d.get("tv_roi_align_bwd_us")), sp))
if gains:
    g = sorted(v for _, v in gains)
    print("\nRoIPool fwd+bwd against the earlier records: min %.2fx, median %.2fx, max %.2fx over %d points."
          % (g[0], g[len(g) // 2], g[-1], len(g)))
    for dt in ("float32", "bfloat16"):
        gg = sorted(v for k, v in gains if k[3] == dt)
        if gg:
            print("%s: median %.2fx (min %.2fx, max %.2fx)." % (dt, gg[len(gg) // 2], gg[0], gg[-1]))
