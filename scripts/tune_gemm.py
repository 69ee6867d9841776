"""Measure every GEMM / conv shape of the SD-1.5 task under each (block_n, splitk) and keep the fastest.

Runs the real pipeline once with recording on to collect the shape keys, then times each shape inside a
CUDA graph (16 back-to-back calls rotating over weight copies so weights come from HBM, as in the real
step) for every legal configuration. Writes arbius_b200/ops/tuned_sm100.json (if --write) and a report.
"""
import argparse
import json
import sys
from collections import Counter

import torch

sys.path.insert(0, ".")
import arbius_b200.ops  # noqa: E402,F401
from arbius_b200 import ops  # noqa: E402

G = sys.modules["arbius_b200.ops.gemm"]
BNS = [32, 64, 96, 128, 160, 256]
SPLITS = [1, 2, 3, 4, 6, 8]
REPS = 16


def time_graph(fn, iters=5):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best * 1e3 / REPS   # us per call


def bench_key(key, dt=torch.float16, only_splitk=None):
    parts = key.split(":")
    kind = parts[0]
    res = {}
    if kind in ("lin", "geglu"):
        m, n, k = int(parts[1]), int(parts[2]), int(parts[3])
        wbytes = n * k * 2
        ncopy = max(2, min(REPS, (300 << 20) // max(wbytes, 1)))
        x = torch.randn(m, k, device="cuda", dtype=dt)
        ws = [torch.randn(n, k, device="cuda", dtype=dt) / k ** 0.5 for _ in range(ncopy)]
        b = torch.randn(n, device="cuda", dtype=dt)
        kc = (k + 63) // 64
        flops = 2.0 * m * n * k
        for bn in BNS:
            if kind == "geglu" and bn % 32:
                continue
            if bn > 64 and bn >= 2 * n:
                continue
            for sk in ([only_splitk] if only_splitk else SPLITS):
                if only_splitk is None and sk > 1 and (kc // sk < 4 or m * n * sk * 4 > (512 << 20)):
                    continue
                def fn():
                    for i in range(REPS):
                        ops.gemm(x, ws[i % ncopy], b, geglu=(kind == "geglu"), block_n=bn, splitk=sk)
                try:
                    res[(bn, sk)] = time_graph(fn)
                except RuntimeError as e:
                    res[(bn, sk)] = float("inf")
    elif kind == "tconv":
        bfhw, n, c = parts[1], int(parts[2]), int(parts[3])
        b_, f, hw = (int(v) for v in bfhw.split("x"))
        wbytes = n * 3 * c * 2
        ncopy = max(2, min(REPS, (300 << 20) // wbytes))
        x = torch.randn(b_, f, hw, c, device="cuda", dtype=dt)
        ws = [torch.randn(n, 3 * c, device="cuda", dtype=dt) / (3 * c) ** 0.5 for _ in range(ncopy)]
        bias = torch.randn(n, device="cuda", dtype=dt)
        flops = 2.0 * b_ * f * hw * n * 3 * c
        for bn in BNS:
            if bn > 64 and bn >= 2 * n:
                continue
            for sk in ([only_splitk] if only_splitk else [1]):
                def fn():
                    for i in range(REPS):
                        ops.conv_temporal(x, ws[i % ncopy], bias, block_n=bn, splitk=sk)
                try:
                    res[(bn, sk)] = time_graph(fn)
                except RuntimeError:
                    res[(bn, sk)] = float("inf")
    else:
        bhw, n, c = parts[1], int(parts[2]), int(parts[3])
        b_, h, w = (int(v) for v in bhw.split("x"))
        wbytes = n * 9 * c * 2
        ncopy = max(2, min(REPS, (300 << 20) // wbytes))
        x = torch.randn(b_, h, w, c, device="cuda", dtype=dt)
        ws = [torch.randn(n, 9 * c, device="cuda", dtype=dt) / (9 * c) ** 0.5 for _ in range(ncopy)]
        bias = torch.randn(n, device="cuda", dtype=dt)
        kc = 9 * c // 64
        m = b_ * h * w
        flops = 2.0 * m * n * 9 * c
        for bn in BNS:
            if bn > 64 and bn >= 2 * n:
                continue
            for sk in ([only_splitk] if only_splitk else SPLITS):
                if only_splitk is None and sk > 1 and (kc // sk < 4 or m * n * sk * 4 > (512 << 20)):
                    continue
                def fn():
                    for i in range(REPS):
                        ops.conv3x3(x, ws[i % ncopy], bias, block_n=bn, splitk=sk)
                try:
                    res[(bn, sk)] = time_graph(fn)
                except RuntimeError:
                    res[(bn, sk)] = float("inf")
    return res, flops, wbytes


def _load_table():
    """The table being built in this gpurun call (gpurun_out/) if there is one, else the tree's."""
    import os
    for path in ("gpurun_out/tuned_sm100.json", "arbius_b200/ops/tuned_sm100.json"):
        if os.path.exists(path):
            return json.load(open(path))
    return {}


def tune_batched(a):
    """For every tuned canonical shape: the same op with task_batch x the rows, split-K fixed, best block_n."""
    canon = _load_table()
    out = {}
    # only the shapes AnythingV3 itself runs (the table also holds Kandinsky's, their row bands and earlier batches)
    from arbius_b200.models.sd15.pipeline import SD15Config, SD15Pipeline
    pipe = SD15Pipeline(SD15Config(), backend="native", device="cuda", dtype=torch.float16, use_graph=False)
    G.RECORD = []
    pipe.generate("cat", "", a.size, a.size, 2, 7.5, "DPMSolverMultistep", seed=1)
    torch.cuda.synchronize()
    mine = set(k for k in G.RECORD if k.split(":")[0] in ("lin", "geglu", "conv"))
    G.RECORD = None
    del pipe
    torch.cuda.empty_cache()
    for key, (bn0, sk) in sorted(canon.items()):
        if key not in mine:
            continue
        parts = key.split(":")
        if parts[0] == "conv":
            b_, h, w = (int(v) for v in parts[1].split("x"))
            if b_ != 2:
                continue                                   # only the UNet's shapes (batch 2) grow with the task batch
            big = f"conv:{b_ * a.task_batch}x{h}x{w}:{parts[2]}:{parts[3]}"
        elif parts[0] in ("lin", "geglu"):
            big = f"{parts[0]}:{int(parts[1]) * a.task_batch}:{parts[2]}:{parts[3]}"
        else:
            continue
        res, flops, _ = bench_key(big, only_splitk=sk)
        (bn, _), us = min(res.items(), key=lambda kv: kv[1])
        base = res.get((bn0, sk), float("nan"))
        out[big] = [bn, sk]
        print(json.dumps(dict(key=big, canonical=[bn0, sk], best_bn=bn, us=round(us, 2), canonical_bn_us=round(base, 2),
                              tflops=round(flops / us / 1e6, 1))), flush=True)
    json.dump(out, open(a.out.replace(".json", f"_tb{a.task_batch}.json"), "w"), indent=0, sort_keys=True)
    canon.update(out)
    json.dump(canon, open("gpurun_out/tuned_sm100.json", "w"), indent=0, sort_keys=True)


def tune_kandinsky(a):
    """Kandinsky 2.1's UNet shapes: (block_n, splitk) for every recorded shape of the undivided task — this FIXES the
    canonical split-K of those shapes — and then, for the row bands of patch parallelism (1/P of the rows, P in
    --bands), the best tile width alone with split-K held at the canonical shape's. Writes the merged table."""
    from arbius_b200.models.kandinsky.pipeline import Kandinsky2Config, Kandinsky2Pipeline
    cfg = Kandinsky2Config()
    cfg.steps, cfg.prior_steps = 1, 1
    pipe = Kandinsky2Pipeline(cfg, backend="native", device="cuda", dtype=torch.float16, use_graph=False)
    G.RECORD = []
    pipe.generate("cat", 768, 768, seed=1)
    torch.cuda.synchronize()
    counts = Counter(k for k in G.RECORD if k.split(":")[0] in ("lin", "geglu", "conv"))
    G.RECORD = None
    del pipe
    torch.cuda.empty_cache()
    table = _load_table()
    new, report = {}, []
    for key, cnt in sorted(counts.items(), key=lambda kv: -kv[1]):
        parts = key.split(":")
        if parts[0] == "conv":
            b_, h, w = (int(v) for v in parts[1].split("x"))
            m = b_ * h * w
        else:
            m = int(parts[1])
        if m < 256 or (key in table and not a.retune):
            continue                                   # tiny per-task GEMMs (embeddings, prior): the heuristic is fine
        res, flops, _ = bench_key(key)
        (bn, sk), us = min(res.items(), key=lambda kv: kv[1])
        new[key] = [bn, sk]
        report.append(dict(key=key, calls=cnt, best=[bn, sk], us=round(us, 2), tflops=round(flops / us / 1e6, 1),
                           all={f"{b2}/{s2}": round(v, 2) for (b2, s2), v in sorted(res.items())}))
        print(json.dumps({k: v for k, v in report[-1].items() if k != "all"}), flush=True)
        for P in a.bands:
            if parts[0] == "conv":
                if h % 8 or ((h // 8) * w) % 8 or h % P:
                    continue
                band = f"conv:{b_}x{h // P}x{w}:{parts[2]}:{parts[3]}"
            else:
                if m % (P * 8):
                    continue
                band = f"{parts[0]}:{m // P}:{parts[2]}:{parts[3]}"
            if band in new or (band in table and not a.retune):
                continue
            r2, f2, _ = bench_key(band, only_splitk=sk)
            (bn2, _), us2 = min(r2.items(), key=lambda kv: kv[1])
            new[band] = [bn2, sk]
            report.append(dict(key=band, band_of=key, P=P, best=[bn2, sk], us=round(us2, 2), tflops=round(f2 / us2 / 1e6, 1),
                               all={f"{b2}/{s2}": round(v, 2) for (b2, s2), v in sorted(r2.items())}))
            print(json.dumps({k: v for k, v in report[-1].items() if k != "all"}), flush=True)
    table.update(new)
    json.dump(report, open(a.out.replace(".json", "_kandinsky.json"), "w"), indent=1)
    json.dump(table, open("gpurun_out/tuned_sm100.json", "w"), indent=0, sort_keys=True)
    print(f"tuned {len(new)} shapes; merged table in gpurun_out/tuned_sm100.json", flush=True)


def tune_video(a):
    """ZeroScope XL's UNet3D shapes (1024x576x24): the best TILE WIDTH for every recorded shape with split-K held at what
    the heuristic gives today, so the bytes of a task do not change. Writes the merged table."""
    from arbius_b200.models.video.pipeline import TextToVideoConfig, TextToVideoPipeline
    pipe = TextToVideoPipeline(TextToVideoConfig.zeroscope_xl(), backend="native", device="cuda", dtype=torch.float16,
                               family="zeroscopev2xl", use_graph=False)
    G.RECORD = []
    pipe.generate("cat", num_inference_steps=1, seed=1)
    torch.cuda.synchronize()
    counts = Counter(k for k in G.RECORD if k.split(":")[0] in ("lin", "geglu", "conv", "tconv"))
    G.RECORD = None
    del pipe
    torch.cuda.empty_cache()
    table = _load_table()
    new, report = {}, []
    for key, cnt in sorted(counts.items(), key=lambda kv: -kv[1]):
        parts = key.split(":")
        n = int(parts[2])
        if parts[0] == "conv":
            b_, h, w = (int(v) for v in parts[1].split("x"))
            m, kc = b_ * h * w, 9 * int(parts[3]) // 64
        elif parts[0] == "tconv":
            b_, f, hw = (int(v) for v in parts[1].split("x"))
            m, kc = b_ * f * hw, 3 * int(parts[3]) // 64
        else:
            m, kc = int(parts[1]), (int(parts[3]) + 63) // 64
        if m < 4096 or (key in table and not a.retune):
            continue
        sk = G.auto_splitk(m, n, kc)
        res, flops, _ = bench_key(key, only_splitk=sk)
        (bn, _), us = min(res.items(), key=lambda kv: kv[1])
        base = min((v for (b2, s2), v in res.items()), default=float("nan"))
        new[key] = [bn, sk]
        report.append(dict(key=key, calls=cnt, best=[bn, sk], us=round(us, 2), tflops=round(flops / us / 1e6, 1),
                           all={f"{b2}/{s2}": round(v, 2) for (b2, s2), v in sorted(res.items())}))
        print(json.dumps(report[-1]), flush=True)
    table.update(new)
    json.dump(report, open(a.out.replace(".json", "_zeroscope.json"), "w"), indent=1)
    json.dump(table, open("gpurun_out/tuned_sm100.json", "w"), indent=0, sort_keys=True)
    print(f"tuned {len(new)} shapes; merged table in gpurun_out/tuned_sm100.json", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="anythingv3", choices=["anythingv3", "kandinsky2", "zeroscopev2xl"])
    ap.add_argument("--retune", action="store_true", help="measure shapes again even if the table has them")
    ap.add_argument("--bands", type=lambda v: [int(x) for x in v.split(",")], default=[2, 4, 8])
    ap.add_argument("--write", action="store_true")
    ap.add_argument("--out", default="gpurun_out/tune_report.json")
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--task-batch", type=int, default=0,
                    help="tune the tile width (only) of every recorded shape at this many tasks per batch; split-K stays "
                         "the canonical entry's, so the bytes per task do not change")
    a = ap.parse_args()
    if a.task_batch > 1:
        return tune_batched(a)
    if a.model == "kandinsky2":
        return tune_kandinsky(a)
    if a.model == "zeroscopev2xl":
        return tune_video(a)
    from arbius_b200.models.sd15.pipeline import SD15Config, SD15Pipeline
    pipe = SD15Pipeline(SD15Config(), backend="native", device="cuda", dtype=torch.float16, use_graph=False)
    G.RECORD = []
    pipe.generate("cat", "", a.size, a.size, 2, 7.5, "DPMSolverMultistep", seed=1)
    torch.cuda.synchronize()
    counts = Counter(k for k in G.RECORD if k.split(":")[0] in ("lin", "geglu", "conv"))
    G.RECORD = None
    del pipe
    torch.cuda.empty_cache()
    report, table = [], {}
    for key, cnt in sorted(counts.items(), key=lambda kv: -kv[1]):
        res, flops, wbytes = bench_key(key)
        (bn, sk), us = min(res.items(), key=lambda kv: kv[1])
        parts = key.split(":")
        if parts[0] == "conv":
            b_, h, w = (int(v) for v in parts[1].split("x"))
            m, n, kc = b_ * h * w, int(parts[2]), 9 * int(parts[3]) // 64
        else:
            m, n, kc = int(parts[1]), int(parts[2]), (int(parts[3]) + 63) // 64
        cur_sk = G.auto_splitk(m, n, kc)
        cur = min((v for (b2, s2), v in res.items() if s2 == cur_sk), default=float("nan"))
        table[key] = [bn, sk]
        report.append(dict(key=key, calls_per_2steps=cnt, best=[bn, sk], us=round(us, 2),
                           tflops=round(flops / us / 1e6, 1), w_gbs=round(wbytes / us / 1e3, 1),
                           heuristic_splitk=cur_sk, heuristic_best_us=round(cur, 2),
                           all={f"{b2}/{s2}": round(v, 2) for (b2, s2), v in sorted(res.items())}))
        print(json.dumps({k: v for k, v in report[-1].items() if k != "all"}), flush=True)
    json.dump(report, open(a.out, "w"), indent=1)
    if a.write:
        merged = _load_table()
        merged.update(table)
        json.dump(merged, open("gpurun_out/tuned_sm100.json", "w"), indent=0, sort_keys=True)


if __name__ == "__main__":
    main()
