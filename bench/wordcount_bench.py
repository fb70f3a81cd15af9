"""BASELINE config 1: WordCount.  Local CPU mode through the public API (LocalJobSubmission-style
in-process job), the C++ host baseline (csrc/host/oracle_cpu.cpp), and — when a GPU is present —
the tokenizer + hash-aggregate kernels.  Prints one JSON line."""
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dryad_b200 as D  # noqa: E402
from dryad_b200.models import wordcount as WC  # noqa: E402

lines = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
text = WC.generate_text(lines)
res = {"bytes": len(text), "lines": lines}
fd, path = tempfile.mkstemp(suffix=".txt")
os.write(fd, text)
os.close(fd)
ctx = D.DryadLinqContext(4)
t0 = time.perf_counter()
api = WC.through_api(ctx, path)
res["api_local_cpu_s"] = round(time.perf_counter() - t0, 3)
res["api_local_cpu_MBps"] = round(len(text) / res["api_local_cpu_s"] / 1e6, 2)
t0 = time.perf_counter()
base = WC.cpu_baseline(text.replace(b"\n", b" "))
res["cpp_baseline_s"] = round(time.perf_counter() - t0, 4)
res["cpp_baseline_MBps"] = round(len(text) / res["cpp_baseline_s"] / 1e6, 1)
res["distinct_words"] = len(api)
try:
    import torch
    if torch.cuda.is_available():
        raw = text.replace(b"\n", b" ")
        t = torch.frombuffer(bytearray(raw), dtype=torch.uint8).cuda()
        for _ in range(2):
            out = WC.gpu_wordcount(t)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            out = WC.gpu_wordcount(t)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        res["gpu_ms"] = round(ms, 3)
        res["gpu_GBps"] = round(len(raw) / ms / 1e6, 2)
        res["gpu_collisions"] = int(out[4])
        # the same query through the PUBLIC API on the GPU backend: text file -> string column ->
        # tokenizer -> hash aggregate with verification -> (word, count) pairs on the host
        gctx = D.DryadLinqContext(1, device="cuda")
        for _ in range(2):
            got = WC.through_api(gctx, path)
        t0 = time.perf_counter()
        reps = 3
        for _ in range(reps):
            got = WC.through_api(gctx, path)
        dt = (time.perf_counter() - t0) / reps
        res["api_gpu_s"] = round(dt, 4)
        res["api_gpu_GBps"] = round(len(text) / dt / 1e9, 2)
        res["api_gpu_host_fallbacks"] = gctx._backend().host_fallbacks
        res["api_gpu_equals_cpu_api"] = got == api
        # same, the text already in HBM (no file read, no H2D): what the engine itself costs
        dtext = torch.frombuffer(bytearray(text), dtype=torch.uint8).cuda()

        def resident():
            q = (gctx.from_text(dtext).select_many(lambda l: l.split(" ")).group_by(lambda w: w)
                 .select(lambda g: D.Pair(g.key, g.count())))
            return q.to_columns()
        for _ in range(2):
            cols = resident()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            cols = resident()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / 5
        res["api_gpu_resident_s"] = round(dt, 5)
        res["api_gpu_resident_GBps"] = round(len(text) / dt / 1e9, 2)
        res["api_gpu_resident_groups"] = int(cols["item2"].numel())
        if os.environ.get("WC_PROFILE"):
            import cProfile, pstats
            pr = cProfile.Profile(); pr.enable(); resident(); torch.cuda.synchronize(); pr.disable()
            pstats.Stats(pr).sort_stats("cumulative").print_stats(25)
except Exception as ex:  # noqa: BLE001
    res["gpu_error"] = str(ex)[:200]
print(json.dumps(res))
