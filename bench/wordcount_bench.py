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
except Exception as ex:  # noqa: BLE001
    res["gpu_error"] = str(ex)[:200]
print(json.dumps(res))
