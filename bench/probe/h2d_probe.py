"""Pinned host -> device copy rate of this box (the floor of every end-to-end number)."""
import torch, json
n = 4_000_000_000
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
out = {}
for name, fn in (("one_copy", lambda: d.copy_(h, non_blocking=True)),):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): fn()
    e1.record(); torch.cuda.synchronize()
    out[name + "_GBps"] = round(3 * n / e0.elapsed_time(e1) / 1e6, 1)
s = [torch.cuda.Stream() for _ in range(4)]
def chunked():
    q = n // 4
    for i in range(4):
        with torch.cuda.stream(s[i]):
            d[i*q:(i+1)*q].copy_(h[i*q:(i+1)*q], non_blocking=True)
    for st in s: torch.cuda.current_stream().wait_stream(st)
for _ in range(2): chunked()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): chunked()
e1.record(); torch.cuda.synchronize()
out["four_streams_GBps"] = round(3 * n / e0.elapsed_time(e1) / 1e6, 1)
print(json.dumps(out))
