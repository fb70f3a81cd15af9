"""Differential run of the random query generators (tests/fuzz_queries.py) on the GPU backend against the
LocalDebug oracle.  Prints one line per mismatch and a summary; meant for a short `gpurun` call.

    python bench/fuzz_gpu.py [first_seed] [last_seed] [seconds]
"""
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from fuzz_columnar import build_columnar  # noqa: E402
from fuzz_queries import build, rand_pipeline  # noqa: E402
from dryad_b200.api.context import DryadLinqContext  # noqa: E402


def _plain(v):
    if isinstance(v, dict):
        return tuple(sorted((k, _plain(x)) for k, x in v.items()))
    if isinstance(v, (list, tuple)):
        return tuple(_plain(x) for x in v)
    if isinstance(v, float):
        return round(v, 6) + 0.0
    return v


def norm(x, ordered):
    if isinstance(x, tuple) and x and x[0] == "ERR":
        return ("ERR", x[1])
    if isinstance(x, list):
        x = [_plain(v) for v in x]
        return x if ordered is True else sorted(x, key=repr)     # "keyed" order: ties may differ, compare as multisets
    return _plain(x)


def ev(make, seed, ctx):
    rng = random.Random(seed)
    try:
        r = make(rng, ctx)
        kind, q, ordered = (r[1], r[0], r[2]) if make is rand_pipeline else r
        if make is build_columnar:
            kind, q, ordered = r
        if q is None:
            return None
        if isinstance(q, tuple):
            return kind, ordered, q[1]()
        return kind, ordered, list(q)
    except Exception as ex:  # noqa: BLE001
        return "?", True, ("ERR", getattr(ex, "error_code", type(ex).__name__), str(ex)[:100])


MAKERS = (rand_pipeline, build)


def main():
    global MAKERS
    if os.environ.get("FUZZ_COLUMNAR"):
        MAKERS = (build_columnar,)
    a, b = int(sys.argv[1]) if len(sys.argv) > 1 else 0, int(sys.argv[2]) if len(sys.argv) > 2 else 40
    budget = float(sys.argv[3]) if len(sys.argv) > 3 else 60.0
    t0 = time.time()
    bad = checked = fallbacks = 0
    for seed in range(a, b):
        for make in MAKERS:
            if time.time() - t0 > budget:
                break
            gpu = DryadLinqContext(3, device=os.environ.get("FUZZ_DEVICE", "cuda:0"), work_dir=f"/tmp/fzg/{seed}")
            got = ev(make, seed, gpu)
            want = ev(make, seed, DryadLinqContext(3, local_debug=True))
            if got is None or want is None:
                continue
            checked += 1
            be = gpu.__dict__.get("_backend_obj") or None
            if norm(got[2], got[1]) != norm(want[2], want[1]):
                bad += 1
                print(f"MISMATCH {make.__name__} seed {seed} {want[0]}\n   want {str(want[2])[:160]}\n   got  {str(got[2])[:160]}", flush=True)
    print(f"checked {checked} mismatches {bad} seconds {time.time() - t0:.1f}", flush=True)


if __name__ == "__main__":
    main()
