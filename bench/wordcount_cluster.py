"""WordCount through the query API on the CPU: one process (LOCAL) against cluster mode (a graph-manager
process + N vertex-host processes), same text, same query.  Prints one JSON line per mode.

    python bench/wordcount_cluster.py [--mb 512] [--hosts 8]
"""
import argparse
import json
import os
import random
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.api.context import DryadLinqContext, PlatformKind  # noqa: E402


def make_text(path, mb):
    rng = random.Random(5)
    vocab = ["w%05d" % i for i in range(50000)]
    with open(path, "wb") as f:
        chunk = []
        size = 0
        while size < (1 << 20):
            line = " ".join(rng.choice(vocab) for _ in range(12)) + "\n"
            chunk.append(line)
            size += len(line)
        block = "".join(chunk).encode()
        for _ in range(mb):
            f.write(block)
    return os.path.getsize(path)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=512)
    ap.add_argument("--hosts", type=int, default=os.cpu_count() or 4)
    a = ap.parse_args()
    d = tempfile.mkdtemp(prefix="wcbench_")
    path = os.path.join(d, "text.txt")
    nbytes = make_text(path, a.mb)
    for mode in ("local", "cluster"):
        ctx = DryadLinqContext(a.hosts if mode == "cluster" else 1, work_dir=os.path.join(d, mode),
                               platform=PlatformKind.LOCAL_CLUSTER if mode == "cluster" else PlatformKind.LOCAL)
        ctx.job_max_nodes = a.hosts
        t0 = time.perf_counter()
        q = (ctx.from_store(f"file://{path}").select_many(lambda l: l.split(" ")).group_by(lambda w: w)
                .select(lambda g: (g.key, g.count())))
        out = q.to_list()
        dt = time.perf_counter() - t0
        rec = {"bench": "wordcount_api_cpu", "mode": mode, "bytes": nbytes, "seconds": round(dt, 3),
               "MB_per_s": round(nbytes / dt / 1e6, 1), "distinct_words": len(out), "total_words": sum(c for _, c in out)}
        if mode == "cluster":
            st = ctx.last_cluster_status
            rec["hosts"] = st["hosts"]
            rec["graph_seconds"] = round(st["seconds"], 3)            # inside the manager: first dispatch to commit
            rec["MB_per_s_graph"] = round(nbytes / st["seconds"] / 1e6, 1)
            rec["vertex_seconds_max"] = round(max(v["seconds"] for v in st["vertices"]), 3)
        else:
            rec["native_vertices"] = ctx.__dict__.get("native_host_vertices", 0)
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
