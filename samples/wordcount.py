"""The canonical WordCount sample (reference: samples/WordCount.cs.pp:22-90) in its three variants:

  python samples/wordcount.py --mode local  [input.txt]      # LocalJobSubmission-style, host CPU
  python samples/wordcount.py --mode debug  [input.txt]      # LocalDebug (LINQ-to-objects oracle)
  python samples/wordcount.py --mode cluster [input.txt]     # a graph-manager process + vertex-host processes
  python samples/wordcount.py --mode gpu    [input.txt]      # one B200: tokenizer + hash-aggregate kernels
  python samples/wordcount.py --mode gpu-api [input.txt]     # one B200, through the query API (text stays in HBM)

The query is the reference's:  lines.SelectMany(l => l.Split(' ')).GroupBy(w => w)
                                    .Select(g => new Pair(g.Key, g.Count())).OrderByDescending(count).Take(20)
"""
import argparse
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dryad_b200 as D  # noqa: E402

SAMPLE = ("the quick brown fox jumps over the lazy dog\n" * 200 +
          "dryad runs dataflow graphs of vertices and channels\n" * 120 +
          "the graph manager schedules vertices close to their inputs\n" * 60)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", choices=["local", "debug", "cluster", "gpu", "gpu-api"], default="local")
    ap.add_argument("--partitions", type=int, default=4)
    ap.add_argument("input", nargs="?")
    a = ap.parse_args(argv)
    path = a.input
    if path is None:
        fd, path = tempfile.mkstemp(suffix=".txt")
        os.write(fd, SAMPLE.encode())
        os.close(fd)
    if a.mode == "gpu":
        import torch
        from dryad_b200.models import wordcount
        raw = open(path, "rb").read().replace(b"\n", b" ")
        text = torch.frombuffer(bytearray(raw), dtype=torch.uint8).cuda()
        rep_start, rep_len, counts, total, collisions = wordcount.gpu_wordcount(text, keep_empty=False)
        if collisions:
            raise SystemExit("two words shared a 64-bit hash: rerun with --mode local")
        d = wordcount.gpu_result_to_dict(raw, rep_start, rep_len, counts)
        top = sorted(((w.decode(), c) for w, c in d.items()), key=lambda wc: (-wc[1], wc[0]))[:20]
    else:
        from dryad_b200.api.context import PlatformKind
        ctx = D.DryadLinqContext(a.partitions, local_debug=(a.mode == "debug"), device="cuda:0" if a.mode == "gpu-api" else "cpu",
                                 platform=PlatformKind.LOCAL_CLUSTER if a.mode == "cluster" else PlatformKind.LOCAL)
        lines = ctx.from_store(f"file://{path}") if hasattr(ctx, "from_store") else ctx.from_enumerable(open(path).read().splitlines(), a.partitions)
        q = (lines.select_many(lambda l: str(l).split(" "))
                  .where(lambda w: w != "")
                  .group_by(lambda w: w)
                  .select(lambda g: D.Pair(g.key, g.count())))
        top = sorted(q.to_list(), key=lambda p: (-p[1], p[0]))[:20]
    for w, c in top:
        print(f"{c:8d}  {w}")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
