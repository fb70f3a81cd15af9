"""WordCount (BASELINE config 1) on the CPU local mode, the oracle, the C++ baseline and the GPU."""
import collections

import pytest
import torch

import dryad_b200 as D
from dryad_b200.models import wordcount as WC


def _expected(text: bytes):
    c = collections.Counter()
    for line in text.split(b"\n")[:-1]:
        c.update(line.split(b" "))
    return c


def test_wordcount_api_local_mode_matches_oracle_and_cpp(tmp_path):
    text = WC.generate_text(400, 9, seed=3, vocab=300) + b"a  b \n\nlast line\n"
    p = tmp_path / "words.txt"
    p.write_bytes(text)
    want = {k.decode(): v for k, v in _expected(text).items()}
    for dbg, parts in ((True, 1), (False, 3)):
        ctx = D.DryadLinqContext(parts, local_debug=dbg)
        assert WC.through_api(ctx, str(p)) == want
    groups, total = WC.cpu_baseline(text, threads=4)
    assert groups == len(want) and total == sum(want.values())


@pytest.mark.gpu
def test_wordcount_gpu_kernels_match_python():
    text = WC.generate_text(5000, 11, seed=5, vocab=2000) + b"x  y \n\n"
    t = torch.tensor(list(text), dtype=torch.uint8, device="cuda")
    rs, rl, cnt, total, collisions = WC.gpu_wordcount(t)
    want = _expected(text)
    assert collisions == 0 and total == sum(want.values())
    assert WC.gpu_result_to_dict(text, rs, rl, cnt) == dict(want)


def test_wordcount_on_the_host_backend_runs_as_a_native_vertex(tmp_path):
    """The traced [SelectMany(split) -> GroupBy(Count)] vertex over a text partition runs in the native
    host library (csrc/host/vertex_native.cpp), not over Python records, and agrees with LocalDebug."""
    import dryad_b200 as D
    from dryad_b200.models import wordcount as WC
    text = WC.generate_text(3000, seed=5, vocab=400) + b"tail  with  double spaces \n\nlast"
    path = tmp_path / "t.txt"
    path.write_bytes(text)
    want = WC.through_api(D.DryadLinqContext(1, local_debug=True), str(path))
    ctx = D.DryadLinqContext(4)
    got = WC.through_api(ctx, str(path))
    assert got == want
    assert ctx.__dict__.get("native_host_vertices", 0) >= 1
    # an opaque lambda in the pipeline takes the record path and still agrees
    q = ctx.from_store("file://" + str(path)).select_many(lambda l: [w for w in l.split(" ")]).group_by(lambda w: w).select(
        lambda g: D.Pair(g.key, g.count()))
    assert dict(q.to_list()) == want
