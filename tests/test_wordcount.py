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
