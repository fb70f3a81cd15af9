"""WordCount (BASELINE.json config 1; samples/WordCount.cs.pp:72-75):
`lines.SelectMany(l => l.Split(' ')).GroupBy(w => w).Select(g => new Pair(g.Key, g.Count()))`.

Three implementations of the same query:
  * `through_api(ctx, path)`   the public query API (host backend = the LocalJobSubmission-style
                               CPU mode; LocalDebug = oracle);
  * `gpu_wordcount(text)`      the fused GPU path: tokenizer kernel (K13) -> 64-bit token hashes
                               (K14) -> hash aggregate with Count + Min(token index) lanes (K6) ->
                               full-key verification of every token against its group's
                               representative (hash-collision check);
  * `cpu_baseline(text)`       multi-threaded C++ (csrc/host/oracle_cpu.cpp), the timed CPU number.
"""
from __future__ import annotations

import ctypes as C
import random
from typing import Dict

import torch

VOCAB = 50_000


def generate_text(n_lines: int, words_per_line: int = 12, seed: int = 1, vocab: int = VOCAB) -> bytes:
    rnd = random.Random(seed)
    words = [("w%x" % (i * 2654435761 % (1 << 28))).encode() for i in range(vocab)]
    # Zipf-like skew: low ranks are much more frequent
    lines = []
    for _ in range(n_lines):
        ws = [words[min(vocab - 1, int(rnd.paretovariate(1.2)) - 1)] for _ in range(words_per_line)]
        lines.append(b" ".join(ws))
    return b"\n".join(lines) + b"\n"


def through_api(ctx, path: str) -> Dict[str, int]:
    import dryad_b200 as D
    q = (ctx.from_store("file://" + path).select_many(lambda l: l.split(" ")).group_by(lambda w: w)
         .select(lambda g: D.Pair(g.key, g.count())))
    return dict(q.to_list())


def gpu_wordcount(text: torch.Tensor, keep_empty: bool = True):
    """text: uint8 CUDA tensor.  Returns (rep_start int64[g], rep_len int32[g], counts int64[g],
    total_tokens, collisions) — collisions > 0 means two different words shared a 64-bit hash and
    the caller must fall back to the host path."""
    from ..ops import core as K
    starts, lens, hashes = K.tokenize(text, " ", keep_empty)
    m = starts.numel()
    if m == 0:
        z = torch.empty(0, dtype=torch.int64, device=text.device)
        return z, torch.empty(0, dtype=torch.int32, device=text.device), z, 0, 0
    tok_idx = torch.arange(m, dtype=torch.int64, device=text.device)
    tbl = K.HashAggTable(min(2 * m, 1 << 24), ["count", "min_i64"], text.device)
    tbl.update(hashes, [None, tok_idx])
    gk, gl = tbl.extract()
    rep_tok = gl[1]
    rep_start, rep_len = starts[rep_tok], lens[rep_tok]
    # verify: every token equals the representative of its hash group
    jt = K.JoinTable(gk.contiguous())
    rows, matched = None, None
    from .. import _native as N
    rows = torch.empty(m, dtype=torch.int32, device=text.device)
    matched = torch.empty(m, dtype=torch.uint8, device=text.device)
    N.check(N.kernels().dryad_join_probe_unique(K.ptr(jt.keys), K.ptr(jt.head), jt.cap, K.ptr(hashes), m, K.ptr(rows),
                                                K.ptr(matched), K.stream()), "join_probe_unique")
    K._count()
    tok_rep_start = K.gather_rows(rep_start.contiguous(), rows)
    tok_rep_len = K.gather_rows(rep_len.contiguous(), rows)
    collisions = K.tokenize_verify(text, starts, lens, tok_rep_start, tok_rep_len)
    return rep_start, rep_len, gl[0], m, collisions


def gpu_result_to_dict(text_bytes: bytes, rep_start, rep_len, counts) -> Dict[bytes, int]:
    s, l, c = rep_start.cpu().tolist(), rep_len.cpu().tolist(), counts.cpu().tolist()
    return {text_bytes[a:a + b]: n for a, b, n in zip(s, l, c)}


def cpu_baseline(text: bytes, threads: int = 0, keep_empty: bool = True):
    import os
    from .. import _native as N
    tot = C.c_int64()
    groups = N.host().cpu_wordcount(text, len(text), threads or (os.cpu_count() or 1), int(keep_empty), C.byref(tot))
    return groups, tot.value
