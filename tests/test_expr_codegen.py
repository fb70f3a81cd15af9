"""Expression IR: tracing, host evaluation, and NVRTC code generation (compiles for sm_100a
without a GPU; execution is covered by the gpu-marked tests)."""
import numpy as np
import pytest
import torch

from dryad_b200.expr import col, eval_numpy, eval_row, lit, trace, where
from dryad_b200.expr import codegen


def test_trace_and_host_eval_agree_with_python():
    f = lambda r: {"a": r.x * 2 + r.y, "b": (r.x % 3 == 0) & (r.y > 1.5), "c": where(r.x > 2, r.x, -r.x)}
    ir = trace(f, {"x": "i", "y": "f"})
    assert set(ir) == {"a", "b", "c"} and ir["a"].kind == "f" and ir["b"].kind == "b"
    xs, ys = np.arange(-5, 6), np.linspace(0, 3, 11)
    got = eval_numpy(ir["a"], {"x": xs, "y": ys}, 11)
    assert np.allclose(got, xs * 2 + ys)
    for x, y in zip(xs.tolist(), ys.tolist()):
        assert eval_row(ir["c"], {"x": x, "y": y}) == (x if x > 2 else -x)
    # the lambdas are Python: // floors and % takes the divisor's sign, on every backend
    assert eval_row(col("x") // 2, {"x": -7}) == -4 and eval_row(col("x") % 3, {"x": -7}) == 2
    assert eval_row(col("x") % -3, {"x": 7}) == -2 and eval_row(col("y", "f") // 2, {"y": -7.5}) == -4.0
    assert trace(lambda r: r.x if r.x > 0 else 0, {"x": "i"}) is None          # data-dependent control flow


@pytest.mark.skipif(not codegen.available(), reason="cuda-python/nvrtc not importable")
def test_generated_cuda_compiles_for_sm100a():
    from cuda.bindings import nvrtc
    exprs = {"a": col("x") * 3 + 1, "b": col("y", "f") * 2.5 - col("x"), "h": col("x").hash(),
             "s": where((col("x") % 2 == 0) | (col("y", "f") < 0.5), col("x") // 3, abs(col("x"))),
             "q": (col("y", "f").sqrt() + col("y", "f").floor()).maximum(1.0)}
    src = codegen._source(exprs, (col("x") > 3) & ~(col("y", "f") >= 2.0), ["x", "y"], [torch.int64, torch.float64])
    err, prog = nvrtc.nvrtcCreateProgram(src.encode(), b"t.cu", 0, [], [])
    opts = [b"--gpu-architecture=sm_100a"]
    (err,) = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    assert int(err) == 0, src
    assert "dryad_fused" in src and src.count("in0[i]") == 1      # every input column is loaded exactly once


@pytest.mark.skipif(not codegen.available(), reason="cuda-python/nvrtc not importable")
def test_fused_probe_gather_reduce_source_compiles():
    """A join probe, a side-array gather and in-kernel reductions are part of the generated
    pipeline: one probe per row even when the node is used twice (CSE), per-block partials out."""
    from cuda.bindings import nvrtc
    from dryad_b200.expr import gather, probe
    hit = probe("dim", col("k"))
    keep = (col("v") >= 500) & (hit >= 0)
    src = codegen._source({"r": where(keep, col("v") * gather("attr", hit), 0), "f": gather("w", hit, "f") * 2.0}, keep,
                          ["k", "v"], [torch.int64, torch.int64], tables=["dim"],
                          arrays=[("attr", torch.int64), ("w", torch.float64)], reduce={"r": "sum", "f": "max"}, count_flags=True)
    assert src.count("dryad_probe<false>(tk0") == 1
    err, prog = nvrtc.nvrtcCreateProgram(src.encode(), b"t.cu", 0, [], [])
    opts = [b"--gpu-architecture=sm_100a"]
    (err,) = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    assert int(err) == 0, src
