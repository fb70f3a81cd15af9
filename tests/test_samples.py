"""The samples run as programs (reference: samples/WordCount.cs.pp and the DryadLINQ sample programs)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    r = subprocess.run([sys.executable, *args], cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    return r.stdout


@pytest.mark.parametrize("mode", ["local", "debug", "cluster"])
def test_wordcount_sample(mode):
    out = _run("samples/wordcount.py", "--mode", mode)
    top = out.splitlines()[0].split()
    assert top == ["460", "the"]                       # 2 x 200 + 1 x 60: "the" is the most frequent word of the built-in text


@pytest.mark.parametrize("mode", ["local", "cluster"])
def test_kmeans_sample_converges(mode):
    out = _run("samples/kmeans.py", "--mode", mode, "--points", "400")
    assert "4 centroids" in out


@pytest.mark.parametrize("mode", ["local", "debug", "cluster"])
def test_sales_report_sample(mode):
    out = _run("samples/sales_report.py", "--mode", mode, "--rows", "2000")
    assert out.splitlines()[0].split() == ["category", "revenue", "orders"] and len(out.splitlines()) >= 9
