"""Comparison baselines: a straightforward PyTorch + NCCL (+ cuBLAS) implementation of the same
queries on the same box (BASELINE.json: "a path that only calls NCCL for the shuffle, aggregation
or broadcast is the baseline, not the product")."""
