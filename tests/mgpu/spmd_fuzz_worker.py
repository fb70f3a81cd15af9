"""Random queries through the SPMD executor (gloo, host backend) against the LocalDebug oracle.
Launched by tests/test_spmd_cpu.py under torchrun; `from_enumerable(data, n)` is mapped to 1 or WORLD
partitions (what SPMD execution supports), results are gathered in rank order before comparing."""
import os
import random
import sys

import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
from fuzz_columnar import build_columnar  # noqa: E402
from fuzz_operators import build_operators  # noqa: E402
from fuzz_queries import build, rand_pipeline  # noqa: E402
from test_fuzz_differential import _norm_typed  # noqa: E402
from dryad_b200.api.context import DryadLinqContext  # noqa: E402


def ev(make, seed, ctx):
    rng = random.Random(seed)
    try:
        r = make(rng, ctx)
        q, ordered = (r[0], r[2]) if make is rand_pipeline else (r[1], r[2])
        if q is None:
            return None
        if isinstance(q, tuple):
            return ordered, q[1]()
        return ordered, list(q)
    except Exception as ex:  # noqa: BLE001
        return True, ("ERR", getattr(ex, "error_code", type(ex).__name__), str(ex)[:160])


def main():
    first, last = int(sys.argv[1]), int(sys.argv[2])
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    ctx = DryadLinqContext(platform="B200", device="cpu")
    plain = ctx.from_enumerable
    ctx.from_enumerable = lambda data, n=None: plain(data, 1 if n == 1 else world)
    bad = checked = unsupported = 0
    faults = os.environ.get("FUZZ_FAULTS") == "1"
    state = {"seed": 0}
    if faults:
        # a transient failure on ONE rank in the first attempt of some stages: every rank must re-run the stage
        def inject(node, r, version):
            if version == 0 and (state["seed"] * 31 + node.uid * 7) % 4 == 0 and r == (state["seed"] + node.uid) % world:
                raise RuntimeError(f"injected transient failure (seed {state['seed']}, stage {node.uid}, rank {r})")
        ctx.fault_injector = inject
    for seed in range(first, last):
        state["seed"] = seed
        for make in (build_columnar, rand_pipeline, build, build_operators):
            got = ev(make, seed, ctx)
            box = [None] * world
            dist.all_gather_object(box, got)
            if all(b is not None and isinstance(b[1], list) for b in box):
                got = (box[0][0], [x for b in box for x in b[1]])          # partitions in rank order
            if rank != 0 or got is None:
                continue
            if isinstance(got[1], tuple) and got[1][:1] == ("ERR",) and "1 or world partitions" in got[1][2]:
                unsupported += 1                                            # explicit partition counts SPMD does not take
                continue
            want = ev(make, seed, DryadLinqContext(3, local_debug=True))
            checked += 1
            if _norm_typed(got[1], got[0]) != _norm_typed(want[1], want[0]):
                bad += 1
                print("MISMATCH", make.__name__, seed, "\n  want", str(want[1])[:200], "\n  got ", str(got[1])[:200], flush=True)
    dist.barrier()
    if rank == 0:
        print(f"spmd fuzz: checked {checked} unsupported {unsupported} mismatches {bad}", flush=True)
    dist.destroy_process_group()
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
