"""Range-partition sampler semantics (LinqToDryad/DryadLinqSampler.cs:42-106, 139-246) and the
left-most bucket rule (LinqToDryad/DryadLinqUtil.cs:112-139)."""
from __future__ import annotations

import random
from typing import Callable, List, Sequence

SAMPLE_RATE = 0.001
MAX_SECOND_PHASE_SAMPLES = 1024 * 1024


def phase1_sample(keys: Sequence, seed: int, rate: float = SAMPLE_RATE) -> List:
    """Bernoulli(rate) sample seeded by the partition id; all keys if fewer than 10 were drawn."""
    rnd = random.Random(seed)
    samples = [k for k in keys if rnd.random() < rate]
    if len(samples) < 10:
        return list(keys)
    return samples


def select_separators(sorted_samples: Sequence, pcount: int, descending: bool = False) -> List:
    """`pcount-1` separators from ASCENDING-sorted samples (RangeSamplerCore)."""
    cnt = len(sorted_samples)
    if pcount <= 1:
        return []
    if cnt == 0:
        return []
    out = []
    if cnt < pcount:
        if descending:
            out = [sorted_samples[i] for i in range(cnt - 1, -1, -1)]
            out += [sorted_samples[0]] * (pcount - 1 - cnt)
        else:
            out = list(sorted_samples[:cnt])
            out += [sorted_samples[cnt - 1]] * (pcount - 1 - cnt)
        return out[:pcount - 1]
    intv = cnt // pcount
    if descending:
        return [sorted_samples[cnt - intv * (i + 1)] for i in range(pcount - 1)]
    return [sorted_samples[intv * (i + 1)] for i in range(pcount - 1)]


def range_separators(all_keys_by_partition: Sequence[Sequence], pcount: int, descending: bool = False,
                     key_fn: Callable = lambda x: x, rate: float = SAMPLE_RATE) -> List:
    samples: List = []
    for pid, part in enumerate(all_keys_by_partition):
        samples.extend(phase1_sample([key_fn(x) for x in part], pid, rate))
    if len(samples) > MAX_SECOND_PHASE_SAMPLES:
        r = random.Random(314159)
        res = samples[:MAX_SECOND_PHASE_SAMPLES]
        for i in range(MAX_SECOND_PHASE_SAMPLES, len(samples)):
            j = r.randrange(i)
            if j < MAX_SECOND_PHASE_SAMPLES:
                res[j] = samples[i]
        samples = res
    samples.sort()
    return select_separators(samples, pcount, descending)


def range_bucket(key, separators: Sequence, descending: bool = False) -> int:
    """Left-most rule: ascending -> number of separators strictly below key; descending ->
    number strictly above (DryadLinqUtil.BinarySearch)."""
    lo, hi = 0, len(separators)
    while lo < hi:
        mid = (lo + hi) // 2
        s = separators[mid]
        right = (s > key) if descending else (s < key)
        if right:
            lo = mid + 1
        else:
            hi = mid
    return lo
