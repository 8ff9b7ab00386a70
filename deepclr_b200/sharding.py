"""Data-parallel partition of independent registration pairs over ranks (one process per GPU, no collective on the
data path: pairs never interact -- SURVEY.md section 8e)."""
from typing import Tuple

import numpy as np


def pair_range(num_pairs: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous [begin, end) slice of the pair index range owned by `rank`; sizes differ by at most one."""
    if not 0 <= rank < world_size:
        raise ValueError('rank out of range')
    base, extra = divmod(num_pairs, world_size)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def shard_batch(x: np.ndarray, world_size: int, rank: int) -> np.ndarray:
    """x [2B,N,C] in the reference layout (templates in rows 0..B-1, sources in rows B..2B-1,
    deepclr/data/build.py:82) -> the local [2*B_rank, N, C] batch in the same layout.  Shards PAIRS, not rows."""
    if x.shape[0] % 2:
        raise ValueError('batch must contain templates then sources')
    B = x.shape[0] // 2
    b, e = pair_range(B, world_size, rank)
    return np.concatenate([x[b:e], x[B + b:B + e]], axis=0)
