"""CPU restatement of the pre-processing either side of the forward.  TEST INFRASTRUCTURE ONLY (checker for
dclr_filter_clouds / dclr_subsample_clouds; never imported by the product path).

systematic_erasing / range_selection / truncate_dimension follow deepclr/data/transforms/transforms.py:244-268, 90-113,
271-283 line by line and are pinned against the reference classes themselves by oracle/gen_golden.py
(tests/golden/preproc_small.npz).  The reference's random subsample (deepclr/models/base.py:130-135,
transforms.py:128-131) draws from torch / numpy global generators, so only its PROPERTIES are part of the contract
(n distinct rows of the cloud, equal sizes untouched); feistel_perm restates the keyed bijection of
deepclr_b200/csrc/preproc.cu so that the chosen rows can also be compared index by index.
"""
import numpy as np

M32 = 0xFFFFFFFF


def systematic_erasing(cloud: np.ndarray, nth: int, start: int = 0) -> np.ndarray:
    if nth == 1:
        return cloud
    return cloud[start::nth, :]


def range_selection(cloud: np.ndarray, min_range: float, max_range: float, dim: int = 3) -> np.ndarray:
    if min_range == 0.0 and np.isinf(max_range):
        return cloud
    cloud_max = np.max(np.abs(cloud[:, :(dim - 1)]), axis=1)
    inliers = (cloud_max >= np.float32(min_range)) & (cloud_max <= np.float32(max_range))
    return cloud[inliers, :]


def truncate_dimension(cloud: np.ndarray, input_dim: int) -> np.ndarray:
    return cloud[:, :input_dim]


def filter_cloud(cloud, nth=1, start=0, min_range=0.0, max_range=np.inf, out_dim=None):
    """The chain of deepclr/data/transforms/build.py:23-24,37-38 (+ TruncateDimension)."""
    c = range_selection(systematic_erasing(cloud, nth, start), min_range, max_range)
    return c if out_dim is None else truncate_dimension(c, out_dim)


def _mix32(x: int) -> int:
    x &= M32
    x ^= x >> 16
    x = (x * 0x7feb352d) & M32
    x ^= x >> 15
    x = (x * 0x846ca68b) & M32
    x ^= x >> 16
    return x


def cloud_key(seed: int, b: int) -> int:
    lo, hi = seed & M32, (seed >> 32) & M32
    return _mix32(lo ^ _mix32((hi + b) & M32))


def feistel_perm(j: int, n: int, key: int) -> int:
    if n <= 1:
        return 0
    bits = max((n - 1).bit_length(), 2)
    hb = bits >> 1
    ha = bits - hb
    ma, mb = (1 << ha) - 1, (1 << hb) - 1
    x = j
    while True:
        L, R = x >> hb, x & mb
        for r in range(4):
            F = _mix32(R ^ ((key + 0x9e3779b9 * (r + 1)) & M32)) & (mb if r & 1 else ma)
            L, R = R, L ^ F
        x = (L << hb) | R
        if x < n:
            return x


def _mix32_np(x):
    x = x & M32
    x ^= x >> 16
    x = (x * 0x7feb352d) & M32
    x ^= x >> 15
    x = (x * 0x846ca68b) & M32
    x ^= x >> 16
    return x


def feistel_perm_np(j: np.ndarray, n: int, key: int) -> np.ndarray:
    """Vectorised feistel_perm (uint64 arithmetic) for full-size checks."""
    j = np.asarray(j, dtype=np.uint64)
    if n <= 1:
        return np.zeros(j.shape, np.int64)
    bits = max((n - 1).bit_length(), 2)
    hb = bits >> 1
    ha = bits - hb
    ma, mb = np.uint64((1 << ha) - 1), np.uint64((1 << hb) - 1)
    x = j.copy()
    todo = np.ones(x.shape, bool)
    while todo.any():
        cur = x[todo]
        L, R = cur >> np.uint64(hb), cur & mb
        for r in range(4):
            F = _mix32_np(R ^ np.uint64((key + 0x9e3779b9 * (r + 1)) & M32)) & (mb if r & 1 else ma)
            L, R = R, L ^ F
        cur = (L << np.uint64(hb)) | R
        x[todo] = cur
        todo[todo] = cur >= n
    return x.astype(np.int64)


def subsample_rows(n: int, n_out: int, seed: int, b: int, ordered_if_equal: bool = True) -> np.ndarray:
    if n == 0:
        return np.full(n_out, -1, np.int32)
    if n == n_out and ordered_if_equal:
        return np.arange(n_out, dtype=np.int32)
    key = cloud_key(seed, b)
    return feistel_perm_np(np.arange(n_out) % n, n, key).astype(np.int32)
