"""The integer images of fp64 thresholds used by the grouping pass (csrc/kp_snapshot.cu: limit_cpu / limit_mem),
restated in Python with the same IEEE operations and checked against their definition by brute force:
M = the largest integer m with fl(m / 1000.0) <= T (resp. m * 2^-30 <= T), so that "request <= M" is the reference's
own `cpu_req <= free_cpu` compare (scheduler.py:232-233 operands, :742-753 request units) for every int64 request with
|x| < 2^53."""
import math

import numpy as np

EXACT_MAX = (1 << 53) - 1


def limit_cpu(T: float) -> int:
    if EXACT_MAX / 1000.0 <= T:
        return EXACT_MAX
    if not (-EXACT_MAX / 1000.0 <= T):
        return -EXACT_MAX - 1
    c = int(math.floor(T * 1000.0)) + 2
    c = min(c, EXACT_MAX)
    while c > -EXACT_MAX and c / 1000.0 > T:
        c -= 1
    return c


def limit_mem(T: float) -> int:
    g = math.floor(T * 1073741824.0) if math.isfinite(T * 1073741824.0) else T * 1073741824.0
    if g >= float(EXACT_MAX):
        return EXACT_MAX
    if g < -float(EXACT_MAX):
        return -EXACT_MAX - 1
    return int(g)


def thresholds():
    rng = np.random.default_rng(1)
    ts = [0.0, -0.0, 1e-300, -1e-300, 0.001, 0.0005, 1.0, 8.0, 0.32, 31.712, float("inf"), float("-inf"),
          EXACT_MAX / 1000.0, -EXACT_MAX / 1000.0, 9.1e12, -9.1e12, 9.007199254740991e12, 2.0 ** 43, 2.0 ** 42 + 0.5]
    for m in rng.integers(-10 ** 7, 10 ** 7, 400).tolist() + rng.integers(-EXACT_MAX, EXACT_MAX, 400).tolist():
        t = m / 1000.0                                   # thresholds that sit exactly on, just below and just above an image
        ts += [t, np.nextafter(t, -np.inf), np.nextafter(t, np.inf)]
    ts += (rng.standard_normal(400) * 10.0 ** rng.integers(-6, 13, 400)).tolist()
    return [float(t) for t in ts]


def test_limit_cpu_is_the_largest_request_that_fits():
    for T in thresholds():
        M = limit_cpu(T)
        assert -EXACT_MAX - 1 <= M <= EXACT_MAX
        if M >= -EXACT_MAX:
            assert M / 1000.0 <= T, (T, M)
        if M < EXACT_MAX:
            assert not ((M + 1) / 1000.0 <= T), (T, M)
        # monotone image: everything below M fits, everything above does not (spot checks around M)
        for d in (1, 2, 3, 1000, 10 ** 6):
            if M - d >= -EXACT_MAX:
                assert (M - d) / 1000.0 <= T
            if M + d <= EXACT_MAX:
                assert not ((M + d) / 1000.0 <= T)


def test_limit_mem_is_the_largest_request_that_fits():
    s = 2.0 ** -30
    for T in thresholds():
        M = limit_mem(T)
        if M >= -EXACT_MAX:
            assert M * s <= T, (T, M)
        if M < EXACT_MAX:
            assert not ((M + 1) * s <= T), (T, M)
