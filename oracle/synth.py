"""TEST INFRASTRUCTURE ONLY (oracle).

NumPy restatement of the deterministic synthetic COO generator of SURVEY.md 8d — the same counter
hash the device generator computes (paddle_sparse_b200/csrc/synth.cu), so the CPU oracle and the
GPU path can be fed identical inputs without moving them between host and device:

    h = splitmix64(seed + i); row = (h >> 32) % M; col = (h & 0xffffffff) % N
    value = fp32((splitmix64((seed ^ 0x9e3779b97f4a7c15) + i) >> 40) * 2**-24)
"""
import numpy as np

_U64 = np.uint64
_GOLD = _U64(0x9E3779B97F4A7C15)


def splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        x = x.astype(_U64) + _GOLD
        x = (x ^ (x >> _U64(30))) * _U64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> _U64(27))) * _U64(0x94D049BB133111EB)
        return x ^ (x >> _U64(31))


def synth_coo(seed: int, first: int, n: int, M: int, N: int, with_value: bool = True):
    """Entries [first, first + n) of the stream: (row int64, col int64, value float32 | None)."""
    with np.errstate(over="ignore"):
        c = np.arange(first, first + n, dtype=_U64)
        h = splitmix64(_U64(seed) + c)
        row = ((h >> _U64(32)) % _U64(M)).astype(np.int64)
        col = ((h & _U64(0xFFFFFFFF)) % _U64(N)).astype(np.int64)
        value = None
        if with_value:
            g = splitmix64((_U64(seed) ^ _GOLD) + c)
            value = ((g >> _U64(40)).astype(np.float32) * np.float32(2.0 ** -24)).astype(np.float32)
    return row, col, value
