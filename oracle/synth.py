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


_RMAT_A, _RMAT_AB, _RMAT_ABC = 2448131358, 3264175144, 4080218931  # floor(0.57 / 0.76 / 0.95 * 2**32)


def synth_rmat(seed: int, first: int, n: int, M: int, N: int, with_value: bool = True):
    """The skewed variant of SURVEY.md 8d: R-MAT (a, b, c, d) = (0.57, 0.19, 0.19, 0.05) on the same counter
    stream (paddle_sparse_b200/csrc/synth.cu: synth_rmat_kernel).  Level l of entry c reads the high (l even) or
    low (l odd) 32 bits of splitmix64(seed + 16 c + l // 2); values as in synth_coo."""
    levels = max(int(max(M, N) - 1).bit_length(), 0)
    with np.errstate(over="ignore"):
        c = np.arange(first, first + n, dtype=_U64)
        r = np.zeros(n, dtype=_U64)
        q = np.zeros(n, dtype=_U64)
        h = None
        for lvl in range(levels):
            if lvl % 2 == 0:
                h = splitmix64(_U64(seed) + _U64(16) * c + _U64(lvl // 2))
            u = (h & _U64(0xFFFFFFFF)) if lvl % 2 else (h >> _U64(32))
            rb = (u >= _U64(_RMAT_AB)).astype(_U64)
            cb = (((u >= _U64(_RMAT_A)) & (u < _U64(_RMAT_AB))) | (u >= _U64(_RMAT_ABC))).astype(_U64)
            r = (r << _U64(1)) | rb
            q = (q << _U64(1)) | cb
        row = (r % _U64(M)).astype(np.int64)
        col = (q % _U64(N)).astype(np.int64)
        value = None
        if with_value:
            g = splitmix64((_U64(seed) ^ _GOLD) + c)
            value = ((g >> _U64(40)).astype(np.float32) * np.float32(2.0 ** -24)).astype(np.float32)
    return row, col, value
