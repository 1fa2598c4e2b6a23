"""Host-side minimal-sample generator: a numpy restatement of the reference's seeded
``UniformSampling`` (RansacLib/sampling.h:47-135) so that batches of K samples are the
SAME index sets the reference's sequential loop would draw (the stream is a pure
function of (seed, N, sample_size): the sampler owns a private std::mt19937,
sampling.h:49-55,126).

Ingredients: ``std::mt19937::seed(s)`` is the classic init_genrand (numpy's legacy
seeding); libstdc++'s ``uniform_int_distribution<int>(0, N-1)`` on a 32-bit engine is
Lemire's multiply-shift with rejection (bits/uniform_int_dist.h, GCC >= 11).
"""
from __future__ import annotations

import math

import numpy as np


class Mt19937:
    """Raw 32-bit stream of std::mt19937 seeded with ``seed``."""

    def __init__(self, seed: int):
        self._bg = np.random.MT19937()
        self._bg._legacy_seeding(int(seed) & 0xFFFFFFFF)

    def raw(self, n: int) -> np.ndarray:
        return self._bg.random_raw(n).astype(np.uint64)


def _lemire(raw: np.ndarray, rng_range: int):
    """(values, accepted) for a batch of raw draws and range = b - a + 1."""
    prod = raw * np.uint64(rng_range)
    low = prod & np.uint64(0xFFFFFFFF)
    thresh = np.uint64(((1 << 32) - rng_range) % rng_range)
    return (prod >> np.uint64(32)).astype(np.int64), low >= thresh


class UniformSampling:
    def __init__(self, seed: int, num_data: int, sample_size: int):
        self.n, self.s = int(num_data), int(sample_size)
        self._mt = Mt19937(seed)
        self._buf = np.empty(0, np.int64)
        # sampling.h:69-78
        self.draw = (self.n / (self.n - self.s)) < math.e if self.n > self.s else False

    def _values(self, count: int) -> np.ndarray:
        """`count` accepted uniform ints in [0, n-1], in stream order."""
        while len(self._buf) < count:
            v, ok = _lemire(self._mt.raw(max(4096, 2 * count)), self.n)
            self._buf = np.concatenate([self._buf, v[ok]])
        out, self._buf = self._buf[:count], self._buf[count:]
        return out

    def sample(self, k: int) -> np.ndarray:
        """K consecutive minimal samples, shape (k, s), int32."""
        if not self.draw:
            return self._sample_shuffle(k)
        out = np.empty((k, self.s), np.int32)
        for i in range(k):  # rejection of duplicates is sequential by nature (sampling.h:81-97)
            row = []
            while len(row) < self.s:
                v = int(self._values(1)[0])
                if v not in row:
                    row.append(v)
            out[i] = row
        return out

    def sample_fast(self, k: int) -> np.ndarray:
        """Vectorised path for N >> s: draws k*s values at once and falls back to the
        sequential path only from the first row that contains a duplicate."""
        if not self.draw:
            return self._sample_shuffle(k)
        out = np.empty((k, self.s), np.int32)
        done = 0
        while done < k:
            need = k - done
            vals = self._values(need * self.s)
            rows = vals.reshape(need, self.s)
            srt = np.sort(rows, axis=1)
            dup = (srt[:, 1:] == srt[:, :-1]).any(axis=1)
            if not dup.any():
                out[done:] = rows
                return out
            first = int(np.argmax(dup))
            out[done:done + first] = rows[:first]
            done += first
            # give the unused values back and resolve this one row sequentially
            self._buf = np.concatenate([vals[first * self.s:], self._buf])
            out[done:done + 1] = self.sample(1)
            done += 1
        return out

    def _sample_shuffle(self, k: int) -> np.ndarray:
        # sampling.h:102-135: Fisher-Yates over iota(n) with a fresh distribution per position
        out = np.empty((k, self.s), np.int32)
        for r in range(k):
            a = list(range(self.n))
            if self.s != self.n:
                for i in range(self.n - 1):
                    rr = self.n - i
                    while True:
                        v, ok = _lemire(self._mt.raw(1), rr)
                        if ok[0]:
                            break
                    j = i + int(v[0])
                    a[i], a[j] = a[j], a[i]
            out[r] = a[:self.s]
        return out
