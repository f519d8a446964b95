"""Test-side restatement of the reference's seeded draw stream, independent of the product's csrc/mt19937.h (numpy's MT19937
bit generator supplies the raw 32-bit outputs): MathNet `MersenneTwister(seed)` + `Normal.Samples` + Cmdty.Core's
`MersenneTwisterGenerator(seed, antithetic)` as used by `MultiFactorSpotPriceSimulator.Simulate` — see the header of
storage_b200/csrc/mt19937.h for what is known and how it was pinned (scripts/mt19937_search.py found the draw order)."""
import math

import numpy as np

INV_PI = 0.31830988618379067154


class MathNetMersenneTwister:
    def __init__(self, seed):
        self._bg = np.random.MT19937()
        self._bg._legacy_seeding(int(seed) & 0xFFFFFFFF)   # init_genrand(seed)
        self._buf, self._pos = np.zeros(0, dtype=np.uint64), 0

    def next_double(self):   # genrand_int32() / 2^32
        if self._pos >= len(self._buf):
            self._buf, self._pos = self._bg.random_raw(4096), 0
        v = float(self._buf[self._pos]) * (1.0 / 4294967296.0)
        self._pos += 1
        return v


def _polar(a, b):
    v1, v2 = 2.0 * a - 1.0, 2.0 * b - 1.0
    r = v1 * v1 + v2 * v2
    if r >= 1.0 or r == 0.0:
        return None
    fac = math.sqrt(-2.0 * math.log(r) / r)
    return v1 * fac, v2 * fac


def normal_samples(rnd, n):
    """MathNet Normal.Samples(rnd, values[n], 0, 1)."""
    out = []
    m = int(math.ceil(n * 4 * INV_PI))
    if m % 2:
        m += 1
    u = [rnd.next_double() for _ in range(m)]
    for i in range(0, m, 2):
        p = _polar(u[i], u[i + 1])
        if p is None:
            continue
        for v in p:
            out.append(v)
            if len(out) == n:
                return out
    while len(out) < n:
        p = _polar(rnd.next_double(), rnd.next_double())
        if p is None:
            continue
        for v in p:
            out.append(v)
            if len(out) == n:
                return out
    return out


def path_normals(seed, antithetic, n_sims, n_steps, n_factors, rnd=None):
    """[n_steps][n_factors][n_sims]: one generator call per path fills [step][factor]; with antithetic every second call is the
    negation of the call before it.  Returns (normals, generator) so that a second set can continue the sequence."""
    rnd = rnd or MathNetMersenneTwister(seed)
    per = n_steps * n_factors
    z = np.empty((n_sims, per))
    n = 0
    while n < n_sims:
        fresh = normal_samples(rnd, per)
        z[n] = fresh
        n += 1
        if antithetic and n < n_sims:
            z[n] = -np.asarray(fresh)
            n += 1
    return np.ascontiguousarray(np.transpose(z.reshape(n_sims, n_steps, n_factors), (1, 2, 0))), rnd
