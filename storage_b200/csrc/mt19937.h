// mt19937.h — the reference's seeded draw stream, restated on the host.
//
// LsmcValuationParameters.cs:181-191 hands the simulator a Cmdty.Core `MersenneTwisterGenerator(seed, antithetic: true)`;
// Cmdty.Core (0.1.0-alpha2020101802) and MathNet.Numerics are not in the tree, so the stream below is their published /
// observable behaviour, and it is PINNED by the reference's own golden vectors (tests/test_reference_goldens.py):
// the twelve seeded spot prices of src/Cmdty.Storage.Python/tests/test_multi_factor.py:74-91 and the two seeded NPVs at
// :226 and :298 come out bit for bit (the NPVs through the CPU oracle's Householder-QR valuation).
//   * MathNet MersenneTwister(seed): init_genrand(seed) of Matsumoto & Nishimura's mt19937ar; NextDouble() = genrand_int32() / 2^32.
//   * MathNet Normal.Samples(rnd, values): n = ceil(len * 4 / pi) uniforms (made even) are drawn UP FRONT, consecutive pairs go
//     through Marsaglia's polar transform (v = 2u - 1, reject r >= 1 or r == 0, factor sqrt(-2 ln r / r)), both outputs used in
//     order; unused uniforms of the batch are dropped; if the batch runs dry, further pairs are drawn one at a time.
//   * MultiFactorSpotPriceSimulator.Simulate(numSims): one generator call per simulated path, filling [step][factor] (factor
//     fastest) normals.
//   * antithetic: every second call returns the negated values of the call before it, so paths 2m and 2m + 1 are partners.
//     (Which of the two common layouts Cmdty.Core uses — neighbours or halves — cannot be seen in any number the reference
//     asserts: its seeded NPVs are invariant under a permutation of the paths, and its one per-path golden vector is drawn
//     without antithetic.  Neighbours it is, like the device Philox stream.)
// Host only, serial by construction (one generator walks all paths): a compatibility stream for seeded reproduction, about
// 15 ns per normal; the device Philox stream stays the fast path.
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <vector>

namespace lsmc {

struct Mt19937 {
    uint32_t mt[624];
    int idx = 625;          // 625: not seeded
    int odd_call = 0;       // antithetic generator: the next call returns the negated cache
    std::vector<double> cache;

    void seed(uint32_t s) {
        mt[0] = s;
        for (int i = 1; i < 624; i++) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (uint32_t)i;
        idx = 624;
        odd_call = 0;
        cache.clear();
    }
    bool seeded() const { return idx <= 624; }
    uint32_t next_u32() {
        if (idx >= 624) {
            for (int k = 0; k < 624; k++) {
                const uint32_t y = (mt[k] & 0x80000000u) | (mt[(k + 1) % 624] & 0x7fffffffu);
                mt[k] = mt[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
            }
            idx = 0;
        }
        uint32_t y = mt[idx++];
        y ^= y >> 11;
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= y >> 18;
        return y;
    }
    double next_double() { return (double)next_u32() * (1.0 / 4294967296.0); }

    static bool polar(double a, double b, double& x, double& y) {
        const double v1 = 2.0 * a - 1.0, v2 = 2.0 * b - 1.0, r = v1 * v1 + v2 * v2;
        if (r >= 1.0 || r == 0.0) return false;
        const double fac = std::sqrt(-2.0 * std::log(r) / r);
        x = v1 * fac; y = v2 * fac;
        return true;
    }

    // MathNet Normal.Samples(rnd, values[0..n), 0, 1)
    void normal_samples(double* values, size_t n) {
        if (n == 0) return;
        size_t m = (size_t)std::ceil((double)n * 4.0 * 0.31830988618379067154);
        if (m & 1) m++;
        std::vector<double> u(m);
        for (size_t i = 0; i < m; i++) u[i] = next_double();
        size_t index = 0;
        double x, y;
        for (size_t i = 0; i < m && index < n; i += 2) {
            if (!polar(u[i], u[i + 1], x, y)) continue;
            values[index++] = x;
            if (index == n) return;
            values[index++] = y;
            if (index == n) return;
        }
        while (index < n) {
            const double a = next_double(), b = next_double();
            if (!polar(a, b, x, y)) continue;
            values[index++] = x;
            if (index == n) return;
            values[index++] = y;
        }
    }

    // One call of Cmdty.Core's generator: `per_call` standard normals, every second call the negation of the call before it
    void generate(double* out, size_t per_call, bool antithetic) {
        if (antithetic && odd_call) {
            for (size_t i = 0; i < per_call; i++) out[i] = -cache[i];
            odd_call = 0;
            return;
        }
        normal_samples(out, per_call);
        if (antithetic) { cache.assign(out, out + per_call); odd_call = 1; }
    }
};

// Normals of paths [first, first + count) of a run of generator calls that starts at the generator's current state, in the
// device layout [S][F][count] (one call per path fills [step][factor]); the generator is left after call `total` so that a
// following simulation set continues the sequence the way the reference's shared generator does.
inline void mt19937_path_normals(Mt19937& g, bool antithetic, int64_t total, int64_t first, int64_t count, int S, int F, double* out) {
    const size_t per = (size_t)S * F;
    std::vector<double> z(per);
    for (int64_t n = 0; n < total; n++) {
        g.generate(z.data(), per, antithetic);
        if (n < first || n >= first + count) continue;
        const int64_t col = n - first;
        for (size_t e = 0; e < per; e++) out[e * (size_t)count + (size_t)col] = z[e];   // e = step * F + factor
    }
}

}  // namespace lsmc
