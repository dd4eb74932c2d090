// ORACLE — TEST INFRASTRUCTURE ONLY (see vec.hpp header).
// Deterministic counter-based RNG replacing the reference's entropy-seeded
// StdRng (rand 0.8.3 ChaCha12; src/renderer.rs:163, src/photon.rs:670,699,710).
// The reference is irreproducible by construction, so only the
// DISTRIBUTIONS of gen / gen_range / UnitDisc / Uniform(0..n) are restated; the stream
// layout below is this project's own and is mirrored bit-for-bit by the CUDA kernels
// (rpt_b200/csrc/rng.cuh) so that same-seed images can be compared sample by sample.
#pragma once
#include <cstdint>

namespace orc {

constexpr uint64_t RNG_GAMMA = 0x9E3779B97F4A7C15ull;

inline uint64_t mix64(uint64_t z) {  // splitmix64 finaliser
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

enum RngDomain : uint64_t { DOMAIN_CAMERA = 0, DOMAIN_PHOTON = 1, DOMAIN_GATHER = 2, DOMAIN_BUILD = 3 };

struct Rng {
    uint64_t state;
    // stream (seed, domain, a, b): camera paths use a = pixel index (y*w+x), b = global sample
    // index; photons use a = photon index, b = 0.
    Rng(uint64_t seed, uint64_t domain, uint64_t a, uint64_t b) {
        uint64_t s = mix64(seed + (domain + 1) * RNG_GAMMA);
        s = mix64(s + a);
        state = mix64(s ^ b);
    }
    uint64_t next_u64() {
        state += RNG_GAMMA;
        return mix64(state);
    }
    // rng.gen::<f64>(): 53 random bits -> [0,1)
    double gen() { return (double)(next_u64() >> 11) * (1.0 / 9007199254740992.0); }
    // rng.gen_range(lo..hi) for f64 (distribution only: lo + (hi-lo)*U[0,1))
    double gen_range(double lo, double hi) { return lo + (hi - lo) * gen(); }
    // rng.sample(Uniform::from(0..n)) (distribution only)
    uint64_t gen_index(uint64_t n) {
        uint64_t i = (uint64_t)(gen() * (double)n);
        return i < n ? i : n - 1;
    }
    // rand_distr::UnitDisc: rejection from the square
    void unit_disc(double* x, double* y) {
        for (;;) {
            double a = gen_range(-1.0, 1.0);
            double b = gen_range(-1.0, 1.0);
            if (a * a + b * b <= 1.0) { *x = a; *y = b; return; }
        }
    }
    // rand_distr::UnitCircle (0.4.0): x1, x2 uniform in (-1,1) until x1^2 + x2^2 < 1, then
    // ((x1^2 - x2^2) / sum, 2 x1 x2 / sum)
    void unit_circle(double* x, double* y) {
        double x1, x2, sum;
        do {
            x1 = gen_range(-1.0, 1.0);
            x2 = gen_range(-1.0, 1.0);
            sum = x1 * x1 + x2 * x2;
        } while (sum >= 1.0);
        *x = (x1 * x1 - x2 * x2) / sum;
        *y = 2.0 * x1 * x2 / sum;
    }
};

}  // namespace orc
