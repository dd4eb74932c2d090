// Counter-based RNG streams of the B200 core.  The reference seeds rand::StdRng from OS
// entropy per image row / per photon (src/renderer.rs:163, src/photon.rs:670) and is therefore
// irreproducible; here every camera path (pixel, sample) and every photon owns a splitmix64
// stream so that results do not depend on scheduling or on the number of GPUs.
#pragma once
#include "common.cuh"

namespace rptx {

constexpr uint64_t kRngGamma = 0x9E3779B97F4A7C15ull;
enum RngDomain : uint64_t { kDomainCamera = 0, kDomainPhoton = 1, kDomainGather = 2, kDomainBuild = 3, kDomainThin = 4 };

RPTX_HD uint64_t mix64(uint64_t z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27; z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

struct Rng {
    uint64_t state;
    RPTX_HD static Rng make(uint64_t seed, uint64_t domain, uint64_t a, uint64_t b) {
        Rng r;
        uint64_t s = mix64(seed + (domain + 1) * kRngGamma);
        s = mix64(s + a);
        r.state = mix64(s ^ b);
        return r;
    }
    RPTX_HD uint64_t next_u64() {
        state += kRngGamma;
        return mix64(state);
    }
    // rng.gen::<f64>() in [0,1)
    RPTX_HD double gen() { return (double)(next_u64() >> 11) * (1.0 / 9007199254740992.0); }
    // rng.gen_range(lo..hi)
    RPTX_HD double gen_range(double lo, double hi) { return lo + (hi - lo) * gen(); }
    // rng.sample(Uniform::from(0..n))
    RPTX_HD uint64_t gen_index(uint64_t n) {
        uint64_t i = (uint64_t)(gen() * (double)n);
        return i < n ? i : n - 1;
    }
    // rand_distr::UnitDisc
    RPTX_HD void unit_disc(double* x, double* y) {
        for (;;) {
            double a = gen_range(-1.0, 1.0);
            double b = gen_range(-1.0, 1.0);
            if (a * a + b * b <= 1.0) { *x = a; *y = b; return; }
        }
    }
    // rand_distr::UnitCircle
    RPTX_HD void unit_circle(double* x, double* y) {
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

}  // namespace rptx
