// loader.cpp -- see loader.h
#include "loader.h"

#include <cmath>

namespace opal_driver {

namespace {
inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
}

Xoshiro256StarStar::Xoshiro256StarStar(uint64_t seed) {
    for (int i = 0; i < 4; i++) {                       // SplitMix64
        seed += 0x9e3779b97f4a7c15ULL;
        uint64_t z = seed;
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
        s_[i] = z ^ (z >> 31);
    }
}

uint64_t Xoshiro256StarStar::next_u64() {
    const uint64_t result = rotl(s_[1] * 5, 7) * 9;
    const uint64_t t = s_[1] << 17;
    s_[2] ^= s_[0];
    s_[3] ^= s_[1];
    s_[1] ^= s_[2];
    s_[0] ^= s_[3];
    s_[2] ^= t;
    s_[3] = rotl(s_[3], 45);
    return result;
}

double Xoshiro256StarStar::standard_normal() {
    const double u1 = 1.0 - gen_f64();                  // (0,1]
    const double u2 = gen_f64();
    return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586476925 * u2);
}

double Xoshiro256StarStar::exp1() { return -std::log1p(-gen_f64()); }

ParticleArrays load_population(size_t npc, const std::function<double(double)>& number_density,
                               const std::function<double(double, double, double)>& ux, const std::function<double(double, double, double)>& uy,
                               const std::function<double(double, double, double)>& uz, int64_t ncells, double slab_xmin, double dx,
                               Xoshiro256StarStar& rng, bool with_tau) {
    ParticleArrays p;
    for (int64_t c = 0; c < ncells; c++) {                               // mod.rs:179
        const double xc = slab_xmin + ((double)c + 0.5) * dx;            // mod.rs:180
        const double nreal = number_density(xc) * dx;                    // mod.rs:181
        const double weight = nreal / (double)npc;                       // mod.rs:182
        const size_t nsuper = nreal > 0.0 ? npc : 0;                     // mod.rs:183
        for (size_t i = 0; i < nsuper; i++) {
            const double x = rng.gen_f64();                              // mod.rs:187
            const double real_x = slab_xmin + ((double)c + x) * dx;      // mod.rs:188
            double u[3];
            const std::function<double(double, double, double)>* f[3] = {&ux, &uy, &uz};
            for (int k = 0; k < 3; k++) {                                // mod.rs:189-193: gen() then sample(StandardNormal), per component
                const double urand = rng.gen_f64();
                const double nrand = rng.standard_normal();
                u[k] = (*f[k])(real_x, urand, nrand);
            }
            p.cell.push_back(c);
            p.x.push_back(x);
            p.ux.push_back(u[0]); p.uy.push_back(u[1]); p.uz.push_back(u[2]);
            p.weight.push_back(weight);
            if (with_tau) p.tau.push_back(rng.exp1());                   // mod.rs:194 with_optical_depths
        }
    }
    return p;
}

}  // namespace opal_driver
