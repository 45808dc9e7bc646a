// loader.h -- initial particle loading: Population<T>::new, src/particle/mod.rs:172-203.
//
// Random numbers.  The reference draws from rand_xoshiro 0.4.0's Xoshiro256StarStar seeded with the MPI rank
// (main.rs:53) and from rand_distr 0.2.2's StandardNormal / Exp1 ziggurats.  Neither crate is vendored under
// /root/reference and no reference test pins a random output ("parity unpinned"), so this follows the
// published xoshiro256** / SplitMix64 algorithms for the uniform stream, and draws normals by Box-Muller and
// exponentials by inversion: the loaded plasma is statistically, not bitwise, the reference's.  The draw ORDER per
// particle is the reference's: x; then (urand, nrand) for ux, uy, uz; then the optical depth (mod.rs:186-193).
#pragma once
#include <cstdint>
#include <functional>
#include <vector>

namespace opal_driver {

class Xoshiro256StarStar {
public:
    explicit Xoshiro256StarStar(uint64_t seed);      // seed_from_u64: SplitMix64 fill
    uint64_t next_u64();
    double gen_f64() { return (double)(next_u64() >> 11) * (1.0 / 9007199254740992.0); }   // rand 0.7 Standard for f64: 53 bits, [0,1)
    double standard_normal();                         // Box-Muller (one value per call, second discarded)
    double exp1();                                    // -ln(1 - U)

private:
    uint64_t s_[4];
};

struct ParticleArrays {       // SoA, slab-local cells: the arguments of opal_particles_upload
    std::vector<int64_t> cell;
    std::vector<double> x, ux, uy, uz, weight, tau;
    size_t size() const { return cell.size(); }
};

// Population::new for the cells [0, ncells) of a slab whose first cell starts at slab_xmin (Grid::xmin, yee.rs:531).
// with_tau: Electron::with_optical_depths draws tau = Exp1 (electron.rs:170); ions draw nothing (ion.rs:160).
ParticleArrays load_population(size_t npc, const std::function<double(double)>& number_density,
                               const std::function<double(double, double, double)>& ux, const std::function<double(double, double, double)>& uy,
                               const std::function<double(double, double, double)>& uz, int64_t ncells, double slab_xmin, double dx,
                               Xoshiro256StarStar& rng, bool with_tau);

}  // namespace opal_driver
