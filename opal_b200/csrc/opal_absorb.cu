// opal_absorb.cu -- photon absorption and stimulated emission (particle/interactions.rs:145-340, photon.rs:312-362,
// qed/photon_absorption.rs, qed/stimulated_emission.rs, qed/special_functions/airy.rs).
//
// Compiled with --fmad=false (see opal_absorb.cuh).
#include <cstdint>

#include "opal_absorb.cuh"
#include "opal_b200.h"

namespace opal {
namespace pairs {

__global__ void k_airy(int n, const double* __restrict__ x, double* __restrict__ ai, int* __restrict__ defined) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = 0.0;
    const bool ok = airy_ai_for_positive(x[i], &v);
    ai[i] = ok ? v : 0.0;
    defined[i] = ok ? 1 : 0;
}

__global__ void k_pair_cross_sections(int n, const double* __restrict__ k4, const double* __restrict__ p4, const double* __restrict__ chi_gamma,
                                      const double* __restrict__ chi_e, double* __restrict__ sigma_abs, double* __restrict__ sigma_st,
                                      int* __restrict__ defined) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double k[4] = {k4[4 * i], k4[4 * i + 1], k4[4 * i + 2], k4[4 * i + 3]};
    const double p[4] = {p4[4 * i], p4[4 * i + 1], p4[4 * i + 2], p4[4 * i + 3]};
    double sa = 0.0, ss = 0.0;
    const bool a = pair_cross_section<+1>(k, p, chi_gamma[i], chi_e[i], &sa);
    const bool s = pair_cross_section<-1>(k, p, chi_gamma[i], chi_e[i], &ss);
    sigma_abs[i] = a ? sa : 0.0;
    sigma_st[i] = s ? ss : 0.0;
    defined[i] = (a ? 1 : 0) | (s ? 2 : 0);
}

namespace {
// a few device arrays for the lifetime of one pure-function call
struct Scratch {
    void* p[8] = {nullptr};
    int n = 0;
    bool ok = true;
    template <class T>
    T* in(const T* host, size_t count) {
        T* d = out<T>(count);
        if (d && cudaMemcpy(d, host, count * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) ok = false;
        return d;
    }
    template <class T>
    T* out(size_t count) {
        void* d = nullptr;
        if (cudaMalloc(&d, count * sizeof(T)) != cudaSuccess) { ok = false; return nullptr; }
        p[n++] = d;
        return (T*)d;
    }
    template <class T>
    void back(T* host, const T* dev, size_t count) {
        if (ok && cudaMemcpy(host, dev, count * sizeof(T), cudaMemcpyDeviceToHost) != cudaSuccess) ok = false;
    }
    ~Scratch() { for (int i = 0; i < n; i++) cudaFree(p[i]); }
};
}  // namespace

}  // namespace pairs
}  // namespace opal

using namespace opal::pairs;

extern "C" {

opal_status opal_qed_airy(int64_t n, const double* x, double* ai, int* defined) {
    if (n < 0 || n > INT32_MAX || (n > 0 && (!x || !ai || !defined))) return OPAL_EINVAL;
    if (n == 0) return OPAL_OK;
    Scratch s;
    const double* d_x = s.in(x, (size_t)n);
    double* d_ai = s.out<double>((size_t)n);
    int* d_ok = s.out<int>((size_t)n);
    if (!s.ok) return OPAL_ECUDA;
    k_airy<<<(unsigned)((n + 127) / 128), 128>>>((int)n, d_x, d_ai, d_ok);
    if (cudaGetLastError() != cudaSuccess) return OPAL_ECUDA;
    s.back(ai, d_ai, (size_t)n);
    s.back(defined, d_ok, (size_t)n);
    return s.ok ? OPAL_OK : OPAL_ECUDA;
}

opal_status opal_qed_pair_cross_sections(int64_t n, const double* k4, const double* p4, const double* chi_gamma, const double* chi_e,
                                         double* sigma_abs, double* sigma_st, int* defined) {
    if (n < 0 || n > INT32_MAX / 4 || (n > 0 && (!k4 || !p4 || !chi_gamma || !chi_e || !sigma_abs || !sigma_st || !defined))) return OPAL_EINVAL;
    if (n == 0) return OPAL_OK;
    Scratch s;
    const double* d_k = s.in(k4, (size_t)n * 4);
    const double* d_p = s.in(p4, (size_t)n * 4);
    const double* d_cg = s.in(chi_gamma, (size_t)n);
    const double* d_ce = s.in(chi_e, (size_t)n);
    double* d_sa = s.out<double>((size_t)n);
    double* d_ss = s.out<double>((size_t)n);
    int* d_ok = s.out<int>((size_t)n);
    if (!s.ok) return OPAL_ECUDA;
    k_pair_cross_sections<<<(unsigned)((n + 127) / 128), 128>>>((int)n, d_k, d_p, d_cg, d_ce, d_sa, d_ss, d_ok);
    if (cudaGetLastError() != cudaSuccess) return OPAL_ECUDA;
    s.back(sigma_abs, d_sa, (size_t)n);
    s.back(sigma_st, d_ss, (size_t)n);
    s.back(defined, d_ok, (size_t)n);
    return s.ok ? OPAL_OK : OPAL_ECUDA;
}

}  // extern "C"
