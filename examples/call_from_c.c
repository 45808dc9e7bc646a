/* call_from_c.c -- the C ABI of include/opal_b200.h from plain C99: one periodic slab, a thermal electron population,
 * 50 fused steps of the loop body main.rs:238-267, the four conservation energies before and after.
 *
 *   gcc -std=c99 -Iinclude examples/call_from_c.c -Lopal_b200/csrc -lopal_b200 -Wl,-rpath,$PWD/opal_b200/csrc -lm -o call_from_c
 *
 * The same source runs over the CPU oracle (test infrastructure) with
 *   gcc -std=c99 -DOPAL_PREFIX=opal_ref_ -Iinclude examples/call_from_c.c -Loracle -lopal_oracle -Wl,-rpath,$PWD/oracle -lm ...
 * which is how tests/test_abi_symbols.py exercises it without a GPU.  Without a CUDA device the product library
 * answers OPAL_ECUDA from opal_create: there is no CPU fallback. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "opal_b200.h"

#define CK(call)                                                                                  \
    do {                                                                                          \
        opal_status st_ = (call);                                                                 \
        if (st_ != OPAL_OK) {                                                                     \
            fprintf(stderr, "%s -> status %d: %s\n", #call, (int)st_, ctx ? OPAL_FN(last_error)(ctx) : "(no context)"); \
            return 1;                                                                             \
        }                                                                                         \
    } while (0)

static double uniform(unsigned long long* s) {                 /* splitmix64: any generator will do for a demo */
    unsigned long long z = (*s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return (double)((z ^ (z >> 31)) >> 11) * (1.0 / 9007199254740992.0);
}

int main(void) {
    const int64_t nx = 256, npc = 64, n = nx * npc;
    const double dx = 500.0, c = 2.997925e8, dt = 0.95 * dx / c, u_th = 0.03128, ne = 20.0;
    const double two_pi = 6.283185307179586;
    opal_ctx* ctx = NULL;
    CK(OPAL_FN(create)(&ctx, 0, nx, 0.0, dx, dt, OPAL_BDY_INTERNAL, OPAL_BDY_INTERNAL, 0, 1, 12345u, 0u));

    int64_t* cell = malloc((size_t)n * sizeof *cell);
    double* a = malloc((size_t)n * 5 * sizeof *a);
    double *x = a, *ux = a + n, *uy = a + 2 * n, *uz = a + 3 * n, *w = a + 4 * n;
    unsigned long long seed = 1;
    for (int64_t i = 0; i < n; i++) {
        cell[i] = i / npc;
        x[i] = uniform(&seed);
        const double r1 = sqrt(-2.0 * log(1.0 - uniform(&seed))), r2 = sqrt(-2.0 * log(1.0 - uniform(&seed)));
        const double p1 = two_pi * uniform(&seed), p2 = two_pi * uniform(&seed);
        ux[i] = u_th * r1 * cos(p1);
        uy[i] = u_th * r1 * sin(p1);
        uz[i] = u_th * r2 * cos(p2);
        w[i] = ne * dx / (double)npc;
    }
    CK(OPAL_FN(particles_upload)(ctx, OPAL_ELECTRON, n, cell, x, ux, uy, uz, w, NULL));

    double e0[4], e1[4];
    CK(OPAL_FN(energy)(ctx, &e0[0], &e0[1], &e0[2], &e0[3]));
    CK(OPAL_FN(run_steps)(ctx, 50, 0.0, NULL, NULL, 1, 0, NULL));
    CK(OPAL_FN(energy)(ctx, &e1[0], &e1[1], &e1[2], &e1[3]));
    int64_t left = 0, right = 0, count = 0;
    CK(OPAL_FN(last_leavers)(ctx, OPAL_ELECTRON, &left, &right));
    CK(OPAL_FN(particles_count)(ctx, OPAL_ELECTRON, &count));
    printf("particles %lld (wrapped around the ring in the last step: %lld left, %lld right)\n", (long long)count, (long long)left, (long long)right);
    printf("em_field  %.6e -> %.6e J\nelectrons %.6e -> %.6e J\n", e0[0], e1[0], e0[1], e1[1]);
    const double tot0 = e0[0] + e0[1], tot1 = e1[0] + e1[1];
    printf("total energy drift over 50 steps: %.2e\n", tot1 / tot0 - 1.0);
    free(cell);
    free(a);
    CK(OPAL_FN(destroy)(ctx));
    return count == n && fabs(tot1 / tot0 - 1.0) < 1e-3 ? 0 : 2;
}
