/* oracle.h -- C interface of liboracle.so, the CPU restatement of the Chrono::Multicore granular-contact hot path.
 *
 * TEST INFRASTRUCTURE, NOT PRODUCT CODE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this.  It restates, in plain scalar C++ (own containers, std::sort instead of
 * Thrust), the reference algorithms cited function by function in oracle.cpp, and is pinned against
 *   - the reference's own collision code built from /root/reference (oracle/_ref, tests/test_oracle_vs_ref.py),
 *   - golden vectors taken from the reference's unit tests (tests/golden/).
 * Scope: spheres, boxes and the other sphere-analytic shapes; SMC force models; body integration.
 */
#ifndef CB200_ORACLE_H
#define CB200_ORACLE_H
#include <stdint.h>
#include "../include/chrono_b200.h" /* settings struct only (the boundary contract) */

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_system orc_system;

orc_system* orc_create(void);
void orc_destroy(orc_system*);
void orc_set_threads(int n); /* OpenMP threads for the embarrassingly parallel loops; 1 = scalar port */
int orc_max_threads(void);
void orc_set_settings(orc_system*, const cb200_settings*);
void orc_default_settings(cb200_settings*);
void orc_set_bodies(orc_system*, int nb, const double* mass, const double* inertia_xx, const char* active,
                    const char* collide);
void orc_set_shapes(orc_system*, int ns, const int32_t* type, const uint32_t* body, const int16_t* fam,
                    const double* lpos, const double* lrot, const double* dims, const int32_t* mat);
void orc_set_materials_smc(orc_system*, int nm, const float* table);
void orc_set_state(orc_system*, const double* pos, const double* rot, const double* vel);
void orc_get_state(orc_system*, double* pos, double* rot, double* vel);

void orc_collide(orc_system*);          /* AABB + broadphase + narrowphase */
void orc_smc_step(orc_system*, int n);  /* n full SMC steps */

/* counts: N_i, N_a, N_p, N_c, bins x/y/z, num_bins */
void orc_get_counts(orc_system*, uint32_t* c8);
void orc_get_grid(orc_system*, double* g12); /* origin[3], max[3], bin_size[3], inv_bin_size[3] */
void orc_get_aabb(orc_system*, double* mn, double* mx);
void orc_get_pairs(orc_system*, int64_t* pairs);
void orc_get_contacts(orc_system*, double* normal, double* pta, double* ptb, double* depth, double* erad,
                      int32_t* bids, int64_t* shape_ids);
void orc_get_contact_forces(orc_system*, double* ct_force, double* ct_torque); /* 2 rows per contact */
void orc_get_body_forces(orc_system*, double* force, double* torque);          /* dense per body */
void orc_get_hf(orc_system*, double* hf);
/* seconds spent in the last orc_smc_step call, by phase: update, broad, narrow, forces, accumulate, integrate */
void orc_get_timers(orc_system*, double* t6);

#ifdef __cplusplus
}
#endif
#endif
