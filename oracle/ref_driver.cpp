// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// C-ABI driver around the *reference's own* collision translation units, compiled unmodified from
// /root/reference (see oracle/Makefile; outputs go to oracle/_ref/libchrono_ref.so). It is the checker
// the CUDA path and the CPU restatement (oracle/oracle.cpp) are pinned against. Nothing under
// chrono_wisc_b200/ may link or load this.
//
// How the reference code is reached:
//   * ChBroadphase / ChNarrowphase keep `cd_data` private but befriend `chrono::ChCollisionSystemMulticore`
//     (src/chrono/collision/multicore/ChBroadphase.h:60, ChNarrowphase.h:161).  The real class of that name
//     needs ChSystem (Eigen), so this file defines its own class with that name and the one member
//     function we need from it, GenerateAABB(), whose body is the reference's own text, included by
//     line-range extraction at build time (ChCollisionSystemMulticore.cpp:376-545 -> _ref/gen_aabb.inc).
//   * function_CalcContactForces is likewise extracted (ChIterativeSolverMulticoreSMC.cpp:56-552 ->
//     _ref/gen_smc_force.inc) and compiled against a three-enum ChSystemSMC stub.
//   * Orthogonalize / Compute_Jacobian / Cone_generalized_rigid come from
//     ChConstraintUtils.cpp:21-116 -> _ref/gen_cone.inc.
// No reference source text is stored in this repository.

#include <algorithm>
#include <climits>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <vector>

#include "chrono/multicore_math/thrust.h"
#include "chrono/collision/ChCollisionModel.h"
#include "chrono/collision/ChCollisionInfo.h"
#include "chrono/collision/multicore/ChBroadphase.h"
#include "chrono/collision/multicore/ChNarrowphase.h"
#include "chrono/collision/multicore/ChCollisionUtils.h"
#include "chrono/collision/multicore/ChConvexShape.h"
#include "chrono/multicore_math/utility.h"

namespace chrono {

// The friend of ChBroadphase/ChNarrowphase.
class ChCollisionSystemMulticore {
  public:
    ChCollisionSystemMulticore() {
        cd_data = std::make_shared<ChCollisionData>(true);
        broadphase.cd_data = cd_data;
        narrowphase.cd_data = cd_data;
    }
    void GenerateAABB();
    void SetGrid(int type, const int* res, const double* bsize, double density) {
        broadphase.grid_type = static_cast<ChBroadphase::GridType>(type);
        broadphase.grid_resolution = vec3(res[0], res[1], res[2]);
        broadphase.bin_size = real3(bsize[0], bsize[1], bsize[2]);
        broadphase.grid_density = density;
    }
    void SetAlgorithm(int a) { narrowphase.algorithm = static_cast<ChNarrowphase::Algorithm>(a); }

    std::shared_ptr<ChCollisionData> cd_data;
    ChBroadphase broadphase;
    ChNarrowphase narrowphase;
};

#include "gen_aabb.inc"

// ---- SMC force function (extracted) ---------------------------------------------------------------
struct ChSystemSMC {
    enum ContactForceModel { Hooke, Hertz, PlainCoulomb, Flores };
    enum AdhesionForceModel { Constant, DMT, Perko };
    enum TangentialDisplacementModel { None, OneStep, MultiStep };
};
}  // namespace chrono

#define max_shear 20
using namespace chrono;
#include "gen_smc_force.inc"

#define CH_MULTICORE_API
namespace chrono {
#include "gen_cone.inc"
}  // namespace chrono

// =====================================================================================================
// C ABI
// =====================================================================================================

namespace {

struct RefSystem {
    chrono::ChCollisionSystemMulticore cs;
};

inline real3 r3(const double* p) { return real3(p[0], p[1], p[2]); }
inline quaternion q4(const double* p) { return quaternion(p[0], p[1], p[2], p[3]); }

}  // namespace

extern "C" {

void* ref_create() { return new RefSystem(); }
void ref_destroy(void* h) { delete static_cast<RefSystem*>(h); }

void ref_set_num_threads(int n) { omp_set_num_threads(n); }
int ref_get_max_threads() { return omp_get_max_threads(); }

// Body state: pos[3n], rot[4n] as (w,x,y,z), flags.
void ref_set_bodies(void* h, int n, const double* pos, const double* rot, const char* active, const char* collide) {
    auto& cd = *static_cast<RefSystem*>(h)->cs.cd_data;
    cd.state_data.num_rigid_bodies = n;
    cd.state_data.pos_rigid->resize(n);
    cd.state_data.rot_rigid->resize(n);
    cd.state_data.active_rigid->resize(n);
    cd.state_data.collide_rigid->resize(n);
    for (int i = 0; i < n; i++) {
        (*cd.state_data.pos_rigid)[i] = r3(pos + 3 * i);
        (*cd.state_data.rot_rigid)[i] = q4(rot + 4 * i);
        (*cd.state_data.active_rigid)[i] = active[i];
        (*cd.state_data.collide_rigid)[i] = collide[i];
    }
}

// Shapes, appended the way ChCollisionSystemMulticore::Add does (ChCollisionSystemMulticore.cpp:86-187):
// dims[4*i..] holds: sphere (r), box-like (x,y,z), capsule (radius, hlen as stored), rounded (x,y,z,srad),
// triangle: index into tri[] (9 doubles per triangle) in dims[0]; convex: (count, offset) into convex[].
void ref_set_shapes(void* h, int ns, const int* typ, const unsigned* body_id, const short* fam, const double* ObA,
                    const double* ObR, const double* dims, const double* tri, int n_convex, const double* convex) {
    auto& cd = *static_cast<RefSystem*>(h)->cs.cd_data;
    auto& s = cd.shape_data;
    s = shape_container();
    for (int i = 0; i < n_convex; i++)
        s.convex_rigid.push_back(r3(convex + 3 * i));
    std::vector<int> local_count;
    for (int i = 0; i < ns; i++) {
        const double* d = dims + 4 * i;
        int start = -1, length = 1;
        switch (typ[i]) {
            case ChCollisionShape::SPHERE:
                start = (int)s.sphere_rigid.size();
                s.sphere_rigid.push_back(d[0]);
                break;
            case ChCollisionShape::ELLIPSOID:
            case ChCollisionShape::BOX:
            case ChCollisionShape::CYLINDER:
            case ChCollisionShape::CYLSHELL:
            case ChCollisionShape::CONE:
                start = (int)s.box_like_rigid.size();
                s.box_like_rigid.push_back(real3(d[0], d[1], d[2]));
                break;
            case ChCollisionShape::CAPSULE:
                start = (int)s.capsule_rigid.size();
                s.capsule_rigid.push_back(real2(d[0], d[1]));
                break;
            case ChCollisionShape::ROUNDEDBOX:
            case ChCollisionShape::ROUNDEDCYL:
                start = (int)s.rbox_like_rigid.size();
                s.rbox_like_rigid.push_back(real4(d[0], d[1], d[2], d[3]));
                break;
            case ChCollisionShape::CONVEXHULL:
                length = (int)d[0];
                start = (int)d[1];
                break;
            case ChCollisionShape::TRIANGLE: {
                const double* t = tri + 9 * (int)d[0];
                start = (int)s.triangle_rigid.size();
                s.triangle_rigid.push_back(r3(t));
                s.triangle_rigid.push_back(r3(t + 3));
                s.triangle_rigid.push_back(r3(t + 6));
                break;
            }
            default:
                break;
        }
        s.ObA_rigid.push_back(r3(ObA + 3 * i));
        s.ObR_rigid.push_back(q4(ObR + 4 * i));
        s.start_rigid.push_back(start);
        s.length_rigid.push_back(length);
        s.fam_rigid.push_back(S2(fam[2 * i], fam[2 * i + 1]));
        s.typ_rigid.push_back(typ[i]);
        s.id_rigid.push_back(body_id[i]);
        s.local_rigid.push_back(0);
    }
    cd.num_rigid_shapes = ns;
}

void ref_set_collision_settings(void* h, double envelope, int grid_type, const int* res, const double* bin_size,
                                double density, int algorithm) {
    auto* sys = static_cast<RefSystem*>(h);
    sys->cs.cd_data->collision_envelope = envelope;
    sys->cs.SetGrid(grid_type, res, bin_size, density);
    sys->cs.SetAlgorithm(algorithm);
}

// stage mask: 1 = AABB, 2 = broadphase, 4 = narrowphase
void ref_run_collision(void* h, int stages) {
    auto* sys = static_cast<RefSystem*>(h);
    if (stages & 1)
        sys->cs.GenerateAABB();
    if (stages & 2)
        sys->cs.broadphase.Process();
    if (stages & 4)
        sys->cs.narrowphase.Process();
}

// counts: [0] N_i  [1] N_a  [2] N_p  [3] N_c  [4..6] bins_per_axis  [7] num_bins
void ref_get_counts(void* h, unsigned* c) {
    auto& cd = *static_cast<RefSystem*>(h)->cs.cd_data;
    c[0] = cd.num_bin_aabb_intersections;
    c[1] = cd.num_active_bins;
    c[2] = cd.num_possible_collisions;
    c[3] = cd.num_rigid_contacts;
    c[4] = cd.bins_per_axis.x;
    c[5] = cd.bins_per_axis.y;
    c[6] = cd.bins_per_axis.z;
    c[7] = cd.num_bins;
}

// grid: min_bounding_point[3], max_bounding_point[3], bin_size[3], inv_bin_size[3], rigid_min[3], rigid_max[3]
void ref_get_grid(void* h, double* g) {
    auto& cd = *static_cast<RefSystem*>(h)->cs.cd_data;
    const real3* v[6] = {&cd.min_bounding_point, &cd.max_bounding_point,       &cd.bin_size,
                         &cd.inv_bin_size,       &cd.rigid_min_bounding_point, &cd.rigid_max_bounding_point};
    for (int k = 0; k < 6; k++) {
        g[3 * k + 0] = v[k]->x;
        g[3 * k + 1] = v[k]->y;
        g[3 * k + 2] = v[k]->z;
    }
}

void ref_get_aabb(void* h, double* mn, double* mx) {
    auto& cd = *static_cast<RefSystem*>(h)->cs.cd_data;
    for (size_t i = 0; i < cd.aabb_min.size(); i++) {
        mn[3 * i + 0] = cd.aabb_min[i].x, mn[3 * i + 1] = cd.aabb_min[i].y, mn[3 * i + 2] = cd.aabb_min[i].z;
        mx[3 * i + 0] = cd.aabb_max[i].x, mx[3 * i + 1] = cd.aabb_max[i].y, mx[3 * i + 2] = cd.aabb_max[i].z;
    }
}

// Sorted (bin, shape) lists and the run-length encoding of active bins.
void ref_get_bins(void* h, unsigned* bin_number, unsigned* bin_aabb_number, unsigned* bin_active,
                  unsigned* bin_start_index) {
    auto& cd = *static_cast<RefSystem*>(h)->cs.cd_data;
    if (bin_number)
        std::copy(cd.bin_number.begin(), cd.bin_number.begin() + cd.num_bin_aabb_intersections, bin_number);
    if (bin_aabb_number)
        std::copy(cd.bin_aabb_number.begin(), cd.bin_aabb_number.begin() + cd.num_bin_aabb_intersections,
                  bin_aabb_number);
    if (bin_active)
        std::copy(cd.bin_active.begin(), cd.bin_active.begin() + cd.num_active_bins, bin_active);
    if (bin_start_index)
        std::copy(cd.bin_start_index.begin(), cd.bin_start_index.begin() + cd.num_active_bins + 1, bin_start_index);
}

void ref_get_pairs(void* h, long long* pairs) {
    auto& cd = *static_cast<RefSystem*>(h)->cs.cd_data;
    std::copy(cd.pair_shapeIDs.begin(), cd.pair_shapeIDs.begin() + cd.num_possible_collisions, pairs);
}

void ref_get_contacts(void* h, double* norm, double* pta, double* ptb, double* depth, double* erad, int* bids,
                      long long* shape_ids) {
    auto& cd = *static_cast<RefSystem*>(h)->cs.cd_data;
    const unsigned n = cd.num_rigid_contacts;
    for (unsigned i = 0; i < n; i++) {
        const real3& N = cd.norm_rigid_rigid[i];
        const real3& A = cd.cpta_rigid_rigid[i];
        const real3& B = cd.cptb_rigid_rigid[i];
        norm[3 * i] = N.x, norm[3 * i + 1] = N.y, norm[3 * i + 2] = N.z;
        pta[3 * i] = A.x, pta[3 * i + 1] = A.y, pta[3 * i + 2] = A.z;
        ptb[3 * i] = B.x, ptb[3 * i + 1] = B.y, ptb[3 * i + 2] = B.z;
        depth[i] = cd.dpth_rigid_rigid[i];
        erad[i] = cd.erad_rigid_rigid[i];
        bids[2 * i] = cd.bids_rigid_rigid[i].x;
        bids[2 * i + 1] = cd.bids_rigid_rigid[i].y;
        shape_ids[i] = cd.contact_shapeIDs[i];
    }
}

// Single-pair narrowphase through the reference dispatchers (ConvexShapeCustom carries pos/rot/dims/radius the way
// the reference's own unit tests do, src/tests/unit_tests/collision/utest_COLL_narrow_prims.cpp).
// dims = (x,y,z,w); for a sphere x = radius; capsule (radius, hlen); rounded (x,y,z) + w = sweep radius.
// Outputs sized for 8 contacts. Returns nC (or -1 if the pair type is not handled analytically).
int ref_prims_pair(int typeA, const double* posA, const double* rotA, const double* dimsA, int typeB, const double* posB,
                   const double* rotB, const double* dimsB, double separation, double* norm, double* pt1, double* pt2,
                   double* depth, double* erad) {
    ConvexShapeCustom A(typeA, r3(posA), q4(rotA), real3(dimsA[0], dimsA[1], dimsA[2]), dimsA[3]);
    ConvexShapeCustom B(typeB, r3(posB), q4(rotB), real3(dimsB[0], dimsB[1], dimsB[2]), dimsB[3]);
    real3 n[8], p1[8], p2[8];
    real d[8], e[8];
    int nC = 0;
    bool ok = ChNarrowphase::PRIMSCollision(&A, &B, separation, n, p1, p2, d, e, nC);
    if (!ok)
        return -1;
    for (int i = 0; i < nC; i++) {
        norm[3 * i] = n[i].x, norm[3 * i + 1] = n[i].y, norm[3 * i + 2] = n[i].z;
        pt1[3 * i] = p1[i].x, pt1[3 * i + 1] = p1[i].y, pt1[3 * i + 2] = p1[i].z;
        pt2[3 * i] = p2[i].x, pt2[3 * i + 1] = p2[i].y, pt2[3 * i + 2] = p2[i].z;
        depth[i] = d[i];
        erad[i] = e[i];
    }
    return nC;
}

int ref_mpr_pair(int typeA, const double* posA, const double* rotA, const double* dimsA, int typeB, const double* posB,
                 const double* rotB, const double* dimsB, double envelope, double* norm, double* pt1, double* pt2,
                 double* depth) {
    ConvexShapeCustom A(typeA, r3(posA), q4(rotA), real3(dimsA[0], dimsA[1], dimsA[2]), dimsA[3]);
    ConvexShapeCustom B(typeB, r3(posB), q4(rotB), real3(dimsB[0], dimsB[1], dimsB[2]), dimsB[3]);
    real3 n, p1, p2;
    real d;
    if (!ChNarrowphase::MPRCollision(&A, &B, envelope, n, p1, p2, d))
        return 0;
    norm[0] = n.x, norm[1] = n.y, norm[2] = n.z;
    pt1[0] = p1.x, pt1[1] = p1.y, pt1[2] = p1.z;
    pt2[0] = p2.x, pt2[1] = p2.y, pt2[2] = p2.z;
    depth[0] = d;
    return 1;
}

// Reference SMC per-contact force function over a contact list (flat unpadded arrays in, flat out).
// shear_* may be null for OneStep/None. ct_force/ct_torque: [2*nc*3].
void ref_smc_contact_forces(int nc, const int* bids, const int* shape_pairs, int force_model, int adhesion_model,
                            int displ_mode, int use_mat_props, double char_vel, double min_slip_vel,
                            double min_roll_vel, double min_spin_vel, double dT, int nb, const double* mass,
                            const double* pos, const double* rot, const double* vel, const double* fric,
                            const double* modulus, const double* adhesion, const double* cr, const double* smc,
                            const double* pt1, const double* pt2, const double* normal, const double* depth,
                            const double* erad, int* shear_neigh, char* shear_touch, double* shear_disp,
                            double* relvel_init, double* duration, int* ct_bid, double* ct_force, double* ct_torque) {
    std::vector<vec2> bp(nc), sp(nc);
    std::vector<real3> P(nb), F(nc), A(nc), p1(nc), p2(nc), nr(nc), cf(2 * nc), ctq(2 * nc);
    std::vector<quaternion> R(nb);
    std::vector<real2> M(nc);
    std::vector<real4> S(nc);
    std::vector<real> m(mass, mass + nb), v(vel, vel + 6 * nb), c(cr, cr + nc), d(depth, depth + nc),
        er(erad, erad + nc);
    for (int i = 0; i < nb; i++) {
        P[i] = r3(pos + 3 * i);
        R[i] = q4(rot + 4 * i);
    }
    for (int i = 0; i < nc; i++) {
        bp[i] = I2(bids[2 * i], bids[2 * i + 1]);
        sp[i] = shape_pairs ? I2(shape_pairs[2 * i], shape_pairs[2 * i + 1]) : I2(0, 0);
        F[i] = r3(fric + 3 * i);
        M[i] = real2(modulus[2 * i], modulus[2 * i + 1]);
        A[i] = r3(adhesion + 3 * i);
        S[i] = real4(smc[4 * i], smc[4 * i + 1], smc[4 * i + 2], smc[4 * i + 3]);
        p1[i] = r3(pt1 + 3 * i);
        p2[i] = r3(pt2 + 3 * i);
        nr[i] = r3(normal + 3 * i);
    }
    std::vector<vec3> sn;
    std::vector<real3> sd;
    if (shear_neigh) {
        sn.resize(max_shear * nb);
        sd.resize(max_shear * nb);
        for (int i = 0; i < max_shear * nb; i++) {
            sn[i] = vec3(shear_neigh[3 * i], shear_neigh[3 * i + 1], shear_neigh[3 * i + 2]);
            sd[i] = r3(shear_disp + 3 * i);
        }
    }
    std::vector<int> bid(2 * nc);
    // Serial on purpose when history is mutated (the reference's omp loop races there).
#pragma omp parallel for if (!shear_neigh)
    for (int i = 0; i < nc; i++) {
        function_CalcContactForces(
            i, bp.data(), sp.data(), (ChSystemSMC::ContactForceModel)force_model,
            (ChSystemSMC::AdhesionForceModel)adhesion_model, (ChSystemSMC::TangentialDisplacementModel)displ_mode,
            use_mat_props != 0, char_vel, min_slip_vel, min_roll_vel, min_spin_vel, dT, m.data(), P.data(), R.data(),
            v.data(), F.data(), M.data(), A.data(), c.data(), S.data(), p1.data(), p2.data(), nr.data(), d.data(),
            er.data(), sn.data(), shear_touch, sd.data(), relvel_init, duration, bid.data(), cf.data(), ctq.data());
    }
    for (int i = 0; i < 2 * nc; i++) {
        ct_bid[i] = bid[i];
        ct_force[3 * i] = cf[i].x, ct_force[3 * i + 1] = cf[i].y, ct_force[3 * i + 2] = cf[i].z;
        ct_torque[3 * i] = ctq[i].x, ct_torque[3 * i + 1] = ctq[i].y, ct_torque[3 * i + 2] = ctq[i].z;
    }
    if (shear_neigh) {
        for (int i = 0; i < max_shear * nb; i++) {
            shear_neigh[3 * i] = sn[i].x, shear_neigh[3 * i + 1] = sn[i].y, shear_neigh[3 * i + 2] = sn[i].z;
            shear_disp[3 * i] = sd[i].x, shear_disp[3 * i + 1] = sd[i].y, shear_disp[3 * i + 2] = sd[i].z;
        }
    }
}

// Contact frame and Jacobian blocks (ChConstraintUtils.cpp:22-51), cone projection (:69-92).
void ref_orthogonalize(const double* U, double* V, double* W) {
    real3 v, w;
    chrono::Orthogonalize(r3(U), v, w);
    V[0] = v.x, V[1] = v.y, V[2] = v.z;
    W[0] = w.x, W[1] = w.y, W[2] = w.z;
}

void ref_compute_jacobian(const double* quat, const double* U, const double* V, const double* W, const double* point,
                          double* T1, double* T2, double* T3) {
    real3 t1, t2, t3;
    chrono::Compute_Jacobian(q4(quat), r3(U), r3(V), r3(W), r3(point), t1, t2, t3);
    T1[0] = t1.x, T1[1] = t1.y, T1[2] = t1.z;
    T2[0] = t2.x, T2[1] = t2.y, T2[2] = t2.z;
    T3[0] = t3.x, T3[1] = t3.y, T3[2] = t3.z;
}

int ref_cone_generalized(double* gamma_n, double* gamma_u, double* gamma_v, double mu) {
    return chrono::Cone_generalized_rigid(*gamma_n, *gamma_u, *gamma_v, mu) ? 1 : 0;
}

// Math helpers exposed for unit pins of the restatement / device math (MM/real4.cpp:158-188).
void ref_rotate(const double* v, const double* q, double* out) {
    real3 r = Rotate(r3(v), q4(q));
    out[0] = r.x, out[1] = r.y, out[2] = r.z;
}
void ref_rotate_t(const double* v, const double* q, double* out) {
    real3 r = RotateT(r3(v), q4(q));
    out[0] = r.x, out[1] = r.y, out[2] = r.z;
}
void ref_abs_rotate(const double* q, const double* v, double* out) {
    real3 r = AbsRotate(q4(q), r3(v));
    out[0] = r.x, out[1] = r.y, out[2] = r.z;
}
void ref_quat_mult(const double* a, const double* b, double* out) {
    quaternion r = Mult(q4(a), q4(b));
    out[0] = r.w, out[1] = r.x, out[2] = r.y, out[3] = r.z;
}

}  // extern "C"
