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
#include <cmath>
#include <omp.h>

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

// State of the whole-step driver (ref_smc_step): what host_container holds beside the collision data
// (MC/ChDataManager.h:228-262) in the reference's own types.
struct StepState {
    int nb = 0;
    std::vector<real> mass, v, hf;          // v, hf: 6 per body (MC/physics/ChSystemMulticore.cpp:350-362)
    std::vector<real3> inertia, inv_inertia;
    std::vector<float> materials;           // nm x 14 (ChContactMaterialSMC fields, see chrono_b200.h)
    std::vector<int> shape_mat;
    int force_model = 1, adhesion_model = 0, displ_mode = 1, use_mat_props = 1;
    real char_vel = 1, min_slip_vel = 1e-4, min_roll_vel = 1e-4, min_spin_vel = 1e-4, h = 1e-3;
    real3 gravity = real3(0, 0, -9.81);
    double t[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // seconds: update, broad, narrow, materials, forces, accumulate, velocity, integrate
    long steps = 0;
    unsigned ct_body_count = 0;
};

struct RefSystem {
    chrono::ChCollisionSystemMulticore cs;
    StepState st;
};

// thrust::reduce_by_key functor over (force, torque) tuples, as sum_tuples in
// MC/solver/ChIterativeSolverMulticoreSMC.cpp:639-648
struct add_pairs {
    template <typename A, typename B>
    thrust::tuple<real3, real3> operator()(const A& a, const B& b) const {
        return thrust::tuple<real3, real3>(thrust::get<0>(a) + thrust::get<0>(b), thrust::get<1>(a) + thrust::get<1>(b));
    }
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

// ---------------------------------------------------------------------------------------------------------------
// Whole SMC step on the host cores the way the reference runs it: the reference's own collision translation units
// (GenerateAABB, ChBroadphase::Process, ChNarrowphase::Process -- Thrust with the OpenMP backend), its own
// function_CalcContactForces under the OpenMP loop of host_CalcContactForces, Thrust sort_by_key / reduce_by_key for
// ProcessContacts; the parts that live in Eigen / ChBody translation units are restated beside them ("glue"), each
// with the threading the reference gives it.  This is the timed CPU arm of bench.py (kind "_ref+glue").
// ---------------------------------------------------------------------------------------------------------------
void ref_smc_setup(void* h, int nb, const double* mass, const double* inertia, const double* vel, int nm,
                   const float* materials, int ns, const int* shape_mat, const int* isettings, const double* dsettings) {
    auto& st = static_cast<RefSystem*>(h)->st;
    st.nb = nb;
    st.mass.assign(mass, mass + nb);
    st.v.assign(vel, vel + 6 * (size_t)nb);
    st.hf.assign(6 * (size_t)nb, 0.0);
    st.inertia.resize(nb);
    st.inv_inertia.resize(nb);
    for (int i = 0; i < nb; i++) {
        st.inertia[i] = r3(inertia + 3 * i);
        st.inv_inertia[i] = real3(1.0 / inertia[3 * i], 1.0 / inertia[3 * i + 1], 1.0 / inertia[3 * i + 2]);
    }
    st.materials.assign(materials, materials + 14 * (size_t)nm);
    st.shape_mat.assign(shape_mat, shape_mat + ns);
    st.force_model = isettings[0], st.adhesion_model = isettings[1], st.displ_mode = isettings[2];
    st.use_mat_props = isettings[3];
    st.char_vel = dsettings[0], st.min_slip_vel = dsettings[1], st.min_roll_vel = dsettings[2];
    st.min_spin_vel = dsettings[3], st.h = dsettings[4];
    st.gravity = real3(dsettings[5], dsettings[6], dsettings[7]);
    for (double& x : st.t)
        x = 0;
    st.steps = 0;
}

void ref_smc_step(void* hh, int nsteps) {
    auto* sys = static_cast<RefSystem*>(hh);
    auto& st = sys->st;
    auto& cd = *sys->cs.cd_data;
    const int nb = st.nb;
    const real h = st.h;
    std::vector<real3>& pos = *cd.state_data.pos_rigid;
    std::vector<quaternion>& rot = *cd.state_data.rot_rigid;
    std::vector<char>& active = *cd.state_data.active_rigid;
    for (int step = 0; step < nsteps; step++) {
        double t0 = omp_get_wtime(), t1;
        // UpdateRigidBodies (MC/physics/ChSystemMulticore.cpp:329-377, omp over bodies) with ChBody::UpdateForces /
        // ComputeGyro / VariablesFbLoadForces (src/chrono/physics/ChBody.cpp:552-574,423-426,247-257)
#pragma omp parallel for
        for (int i = 0; i < nb; i++) {
            const real3 w(st.v[6 * i + 3], st.v[6 * i + 4], st.v[6 * i + 5]);
            const real3 gyro = Cross(w, st.inertia[i] * w);
            const real3 F = st.gravity * st.mass[i];
            const real3 hF = h * F, hT = h * (real3(0) - gyro);
            st.hf[6 * i] = hF.x, st.hf[6 * i + 1] = hF.y, st.hf[6 * i + 2] = hF.z;
            st.hf[6 * i + 3] = hT.x, st.hf[6 * i + 4] = hT.y, st.hf[6 * i + 5] = hT.z;
        }
        t1 = omp_get_wtime(), st.t[0] += t1 - t0, t0 = t1;
        // ChCollisionSystemMulticore::Run (CMC/ChCollisionSystemMulticore.cpp:289-345): the reference's own code
        sys->cs.GenerateAABB();
        sys->cs.broadphase.Process();
        t1 = omp_get_wtime(), st.t[1] += t1 - t0, t0 = t1;
        sys->cs.narrowphase.Process();
        t1 = omp_get_wtime(), st.t[2] += t1 - t0, t0 = t1;
        const int nc = (int)cd.num_rigid_contacts;
        // composite materials per contact, float arithmetic (ChContactContainerMulticoreSMC::AddContact,
        // MC/collision/ChContactContainerMulticoreSMC.cpp:123-174 + ChContactMaterialCompositeSMC's constructor,
        // src/chrono/physics/ChContactMaterialSMC.cpp:107-130)
        std::vector<real3> fric(nc), adhesion(nc);
        std::vector<real2> modulus(nc);
        std::vector<real> cr(nc);
        std::vector<real4> smc(nc);
#pragma omp parallel for
        for (int i = 0; i < nc; i++) {
            const long long key = cd.contact_shapeIDs[i];
            const float* m1 = st.materials.data() + 14 * st.shape_mat[(int)(key >> 32)];
            const float* m2 = st.materials.data() + 14 * st.shape_mat[(int)(key & 0xffffffff)];
            const float E1 = m1[0], nu1 = m1[1], E2 = m2[0], nu2 = m2[1];
            const float inv_E = (1 - nu1 * nu1) / E1 + (1 - nu2 * nu2) / E2;
            const float inv_G = 2 * (2 - nu1) * (1 + nu1) / E1 + 2 * (2 - nu2) * (1 + nu2) / E2;
            const float E_eff = 1 / inv_E, G_eff = 1 / inv_G;
            fric[i] = real3(std::min(m1[2], m2[2]), std::min(m1[4], m2[4]), std::min(m1[5], m2[5]));
            modulus[i] = real2(E_eff, G_eff);
            adhesion[i] = real3(std::min(m1[7], m2[7]), std::min(m1[8], m2[8]), std::min(m1[9], m2[9]));
            cr[i] = std::min(m1[6], m2[6]);
            smc[i] = real4((float)((m1[10] + m2[10]) / 2), (float)((m1[11] + m2[11]) / 2), (float)((m1[12] + m2[12]) / 2),
                           (float)((m1[13] + m2[13]) / 2));
        }
        t1 = omp_get_wtime(), st.t[3] += t1 - t0, t0 = t1;
        // ProcessContacts step 1 = host_CalcContactForces (MC/solver/ChIterativeSolverMulticoreSMC.cpp:558-602)
        std::vector<int> ct_bid(2 * nc);
        std::vector<real3> ct_force(2 * nc), ct_torque(2 * nc);
        std::vector<vec2> shape_pairs(1);
#pragma omp parallel for
        for (int i = 0; i < nc; i++) {
            function_CalcContactForces(
                i, cd.bids_rigid_rigid.data(), shape_pairs.data(), (ChSystemSMC::ContactForceModel)st.force_model,
                (ChSystemSMC::AdhesionForceModel)st.adhesion_model, (ChSystemSMC::TangentialDisplacementModel)st.displ_mode,
                st.use_mat_props != 0, st.char_vel, st.min_slip_vel, st.min_roll_vel, st.min_spin_vel, h, st.mass.data(),
                pos.data(), rot.data(), st.v.data(), fric.data(), modulus.data(), adhesion.data(), cr.data(), smc.data(),
                cd.cpta_rigid_rigid.data(), cd.cptb_rigid_rigid.data(), cd.norm_rigid_rigid.data(),
                cd.dpth_rigid_rigid.data(), cd.erad_rigid_rigid.data(), nullptr, nullptr, nullptr, nullptr, nullptr,
                ct_bid.data(), ct_force.data(), ct_torque.data());
        }
        t1 = omp_get_wtime(), st.t[4] += t1 - t0, t0 = t1;
        // ProcessContacts steps 2-3 (:696-736): sort by body, reduce by body, hf += h (F, T)
        std::vector<int> ct_body_id(nb);
        std::vector<real3> ct_body_force(nb), ct_body_torque(nb);
        unsigned count = 0;
        if (nc) {
            thrust::sort_by_key(THRUST_PAR ct_bid.begin(), ct_bid.end(),
                                thrust::make_zip_iterator(thrust::make_tuple(ct_force.begin(), ct_torque.begin())));
            auto end_range = thrust::reduce_by_key(
                THRUST_PAR ct_bid.begin(), ct_bid.end(),
                thrust::make_zip_iterator(thrust::make_tuple(ct_force.begin(), ct_torque.begin())), ct_body_id.begin(),
                thrust::make_zip_iterator(thrust::make_tuple(ct_body_force.begin(), ct_body_torque.begin())),
                thrust::equal_to<int>(), add_pairs());
            count = (unsigned)(end_range.first - ct_body_id.begin());
        }
        st.ct_body_count = count;
#pragma omp parallel for
        for (int k = 0; k < (int)count; k++) {  // host_AddContactForces (:611-626)
            const int b = ct_body_id[k];
            const real3 F = h * ct_body_force[k], T = h * ct_body_torque[k];
            st.hf[6 * b] += F.x, st.hf[6 * b + 1] += F.y, st.hf[6 * b + 2] += F.z;
            st.hf[6 * b + 3] += T.x, st.hf[6 * b + 4] += T.y, st.hf[6 * b + 5] += T.z;
        }
        t1 = omp_get_wtime(), st.t[5] += t1 - t0, t0 = t1;
        // ComputeInvMassMatrix + ComputeImpulses (MC/solver/ChIterativeSolverMulticore.cpp:61-147, ...SMC.cpp:854-868):
        // a serial loop over the bodies and a serial (EIGEN_DONT_PARALLELIZE) block-diagonal product in the reference
        for (int i = 0; i < nb; i++) {
            if (!active[i])
                continue;
            const real im = 1.0 / st.mass[i];
            st.v[6 * i] = st.v[6 * i] + im * st.hf[6 * i];
            st.v[6 * i + 1] = st.v[6 * i + 1] + im * st.hf[6 * i + 1];
            st.v[6 * i + 2] = st.v[6 * i + 2] + im * st.hf[6 * i + 2];
            st.v[6 * i + 3] = st.v[6 * i + 3] + st.inv_inertia[i].x * st.hf[6 * i + 3];
            st.v[6 * i + 4] = st.v[6 * i + 4] + st.inv_inertia[i].y * st.hf[6 * i + 4];
            st.v[6 * i + 5] = st.v[6 * i + 5] + st.inv_inertia[i].z * st.hf[6 * i + 5];
        }
        t1 = omp_get_wtime(), st.t[6] += t1 - t0, t0 = t1;
        // scatter + ChBody::VariablesQbIncrementPosition (MC/physics/ChSystemMulticore.cpp:134-153, omp over bodies;
        // src/chrono/physics/ChBody.cpp:288-310; ChMatrix33::SetFromQuaternion, ChQuaternion::SetFromAngleAxis / Cross)
#pragma omp parallel for
        for (int i = 0; i < nb; i++) {
            if (!active[i])
                continue;
            const real* v = &st.v[6 * i];
            pos[i] = real3(pos[i].x + v[0] * h, pos[i].y + v[1] * h, pos[i].z + v[2] * h);
            const real e0 = rot[i].w, e1 = rot[i].x, e2 = rot[i].y, e3 = rot[i].z;
            real ax = ((e0 * e0 + e1 * e1) * 2 - 1) * v[3] + ((e1 * e2 - e0 * e3) * 2) * v[4] + ((e1 * e3 + e0 * e2) * 2) * v[5];
            real ay = ((e1 * e2 + e0 * e3) * 2) * v[3] + ((e0 * e0 + e2 * e2) * 2 - 1) * v[4] + ((e2 * e3 - e0 * e1) * 2) * v[5];
            real az = ((e1 * e3 - e0 * e2) * 2) * v[3] + ((e2 * e3 + e0 * e1) * 2) * v[4] + ((e0 * e0 + e3 * e3) * 2 - 1) * v[5];
            const real len = sqrt(ax * ax + ay * ay + az * az);
            const real angle = len * h;
            if (len < std::numeric_limits<real>::min()) {
                ax = 1, ay = 0, az = 0;
            } else {
                const real s = 1 / len;
                ax *= s, ay *= s, az *= s;
            }
            const real half = angle / 2, sh = sin(half);
            const real d0 = cos(half), d1 = ax * sh, d2 = ay * sh, d3 = az * sh;
            rot[i] = quaternion(d0 * e0 - d1 * e1 - d2 * e2 - d3 * e3, d0 * e1 + d1 * e0 - d3 * e2 + d2 * e3,
                                d0 * e2 + d2 * e0 + d3 * e1 - d1 * e3, d0 * e3 + d3 * e0 - d2 * e1 + d1 * e2);
        }
        t1 = omp_get_wtime(), st.t[7] += t1 - t0;
        st.steps++;
    }
}

void ref_smc_get_state(void* hh, double* pos, double* rot, double* vel) {
    auto* sys = static_cast<RefSystem*>(hh);
    auto& cd = *sys->cs.cd_data;
    for (int i = 0; i < sys->st.nb; i++) {
        const real3 p = (*cd.state_data.pos_rigid)[i];
        const quaternion q = (*cd.state_data.rot_rigid)[i];
        pos[3 * i] = p.x, pos[3 * i + 1] = p.y, pos[3 * i + 2] = p.z;
        rot[4 * i] = q.w, rot[4 * i + 1] = q.x, rot[4 * i + 2] = q.y, rot[4 * i + 3] = q.z;
    }
    memcpy(vel, sys->st.v.data(), sizeof(double) * 6 * (size_t)sys->st.nb);
}

// seconds per phase accumulated since ref_smc_setup (8 values, see StepState::t) and the step count
void ref_smc_get_timers(void* hh, double* t, long* steps) {
    auto& st = static_cast<RefSystem*>(hh)->st;
    for (int i = 0; i < 8; i++)
        t[i] = st.t[i];
    *steps = st.steps;
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
