// smc.cu -- SMC penalty-contact forces, per-body accumulation and the velocity update.
//
// Reference path replaced (MC = src/chrono_multicore):
//   composite materials      MC/collision/ChContactContainerMulticoreSMC.cpp:123-174 +
//                            src/chrono/physics/ChContactMaterialSMC.cpp:107-130  -> host table (float math) + lookup
//   function_CalcContactForces  MC/solver/ChIterativeSolverMulticoreSMC.cpp:56-552 -> k_smc_force
//   ProcessContacts sort_by_key + reduce_by_key + host_AddContactForces  :653-738   -> body->contact incidence lists
//                            (counting sort by body) + ordered per-body sums (k_body_reduce), masked block
//                            reduction for bodies with very many contacts (container walls)
//   ComputeInvMassMatrix / ComputeImpulses  MC/solver/ChIterativeSolverMulticore.cpp:61-147,
//                            ...SMC.cpp:854-868                                      -> k_vel_update
//
// Summation order: each body adds its contacts in ascending contact index, which is the order the reference's
// stable sort produces; no floating-point atomics anywhere, so results are run-to-run reproducible.
#include "common.cuh"
#include "narrow_single.cuh"
#include <float.h>

namespace cb200 {

#define CH_PI_D 3.141592653589793238462643383279
#define CH_2_3_D (2.0 / 3.0)

struct SmcParams {
    int force_model, adhesion_model, displ_mode, use_mat_props;
    double char_vel, min_slip_vel, min_roll_vel, min_spin_vel, dT;
    int nm;
    const double* mat_ovr;  // per-contact composite records set through an AddContactCallback (api.cu), else null
};

struct BodyView {
    d3 pos, v, w;
    q4 rot;
    double mass;
};

__device__ __forceinline__ BodyView load_body(uint32_t b, const double* __restrict__ pos,
                                              const double* __restrict__ rot, const double* __restrict__ vel,
                                              const double* __restrict__ mass) {
    BodyView B;
    B.pos = mk3(pos[3 * b], pos[3 * b + 1], pos[3 * b + 2]);
    const double4 q = reinterpret_cast<const double4*>(rot)[b];
    B.rot = mkq(q.x, q.y, q.z, q.w);
    const double2* v = reinterpret_cast<const double2*>(vel + 6 * (size_t)b);
    const double2 v0 = v[0], v1 = v[1], v2 = v[2];
    B.v = mk3(v0.x, v0.y, v1.x);
    B.w = mk3(v1.y, v2.x, v2.y);
    B.mass = mass[b];
    return B;
}

__device__ __forceinline__ BodyView load_body_nopos(uint32_t b, const double* __restrict__ rot,
                                                    const double* __restrict__ vel, const double* __restrict__ mass) {
    BodyView B;
    B.pos = mk3(0);
    const double4 q = reinterpret_cast<const double4*>(rot)[b];
    B.rot = mkq(q.x, q.y, q.z, q.w);
    const double2* v = reinterpret_cast<const double2*>(vel + 6 * (size_t)b);
    const double2 v0 = v[0], v1 = v[1], v2 = v[2];
    B.v = mk3(v0.x, v0.y, v1.x);
    B.w = mk3(v1.y, v2.x, v2.y);
    B.mass = mass[b];
    return B;
}

// Contact geometry evaluated inside the force kernel (GEN): in a step nobody reads the contact records, so writing
// them in k_contact_gen and reading them back here would be 200 bytes of traffic per contact for nothing.  The
// records are produced on demand instead (ensure_contact_geometry, pairs.cu).
struct GenArgs {
    const int64_t* c_shapes;
    const ShapeMeta* meta;
    const double *s_wpos, *s_wrot, *s_dims;
    const uint32_t* c_src;
    const double* cplx_out;
    double sep;
};

// MultiStep tangential displacement (ChSystemSMC::TangentialDisplacementModel::MultiStep): per body MAX_SHEAR history
// slots (MC/ChDataManager.h:198, host_container::shear_neigh / shear_disp / contact_relvel_init / contact_duration)
#define MAX_SHEAR 20
struct ShearArgs {
    const int64_t* keys;  // contact_shapeIDs: the shape pair of every contact
    int32_t* neigh;       // 3 per slot: partner body, larger shape id, smaller shape id; x == -1: free
    uint8_t* touch;       // slot used in this step
    double* disp;         // 3 per slot
    double* relvel0;      // |normal relative velocity| when the contact began
    double* duration;
};

// cold path: a contact that involves a non-sphere
__device__ __noinline__ void contact_general(uint32_t sA, int tA, double4 wA, uint32_t sB, int tB, double4 wB,
                                             double sep, const double* __restrict__ s_wrot,
                                             const double* __restrict__ s_dims, Contact& c) {
    prims_single(make_view(sA, tA, wA, s_wrot, s_dims), make_view(sB, tB, wB, s_wrot, s_dims), sep, c);
}

template <int FORCE_MODEL, bool USE_MAT, bool GEN, bool MULTI>
__global__ void __launch_bounds__(128, 4) k_smc_force(SmcParams P, GenArgs G, ShearArgs H,
                                                   const int32_t* __restrict__ bids,
                                                   const uint32_t* __restrict__ c_mats,
                                                   const double* __restrict__ c_norm, const double* __restrict__ c_pta,
                                                   const double* __restrict__ c_ptb, const double* __restrict__ c_depth,
                                                   const double* __restrict__ c_erad, const double* __restrict__ pos,
                                                   const double* __restrict__ rot, const double* __restrict__ vel,
                                                   const double* __restrict__ mass, const double* __restrict__ mat,
                                                   const uint8_t* __restrict__ heavy, uint32_t* __restrict__ deg,
                                                   double* __restrict__ ct_force, double* __restrict__ ct_torque,
                                                   DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nc = cnt->n_contacts;
    for (uint32_t i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < nc; i0 += gridDim.x * blockDim.x) {
    // MULTI: the contacts of one shape pair share one history slot; the reference's loop visits them in list order
    // (they are adjacent), so the thread of the pair's first contact walks the whole run
    uint32_t run = 1;
    if (MULTI) {
        const int64_t k0 = H.keys[i0];
        if (i0 > 0 && H.keys[i0 - 1] == k0)
            continue;
        while (i0 + run < nc && H.keys[i0 + run] == k0)
            run++;
    }
    for (uint32_t i = i0; i < i0 + run; i++) {
    d3 force = mk3(0), tq1 = mk3(0), tq2 = mk3(0);
    Contact C;
    uint32_t b1, b2, mm;
    // body and material ids: written beside the key by the store pass of the collision stage (or by k_contact_gen)
    const int2 bb = reinterpret_cast<const int2*>(bids)[i];
    b1 = (uint32_t)bb.x, b2 = (uint32_t)bb.y;
    mm = c_mats[i];
    // everything the force needs from the two bodies but the position, requested before the geometry is known (an
    // SMC contact list holds penetrating contacts almost exclusively)
    BodyView B1 = load_body_nopos(b1, rot, vel, mass), B2 = load_body_nopos(b2, rot, vel, mass);
    const bool hv1 = heavy[b1], hv2 = heavy[b2];
    bool cen1 = false, cen2 = false;  // body-centred sphere: the shape's world position IS the body position
    if (GEN) {
        // DispatchPRIMS on the compacted key list, as k_contact_gen does it
        const int64_t key = G.c_shapes[i];
        const uint32_t sA = (uint32_t)(key >> 32), sB = (uint32_t)(key & 0xffffffff);
        const ShapeMeta mA = G.meta[sA], mB = G.meta[sB];
        const double4 wA = reinterpret_cast<const double4*>(G.s_wpos)[sA];
        const double4 wB = reinterpret_cast<const double4*>(G.s_wpos)[sB];
        const int tA = (mA.info >> 8) & 0xff, tB = (mB.info >> 8) & 0xff;
        const uint32_t src = G.c_src ? G.c_src[i] : 0u;
        if (src) {
            const double* q = G.cplx_out + 11 * (size_t)(src - 1);
            C.n = mk3(q[0], q[1], q[2]), C.p1 = mk3(q[3], q[4], q[5]), C.p2 = mk3(q[6], q[7], q[8]);
            C.depth = q[9], C.erad = q[10];
        } else if (tA == CB200_SPHERE && tB == CB200_SPHERE) {
            sphere_sphere(mk3(wA.x, wA.y, wA.z), wA.w, mk3(wB.x, wB.y, wB.z), wB.w, G.sep, C);
        } else {
            Contact cg;  // only this one has its address taken: the hot path stays in registers
            contact_general(sA, tA, wA, sB, tB, wB, G.sep, G.s_wrot, G.s_dims, cg);
            C.n = cg.n, C.p1 = cg.p1, C.p2 = cg.p2, C.depth = cg.depth, C.erad = cg.erad;
        }
        cen1 = (mA.info & 4u) != 0, cen2 = (mB.info & 4u) != 0;
        if (cen1)
            B1.pos = mk3(wA.x, wA.y, wA.z);
        if (cen2)
            B2.pos = mk3(wB.x, wB.y, wB.z);
    } else {
        C.depth = c_depth[i];
    }
    const double depth = C.depth;
    // body -> contact incidence counts (first half of the counting sort by body that replaces the reference's
    // sort_by_key, ...SMC.cpp:702-703)
    if (!hv1)
        atomicAdd(&deg[b1], 1u);
    if (!hv2)
        atomicAdd(&deg[b2], 1u);
    if (depth < 0) {
        if (!cen1)
            B1.pos = mk3(pos[3 * b1], pos[3 * b1 + 1], pos[3 * b1 + 2]);
        if (!cen2)
            B2.pos = mk3(pos[3 * b2], pos[3 * b2 + 1], pos[3 * b2 + 2]);
        if (!GEN) {
            C.n = mk3(c_norm[3 * i], c_norm[3 * i + 1], c_norm[3 * i + 2]);
            C.p1 = mk3(c_pta[3 * i], c_pta[3 * i + 1], c_pta[3 * i + 2]);
            C.p2 = mk3(c_ptb[3 * i], c_ptb[3 * i + 1], c_ptb[3 * i + 2]);
            C.erad = c_erad[i];
        }
        const d3 n = C.n, pt1 = C.p1, pt2 = C.p2;
        const double erad = C.erad;

        // kinematics (:110-135)
        const d3 pt1_loc = parent_to_local(B1.pos, B1.rot, pt1);
        const d3 pt2_loc = parent_to_local(B2.pos, B2.rot, pt2);
        const d3 vel1 = B1.v + rotate(cross3(B1.w, pt1_loc), B1.rot);
        const d3 vel2 = B2.v + rotate(cross3(B2.w, pt2_loc), B2.rot);
        const d3 relvel = vel2 - vel1;
        const double relvel_n_mag = dot3(relvel, n);
        const d3 relvel_n = relvel_n_mag * n;
        const d3 relvel_t = relvel - relvel_n;
        const double relvel_t_mag = length3(relvel_t);

        // composite material (:140-160)
        const double m_eff = B1.mass * B2.mass / (B1.mass + B2.mass);
        const double* M = P.mat_ovr ? P.mat_ovr + (size_t)i * SMC_MAT_STRIDE
                                    : mat + ((size_t)(mm & 0xffff) * P.nm + (mm >> 16)) * SMC_MAT_STRIDE;
        const double mu_eff = M[0];
        double muRoll_eff = M[1], muSpin_eff = M[2];
        const double E_eff = M[3], G_eff = M[4];
        const double adhesion_eff = M[5], adhesionMultDMT_eff = M[6], adhesionSPerko_eff = M[7];
        double cr_eff = M[8];
        const double user_kn = M[9], user_kt = M[10], user_gn = M[11], user_gt = M[12];

        double kn = 0, kt = 0, gn = 0, gt = 0, kn_simple = 0, gn_simple = 0;
        double t_contact = 0;
        double relvel_init = fabs(relvel_n_mag);
        const double delta_n = -depth;
        d3 delta_t = mk3(0);
        if (P.displ_mode == 1)
            delta_t = relvel_t * P.dT;
        double char_vel = P.char_vel;
        size_t slot = 0;        // MULTI: this contact's history slot
        bool hist_on_b1 = false, have_slot = false;
        if (MULTI) {
            // contact history lives on the body with the larger id, keyed by (partner, larger shape, smaller shape)
            // (:188-250).  Which of the MAX_SHEAR slots a contact gets has no effect on the forces, so slots are
            // claimed with an atomic instead of the reference's first-free scan in contact order.
            delta_t = relvel_t * P.dT;
            const int64_t key = H.keys[i];
            const int s1 = (int)(key >> 32), s2 = (int)(key & 0xffffffff);
            const int sb1 = max((int)b1, (int)b2), sb2 = min((int)b1, (int)b2);
            const int ss1 = max(s1, s2), ss2 = min(s1, s2);
            int cid = -1;
            int32_t* nb_ = H.neigh + 3 * (size_t)MAX_SHEAR * sb1;
            for (int k = 0; k < MAX_SHEAR; k++) {
                if (nb_[3 * k] == sb2 && nb_[3 * k + 1] == ss1 && nb_[3 * k + 2] == ss2) {
                    H.duration[(size_t)MAX_SHEAR * sb1 + k] += P.dT;
                    cid = k;
                    break;
                }
            }
            if (cid < 0) {
                for (int k = 0; k < MAX_SHEAR; k++) {
                    // -2 while the claimer fills the key: neither free nor a match for a concurrent search
                    if (nb_[3 * k] == -1 && atomicCAS(&nb_[3 * k], -1, -2) == -1) {
                        cid = k;
                        const size_t sl = (size_t)MAX_SHEAR * sb1 + k;
                        nb_[3 * k + 1] = ss1, nb_[3 * k + 2] = ss2;
                        H.disp[3 * sl] = 0, H.disp[3 * sl + 1] = 0, H.disp[3 * sl + 2] = 0;
                        H.relvel0[sl] = relvel_init;
                        H.duration[sl] = 0;
                        __threadfence();
                        atomicExch(&nb_[3 * k], sb2);
                        break;
                    }
                }
            }
            have_slot = cid >= 0;
            if (!have_slot) {
                // more than MAX_SHEAR partners: the reference indexes slot -1 here (:232).  Reported as an error (the
                // counter is sticky); nobody else's history is touched: the contact is treated as a new one.
                atomicAdd(&cnt->unsupported, 1u);
                relvel_init = (relvel_init < char_vel) ? char_vel : relvel_init;
            } else {
                slot = (size_t)MAX_SHEAR * sb1 + cid;
                H.touch[slot] = 1;
                hist_on_b1 = sb1 == (int)b1;
                d3 sd = mk3(H.disp[3 * slot], H.disp[3 * slot + 1], H.disp[3 * slot + 2]);
                if (hist_on_b1) {
                    sd += delta_t;
                    sd -= dot3(sd, n) * n;
                    delta_t = sd;
                } else {
                    sd -= delta_t;
                    sd -= dot3(sd, n) * n;
                    delta_t = -sd;
                }
                H.disp[3 * slot] = sd.x, H.disp[3 * slot + 1] = sd.y, H.disp[3 * slot + 2] = sd.z;
                const double r0 = H.relvel0[slot];
                relvel_init = (r0 < char_vel) ? char_vel : r0;
                t_contact = H.duration[slot];
            }
        }
        const double eps = DBL_EPSILON;
        bool done = false;
        // kn_simple / gn_simple only decide whether rolling / spinning friction is switched off (:485-495);
        // without such friction in the material pair they are never looked at
        const bool need_simple = muRoll_eff > eps || muSpin_eff > eps;

        switch (FORCE_MODEL) {
            case 0:  // Hooke (:259-281)
                if (USE_MAT) {
                    const double tmp_k = (16.0 / 15) * sqrt(erad) * E_eff;
                    if (MULTI)
                        char_vel = relvel_init;  // :258
                    const double v2 = char_vel * char_vel;
                    double loge = (cr_eff < eps) ? log(eps) : log(cr_eff);
                    loge = (cr_eff > 1 - eps) ? log(1 - eps) : loge;
                    const double pl = CH_PI_D / loge;
                    const double tmp_g = 1 + pl * pl;
                    kn = tmp_k * pow(m_eff * v2 / tmp_k, 0.2);
                    kt = kn;
                    gn = sqrt(4 * m_eff * kn / tmp_g);
                    gt = gn;
                } else {
                    kn = user_kn;
                    kt = user_kt;
                    gn = m_eff * user_gn;
                    gt = m_eff * user_gt;
                }
                kn_simple = kn;
                gn_simple = gn;
                break;
            case 1:  // Hertz (:283-306)
                if (USE_MAT) {
                    const double sqrt_Rd = sqrt(erad * delta_n);
                    const double Sn = 2 * E_eff * sqrt_Rd;
                    const double St = 8 * G_eff * sqrt_Rd;
                    // beta = loge / sqrt(loge^2 + pi^2), loge = log(max(cr, eps)): a per-material-pair constant,
                    // tabulated on the host (M[13]) instead of one log + sqrt + division per contact
                    const double beta = M[13];
                    kn = CH_2_3_D * Sn;
                    kt = St;
                    gn = -2 * sqrt(5.0 / 6) * beta * sqrt(Sn * m_eff);
                    gt = -2 * sqrt(5.0 / 6) * beta * sqrt(St * m_eff);
                } else {
                    const double tmp = erad * sqrt(delta_n);
                    kn = tmp * user_kn;
                    kt = tmp * user_kt;
                    gn = tmp * m_eff * user_gn;
                    gt = tmp * m_eff * user_gt;
                }
                if (need_simple) {
                    kn_simple = kn / sqrt(delta_n);
                    gn_simple = gn / sqrt(sqrt(delta_n));  // delta_n^0.25
                }
                break;
            case 3:  // Flores (:308-334)
                if (USE_MAT) {
                    const double sqrt_Rd = sqrt(erad * delta_n);
                    const double Sn = 2 * E_eff * sqrt_Rd;
                    const double St = 8 * G_eff * sqrt_Rd;
                    cr_eff = (cr_eff < 0.01) ? 0.01 : cr_eff;
                    cr_eff = (cr_eff > 1.0 - eps) ? 1.0 - eps : cr_eff;
                    if (MULTI)
                        char_vel = relvel_init;  // :312
                    const double loge = log(cr_eff);
                    const double beta = loge / sqrt(loge * loge + CH_PI_D * CH_PI_D);
                    kn = CH_2_3_D * Sn;
                    kt = CH_2_3_D * St;
                    gn = 8.0 * (1.0 - cr_eff) * kn * delta_n / (5.0 * cr_eff * char_vel);
                    gt = -2 * sqrt(5.0 / 6) * beta * sqrt(St * m_eff);
                } else {
                    const double tmp = erad * sqrt(delta_n);
                    kn = tmp * user_kn;
                    kt = tmp * user_kt;
                    gn = tmp * m_eff * user_gn * delta_n;
                    gt = tmp * m_eff * user_gt;
                }
                kn_simple = kn / sqrt(delta_n);
                gn_simple = gn / pow(delta_n, 1.5);
                break;
            default:  // PlainCoulomb (:336-438)
                if (USE_MAT) {
                    const double sqrt_Rd = sqrt(delta_n);
                    const double Sn = 2 * E_eff * sqrt_Rd;
                    const double loge = (cr_eff < eps) ? log(eps) : log(cr_eff);
                    const double beta = loge / sqrt(loge * loge + CH_PI_D * CH_PI_D);
                    kn = CH_2_3_D * Sn;
                    gn = -2 * sqrt(5.0 / 6) * beta * sqrt(Sn * m_eff);
                } else {
                    const double tmp = sqrt(delta_n);
                    kn = tmp * user_kn;
                    gn = tmp * user_gn;
                }
                kn_simple = kn / sqrt(delta_n);
                gn_simple = gn / pow(delta_n, 0.25);
                kt = 0;
                gt = 0;
                {
                    const double forceN_mag = kn * delta_n - gn * relvel_n_mag;
                    const double forceT_mag = mu_eff * tanh(5.0 * relvel_t_mag) * forceN_mag;
                    force = forceN_mag * n;
                    if (relvel_t_mag >= P.min_slip_vel)
                        force -= (forceT_mag / relvel_t_mag) * relvel_t;
                }
                done = true;
                break;
        }

        double forceN_mag = kn * delta_n - gn * relvel_n_mag;
        if (!done) {
            // normal + tangential force with Coulomb clamp (:447-475)
            const d3 forceT_stiff = kt * delta_t;
            const d3 forceT_damp = gt * relvel_t;
            d3 forceT = forceT_stiff + forceT_damp;
            const double forceT_mag = length3(forceT);
            const double delta_t_mag = length3(delta_t);
            const double forceT_slide = mu_eff * fabs(forceN_mag);
            if (forceT_mag > forceT_slide) {
                if (delta_t_mag > eps) {
                    const double ratio = forceT_slide / forceT_mag;
                    forceT *= ratio;
                    if (MULTI && have_slot) {  // the stored displacement is rescaled to the Coulomb limit (:460-467)
                        delta_t = (forceT - forceT_damp) / kt;
                        const d3 sd = hist_on_b1 ? delta_t : -delta_t;
                        H.disp[3 * slot] = sd.x, H.disp[3 * slot + 1] = sd.y, H.disp[3 * slot + 2] = sd.z;
                    }
                } else {
                    forceT = mk3(0);
                }
            }
            force = forceN_mag * n - forceT;
        }

        // torques in the body frames (:482-483)
        const d3 torque1_loc = cross3(pt1_loc, rotate_t(force, B1.rot));
        const d3 torque2_loc = cross3(pt2_loc, rotate_t(force, B2.rot));

        // rolling / spinning friction only for persistent contacts (:488-530)
        if (need_simple) {
            const double d_coeff = gn_simple / (2.0 * m_eff * sqrt(kn_simple / m_eff));
            if (d_coeff < 1.0) {
                const double t_collision = CH_PI_D * sqrt(m_eff / (kn_simple * (1 - d_coeff * d_coeff)));
                if (t_contact <= t_collision) {
                    muRoll_eff = 0.0;
                    muSpin_eff = 0.0;
                }
            }
        }
        d3 m_roll1 = mk3(0), m_roll2 = mk3(0), m_spin1 = mk3(0), m_spin2 = mk3(0);
        if (muRoll_eff > eps || muSpin_eff > eps) {
            const d3 v_rot = rotate(cross3(B2.w, pt2_loc), B2.rot) - rotate(cross3(B1.w, pt1_loc), B1.rot);
            const d3 rel_o = rotate(B2.w, B2.rot) - rotate(B1.w, B1.rot);
            if (length3(v_rot) > P.min_roll_vel && muRoll_eff > eps) {
                m_roll1 = muRoll_eff * cross3(forceN_mag * pt1_loc, rotate_t(v_rot, B1.rot)) / length3(v_rot);
                m_roll2 = muRoll_eff * cross3(forceN_mag * pt2_loc, rotate_t(v_rot, B2.rot)) / length3(v_rot);
            }
            if (length3(rel_o) > P.min_spin_vel && muSpin_eff > eps) {
                const double r1 = length3(pt1_loc), r2 = length3(pt2_loc);
                const double xc = (r1 * r1 - r2 * r2) / (2 * (r1 + r2 - delta_n)) + 0.5 * (r1 + r2 - delta_n);
                double rc = r1 * r1 - xc * xc;
                rc = (rc < eps) ? eps : sqrt(rc);
                const d3 fn = forceN_mag * n;
                m_spin1 = muSpin_eff * rc * rotate_t(dot3(rel_o, fn) * n, B1.rot) / length3(rel_o);
                m_spin2 = muSpin_eff * rc * rotate_t(dot3(rel_o, fn) * n, B2.rot) / length3(rel_o);
            }
        }

        // adhesion is applied after the torques were formed (:533-543)
        if (P.adhesion_model == 0)
            force -= adhesion_eff * n;
        else if (P.adhesion_model == 1)
            force -= adhesionMultDMT_eff * sqrt(erad) * n;
        else
            force -= adhesionSPerko_eff * erad * n;

        tq1 = -torque1_loc + m_roll1 + m_spin1;
        tq2 = torque2_loc - m_roll2 - m_spin2;
    }
    // one 96-byte record per contact, three 32-byte sectors: (force, -) (torque on body 1, -) (torque on body 2, -);
    // the per-body gather then touches exactly two whole sectors per incidence
    double4* o = reinterpret_cast<double4*>(ct_force + CT_STRIDE * (size_t)i);
    o[0] = make_double4(force.x, force.y, force.z, 0.0);
    o[1] = make_double4(tq1.x, tq1.y, tq1.z, 0.0);
    o[2] = make_double4(tq2.x, tq2.y, tq2.z, 0.0);
    }
    }
}

// MultiStep: history slots that saw no contact in this step are released (...SMC.cpp:687-695)
__global__ void __launch_bounds__(256) k_shear_sweep(uint32_t nslots, const uint8_t* __restrict__ touch,
                                                     int32_t* __restrict__ neigh,
                                                     const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nslots && !touch[i])
        neigh[3 * (size_t)i] = -1;
}

// ---- body -> contact incidence lists (counting sort by body id) ------------------------------------------------
// The scanned offsets double as the cursors: every incidence advances its body's offset, so that afterwards off[b] is
// the END of body b's list and off[b - 1] its start (one random 4-byte access per incidence less than with a separate
// cursor array, which also would have to be cleared).
__global__ void __launch_bounds__(256) k_inc_fill(const int32_t* __restrict__ bids,
                                                  const uint8_t* __restrict__ heavy, uint32_t* __restrict__ off,
                                                  uint32_t* __restrict__ inc, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nc = cnt->n_contacts;
    const int lane = threadIdx.x & 31;
    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t i0 = blockIdx.x * blockDim.x + threadIdx.x - lane; i0 < nc; i0 += stride) {
        const uint32_t i = i0 + lane;
        const bool have = i < nc;
        const int2 bb = have ? reinterpret_cast<const int2*>(bids)[i] : make_int2(-1, -1);
        const uint32_t b1 = (uint32_t)bb.x, b2 = (uint32_t)bb.y;
        // Side 1: the list is ordered (bin, shape A, shape B), so the contacts a shape has as A inside one bin sit next
        // to each other: one atomic per run of equal body ids in the warp instead of one per contact
        const bool on1 = have && !heavy[b1], on2 = have && !heavy[b2];
        const uint32_t prev = __shfl_up_sync(0xffffffffu, b1, 1);
        const bool prev_on = __shfl_up_sync(0xffffffffu, (int)on1, 1) != 0;  // (every lane takes part: not inside the || below)
        const bool head = on1 && (lane == 0 || prev != b1 || !prev_on);
        const uint32_t heads = __ballot_sync(0xffffffffu, head || !on1);  // a lane without a slot ends a run too
        const uint32_t upto = (2u << lane) - 1u;  // lanes 0..lane (lane 31: 2u << 31 == 0, all ones)
        const int start = 31 - __clz((int)(heads & upto));
        uint32_t base = 0;
        if (head) {
            const uint32_t above = heads & ~upto;
            const int end = above ? __ffs((int)above) - 1 : 32;
            base = atomicAdd(&off[b1], (uint32_t)(end - lane));
        }
        const uint32_t slot2 = on2 ? atomicAdd(&off[b2], 1u) : 0u;  // (in flight together with the one above)
        base = __shfl_sync(0xffffffffu, base, start);
        if (on1)
            inc[base + (uint32_t)(lane - start)] = i << 1;
        if (on2)
            inc[slot2] = (i << 1) | 1u;
    }
}

// What k_update_forces (integrate.cu) reads: with UPD the body's hf = h (m g + F_applied, T_applied - w x J w) is
// formed here, by the same operations, instead of being written by that kernel and read back.
struct UpdArgs {
    d3 g;
    const double *mass, *inertia, *vel, *fapp;
    const double* inertia_off;  // products of inertia (null: none)
};

// F, T (+)= the record of incidence `code` (contact << 1 | side): force on body 2 of the contact (body 1 gets the
// negative), torque on the coded side
__device__ __forceinline__ void add_record(const double* __restrict__ ct_force, uint32_t code, bool first, d3& F, d3& T) {
    const double4* rec = reinterpret_cast<const double4*>(ct_force + CT_STRIDE * (size_t)(code >> 1));
    const double4 f4 = rec[0], t4 = rec[1 + (code & 1u)];
    const d3 f = mk3(f4.x, f4.y, f4.z);
    if (first) {
        F = (code & 1u) ? f : -f;
        T = mk3(t4.x, t4.y, t4.z);
    } else {
        F = F + ((code & 1u) ? f : -f);
        T = T + mk3(t4.x, t4.y, t4.z);
    }
}

// Batcher's odd-even merge sort as a fixed network over registers (N a power of two; all indices are compile-time)
template <int N>
__device__ __forceinline__ void reduce_sorted(const uint32_t* __restrict__ codes, uint32_t d,
                                              const double* __restrict__ ct_force, d3& F, d3& T) {
    uint32_t c[N];
#pragma unroll
    for (int k = 0; k < N; k++)
        c[k] = (uint32_t)k < d ? codes[k] : 0xffffffffu;
#pragma unroll
    for (int p = 1; p < N; p <<= 1)
#pragma unroll
        for (int k = p; k >= 1; k >>= 1)
#pragma unroll
            for (int j = k % p; j + k < N; j += 2 * k)
#pragma unroll
                for (int i = 0; i < k; i++)
                    if (i + j + k < N && (i + j) / (2 * p) == (i + j + k) / (2 * p)) {
                        const uint32_t lo = min(c[i + j], c[i + j + k]), hi = max(c[i + j], c[i + j + k]);
                        c[i + j] = lo, c[i + j + k] = hi;
                    }
#pragma unroll
    for (int r = 0; r < N; r++)
        if ((uint32_t)r < d)
            add_record(ct_force, c[r], r == 0, F, T);
}

// hf of a body before its contacts: h (m g + F_applied, T_applied - w x J w), the operations of k_update_forces
__device__ __forceinline__ void base_hf(uint32_t b, const UpdArgs& U, double h, d3& bF, d3& bT) {
    const double m = U.mass[b];
    const double* v = U.vel + 6 * (size_t)b;
    const d3 w = mk3(v[3], v[4], v[5]);
    const d3 Jw = sym3_apply(U.inertia, U.inertia_off, b, w);
    const d3 gyro = cross3(w, Jw);
    d3 F = U.g * m;
    d3 T = mk3(0);
    if (U.fapp) {
        const double* a = U.fapp + 6 * (size_t)b;
        F += mk3(a[0], a[1], a[2]);
        T += mk3(a[3], a[4], a[5]);
    }
    bF = h * F;
    bT = h * (T - gyro);
}

// one thread per body: add the body's contacts in ascending contact index, then hf += h * (F, T)
template <bool UPD>
__global__ void __launch_bounds__(256) k_body_reduce(uint32_t nb, const uint8_t* __restrict__ heavy,
                                                     const uint32_t* __restrict__ off, const uint32_t* __restrict__ inc,
                                                     const double* __restrict__ ct_force,
                                                     const double* __restrict__ ct_torque, double h,
                                                     double* __restrict__ b_force, double* __restrict__ hf,
                                                     UpdArgs U, DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    bool touched = false;
    d3 bF = mk3(0), bT = mk3(0);
    if (UPD && b < nb && !heavy[b])
        base_hf(b, U, h, bF, bT);  // a container body's hf is formed by k_heavy_final<true>
    if (b < nb && !heavy[b]) {
        const uint32_t s = b ? off[b - 1] : 0u, d = off[b] - s;  // k_inc_fill moved the offsets to the list ends
        d3 F = mk3(0), T = mk3(0);
        // ascending contact code == the order the reference's stable sort leaves.  Up to 16 incidences (every sphere
        // of a bed): the codes are read once, sorted in registers by a fixed network, and the records gathered with
        // constant register indices (the compiler keeps several gathers in flight).  Longer lists: selection walk.
        if (d <= 8)
            reduce_sorted<8>(inc + s, d, ct_force, F, T);
        else if (d <= 16)
            reduce_sorted<16>(inc + s, d, ct_force, F, T);
        else {
            uint32_t last = 0;
            for (uint32_t r = 0; r < d; r++) {
                uint32_t best = 0xffffffffu;
                for (uint32_t k = 0; k < d; k++) {
                    const uint32_t c = inc[s + k];
                    if ((r == 0 || c > last) && c < best)
                        best = c;
                }
                last = best;
                add_record(ct_force, best, r == 0, F, T);
            }
        }
        double* o = b_force + 6 * (size_t)b;
        o[0] = F.x, o[1] = F.y, o[2] = F.z, o[3] = T.x, o[4] = T.y, o[5] = T.z;
        double* q = hf + 6 * (size_t)b;
        if (d) {
            touched = true;
            const d3 hF = h * F, hT = h * T;
            if (UPD) {
                q[0] = bF.x + hF.x, q[1] = bF.y + hF.y, q[2] = bF.z + hF.z;
                q[3] = bT.x + hT.x, q[4] = bT.y + hT.y, q[5] = bT.z + hT.z;
            } else {
                q[0] += hF.x, q[1] += hF.y, q[2] += hF.z, q[3] += hT.x, q[4] += hT.y, q[5] += hT.z;
            }
        } else if (UPD) {
            q[0] = bF.x, q[1] = bF.y, q[2] = bF.z, q[3] = bT.x, q[4] = bT.y, q[5] = bT.z;
        }
    }
    const unsigned m = __ballot_sync(0xffffffffu, touched);
    if (m && (threadIdx.x & 31) == 0)
        atomicAdd(&cnt->ct_body_count, (uint32_t)__popc(m));
}

__global__ void __launch_bounds__(256) k_heavy_list(uint32_t nb, const uint8_t* __restrict__ heavy,
                                                    uint32_t* __restrict__ list, DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < nb && heavy[b]) {
        const uint32_t k = atomicAdd(&cnt->n_heavy, 1u);
        if (k < MAX_HEAVY)
            list[k] = b;
        else
            atomicOr(&cnt->overflow, (uint32_t)OVF_HEAVY);
    }
}

// Bodies with very many contacts (a container touching 10^5 spheres): fixed-shape masked reduction over the whole
// contact list.  Chunk -> CTA mapping and the in-CTA tree are fixed, hence deterministic.
constexpr int HEAVY_CHUNK = 4096;
__global__ void __launch_bounds__(256) k_heavy_partial(const int32_t* __restrict__ bids,
                                                       const double* __restrict__ ct_force,
                                                       const double* __restrict__ ct_torque,
                                                       const uint32_t* __restrict__ list,
                                                       const DevCounters* __restrict__ cnt, uint32_t chunk_cap,
                                                       double* __restrict__ part) {
    __shared__ double sm[8][7];
    if (cnt->overflow)
        return;
    const uint32_t nc = cnt->n_contacts;
    const uint32_t nchunks = (nc + HEAVY_CHUNK - 1) / HEAVY_CHUNK;
    const uint32_t nh = min(cnt->n_heavy, (uint32_t)MAX_HEAVY);
    for (uint32_t hI = blockIdx.y; hI < nh; hI += gridDim.y)
    for (uint32_t chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x) {
        const uint32_t body = list[hI];
        double acc[7] = {0, 0, 0, 0, 0, 0, 0};
        const uint32_t c0 = chunk * HEAVY_CHUNK;
        // all body-id pairs of the thread first (16 independent coalesced loads), then the few matching records
        int2 bb[HEAVY_CHUNK / 256];
#pragma unroll
        for (int k = 0; k < HEAVY_CHUNK / 256; k++) {
            const uint32_t c = c0 + k * 256 + threadIdx.x;
            bb[k] = c < nc ? reinterpret_cast<const int2*>(bids)[c] : make_int2(-1, -1);
        }
#pragma unroll
        for (int k = 0; k < HEAVY_CHUNK / 256; k++) {
            const uint32_t c = c0 + k * 256 + threadIdx.x;
            if (c < nc) {
                const uint32_t b1 = (uint32_t)bb[k].x, b2 = (uint32_t)bb[k].y;
                const double* rec = ct_force + CT_STRIDE * (size_t)c;
                if (b1 == body) {
                    acc[0] -= rec[0], acc[1] -= rec[1], acc[2] -= rec[2];
                    const double* t = rec + 4;
                    acc[3] += t[0], acc[4] += t[1], acc[5] += t[2];
                    acc[6] += 1;
                }
                if (b2 == body) {
                    acc[0] += rec[0], acc[1] += rec[1], acc[2] += rec[2];
                    const double* t = rec + 8;
                    acc[3] += t[0], acc[4] += t[1], acc[5] += t[2];
                    acc[6] += 1;
                }
            }
        }
#pragma unroll
        for (int q = 0; q < 7; q++)
#pragma unroll
            for (int d = 16; d > 0; d >>= 1)
                acc[q] += __shfl_down_sync(0xffffffffu, acc[q], d);
        const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
        if (lane == 0)
            for (int q = 0; q < 7; q++)
                sm[wid][q] = acc[q];
        __syncthreads();
        if (threadIdx.x < 7) {
            double s = sm[0][threadIdx.x];
            for (int w = 1; w < 8; w++)
                s += sm[w][threadIdx.x];
            part[((size_t)hI * chunk_cap + chunk) * 7 + threadIdx.x] = s;
        }
        __syncthreads();
    }
}

template <bool UPD>
__global__ void __launch_bounds__(256) k_heavy_final(const uint32_t* __restrict__ list,
                                                     const DevCounters* __restrict__ cnt_in, uint32_t chunk_cap,
                                                     const double* __restrict__ part, double h,
                                                     double* __restrict__ b_force, double* __restrict__ hf,
                                                     UpdArgs U, DevCounters* __restrict__ cnt) {
    __shared__ double sm[8][7];
    if (cnt_in->overflow)
        return;
    const uint32_t nchunks = (cnt_in->n_contacts + HEAVY_CHUNK - 1) / HEAVY_CHUNK;
    const uint32_t nh = min(cnt_in->n_heavy, (uint32_t)MAX_HEAVY);
    const uint32_t hI = blockIdx.x;
    if (hI >= nh)
        return;
    double acc[7] = {0, 0, 0, 0, 0, 0, 0};
    for (uint32_t k = threadIdx.x; k < nchunks; k += 256)
        for (int q = 0; q < 7; q++)
            acc[q] += part[((size_t)hI * chunk_cap + k) * 7 + q];
#pragma unroll
    for (int q = 0; q < 7; q++)
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
            acc[q] += __shfl_down_sync(0xffffffffu, acc[q], d);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0)
        for (int q = 0; q < 7; q++)
            sm[wid][q] = acc[q];
    __syncthreads();
    if (threadIdx.x == 0) {
        double s[7];
        for (int q = 0; q < 7; q++) {
            s[q] = sm[0][q];
            for (int w = 1; w < 8; w++)
                s[q] += sm[w][q];
        }
        const uint32_t body = list[hI];
        double* o = b_force + 6 * (size_t)body;
        double* f = hf + 6 * (size_t)body;
        if (UPD) {  // whole step: nobody has written this body's hf yet (k_body_reduce<true> leaves it out)
            d3 bF, bT;
            base_hf(body, U, h, bF, bT);
            const double base[6] = {bF.x, bF.y, bF.z, bT.x, bT.y, bT.z};
            for (int q = 0; q < 6; q++) {
                o[q] = s[q];
                f[q] = s[6] > 0 ? base[q] + h * s[q] : base[q];
            }
        } else {
            for (int q = 0; q < 6; q++) {
                o[q] = s[q];
                if (s[6] > 0)
                    f[q] += h * s[q];
            }
        }
        if (s[6] > 0)
            atomicAdd(&cnt->ct_body_count, 1u);
    }
}

// M_invk = v + M_inv * hf for active bodies (rows of inactive bodies do not exist)
__global__ void __launch_bounds__(256) k_vel_update(uint32_t nb, const uint8_t* __restrict__ bflags,
                                                    const double* __restrict__ mass,
                                                    const double* __restrict__ inv_inertia,
                                                    const double* __restrict__ inv_inertia_off /* may be null */,
                                                    const double* __restrict__ hf, double* __restrict__ vel,
                                                    const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb || !(bflags[b] & 1u))
        return;
    const double inv_mass = 1.0 / mass[b];
    const double* f = hf + 6 * (size_t)b;
    double* v = vel + 6 * (size_t)b;
    v[0] = v[0] + inv_mass * f[0];
    v[1] = v[1] + inv_mass * f[1];
    v[2] = v[2] + inv_mass * f[2];
    const d3 dw = sym3_apply(inv_inertia, inv_inertia_off, b, mk3(f[3], f[4], f[5]));
    v[3] = v[3] + dw.x;
    v[4] = v[4] + dw.y;
    v[5] = v[5] + dw.z;
}

// History arrays for MultiStep: slots of new bodies start free with zero displacement
// (ChSystemMulticoreSMC::AddMaterialSurfaceData, MC/physics/ChSystemMulticoreSMC.cpp:39-50).
int alloc_shear(System* S) {
    const size_t nb = S->nb, n0 = S->sh_bodies;
    if (nb <= n0)
        return CB200_OK;
    if (S->capturing)
        return set_error(CB200_ERR_STATE, "shear history must be allocated before a step is captured");
    if (3 * MAX_SHEAR * nb > S->sh_neigh.cap)
        invalidate_step_graph(S);
    cudaStream_t st = S->stream;
    CB_TRY(S->sh_neigh.reserve(3 * MAX_SHEAR * nb, true, st));
    CB_TRY(S->sh_touch.reserve(MAX_SHEAR * nb, true, st));
    CB_TRY(S->sh_disp.reserve(3 * MAX_SHEAR * nb, true, st));
    CB_TRY(S->sh_relvel.reserve(MAX_SHEAR * nb, true, st));
    CB_TRY(S->sh_dur.reserve(MAX_SHEAR * nb, true, st));
    const size_t a = MAX_SHEAR * n0, n = MAX_SHEAR * (nb - n0);
    CB_CUDA(cudaMemsetAsync(S->sh_neigh.p + 3 * a, 0xff, 3 * n * sizeof(int32_t), st));  // (-1, -1, -1)
    CB_CUDA(cudaMemsetAsync(S->sh_disp.p + 3 * a, 0, 3 * n * sizeof(double), st));
    CB_CUDA(cudaMemsetAsync(S->sh_relvel.p + a, 0, n * sizeof(double), st));
    CB_CUDA(cudaMemsetAsync(S->sh_dur.p + a, 0, n * sizeof(double), st));
    S->sh_bodies = (uint32_t)nb;
    return CB200_OK;
}

int enqueue_smc(System* S) {
    if (S->nb == 0)
        return CB200_OK;
    const uint32_t nb = S->nb;
    const cb200_settings& st = S->set;
    const bool multi = st.tangential_displ_mode == 2;
    if (multi && S->slab_on)
        return set_error(CB200_ERR_UNSUPPORTED, "MultiStep contact history does not migrate between slabs yet");
    ShearArgs H = {};
    if (multi) {
        CB_TRY(alloc_shear(S));
        CB_TRY(ensure_contact_geometry(S));  // the history variant reads the contact records
        H.keys = S->c_shapes.p, H.neigh = S->sh_neigh.p, H.touch = S->sh_touch.p, H.disp = S->sh_disp.p;
        H.relvel0 = S->sh_relvel.p, H.duration = S->sh_dur.p;
        CB_CUDA(cudaMemsetAsync(S->sh_touch.p, 0, (size_t)MAX_SHEAR * nb, S->stream));
    }
    if (S->ns > 0 && S->nm_smc == 0)
        return set_error(CB200_ERR_STATE, "cb200_set_materials_smc was not called");
    SmcParams P;
    P.force_model = st.contact_force_model;
    P.adhesion_model = st.adhesion_force_model;
    P.displ_mode = st.tangential_displ_mode;
    P.use_mat_props = st.use_material_properties;
    P.char_vel = st.characteristic_vel;
    P.min_slip_vel = st.min_slip_vel;
    P.min_roll_vel = st.min_roll_vel;
    P.min_spin_vel = st.min_spin_vel;
    P.dT = st.step_size;
    P.nm = S->nm_smc;
    P.mat_ovr = S->mat_ovr_valid ? S->mat_ovr.p : nullptr;
    const uint32_t chunk_cap = S->cap_contacts / HEAVY_CHUNK + 1;
    const int cgrid = grid_for(S, S->cap_contacts, 256);
    if (!S->body_scratch_clean) {
        CB_CUDA(cudaMemsetAsync(S->b_deg.p, 0, (size_t)(nb + 1) * 4, S->stream));
    }
    S->body_scratch_clean = false;
    const int fgrid = grid_for(S, S->cap_contacts, 128, 16);
    GenArgs G;
    G.c_shapes = S->c_shapes.p, G.meta = S->s_meta.p, G.s_wpos = S->s_wpos.p, G.s_wrot = S->s_wrot.p;
    G.s_dims = S->s_dims.p, G.c_src = use_complex(S) ? S->c_src.p : nullptr, G.cplx_out = S->cplx_out.p;
    G.sep = 2 * S->set.collision_envelope;
    const bool gen = !S->contacts_geom_valid;  // the step skipped k_contact_gen: evaluate the geometry here
#define SMC_LAUNCH2(FM, UM, GN, MS)                                                                       \
    k_smc_force<FM, UM, GN, MS><<<fgrid, 128, 0, S->stream>>>(                                            \
        P, G, H, S->c_bids.p, S->c_mats.p, S->c_norm.p, S->c_pta.p, S->c_ptb.p, S->c_depth.p, S->c_erad.p, \
        S->pos.p, S->rot.p, S->vel.p, S->mass.p, S->mat_smc.p, S->heavy.p, S->b_deg.p, S->ct_force.p,     \
        S->ct_torque.p, S->d_cnt)
#define SMC_LAUNCH(FM, UM)         \
    do {                           \
        if (multi)                 \
            SMC_LAUNCH2(FM, UM, false, true); \
        else if (gen)              \
            SMC_LAUNCH2(FM, UM, true, false);  \
        else                       \
            SMC_LAUNCH2(FM, UM, false, false); \
    } while (0)
    // one instantiation per force model: the runtime switch of the reference (...SMC.cpp:258-440) costs registers
    const bool um = P.use_mat_props != 0;
    CB_TRY(timer_start(S, "k_smc_force"));
    switch (P.force_model) {
        case 0: if (um) SMC_LAUNCH(0, true); else SMC_LAUNCH(0, false); break;
        case 1: if (um) SMC_LAUNCH(1, true); else SMC_LAUNCH(1, false); break;
        case 3: if (um) SMC_LAUNCH(3, true); else SMC_LAUNCH(3, false); break;
        default: if (um) SMC_LAUNCH(2, true); else SMC_LAUNCH(2, false); break;
    }
#undef SMC_LAUNCH
#undef SMC_LAUNCH2
    CB_TRY(timer_stop(S, "k_smc_force"));
    S->launches++;
    if (multi) {
        k_shear_sweep<<<blocks_for(MAX_SHEAR * nb, 256), 256, 0, S->stream>>>(MAX_SHEAR * nb, S->sh_touch.p,
                                                                              S->sh_neigh.p, S->d_cnt);
        S->launches++;
    }
    // side branch: the partial sums of container bodies need the force records only
    cudaStream_t hs;
    CB_TRY(aux_fork(S, 0, &hs));
    k_heavy_list<<<blocks_for(nb, 256), 256, 0, hs>>>(nb, S->heavy.p, S->heavy_list.p, S->d_cnt);
    k_heavy_partial<<<dim3(grid_for(S, chunk_cap, 1, 4), 4), 256, 0, hs>>>(
        S->c_bids.p, S->ct_force.p, S->ct_torque.p, S->heavy_list.p, S->d_cnt, chunk_cap, S->heavy_part.p);
    UpdArgs U;
    U.g = mk3(st.gravity[0], st.gravity[1], st.gravity[2]);
    U.mass = S->mass.p, U.inertia = S->inertia.p, U.vel = S->vel.p, U.fapp = S->has_fapp ? S->fapp.p : nullptr;
    U.inertia_off = S->inertia_off.p;
    if (S->defer_update_forces)  // the body's whole hf is formed there: no dependence on k_body_reduce
        k_heavy_final<true><<<MAX_HEAVY, 256, 0, hs>>>(S->heavy_list.p, S->d_cnt, chunk_cap, S->heavy_part.p,
                                                       st.step_size, S->b_force.p, S->hf.p, U, S->d_cnt);
    if (integrate_fuses_aabb(S))
        CB_TRY(enqueue_reset_bbox(S, hs));
    CB_TRY(scan_u32(S, S->b_deg.p, S->b_deg.p, &S->d_cnt->n_bodies, nb));
    CB_TRY(timer_start(S, "k_inc_fill"));
    k_inc_fill<<<cgrid, 256, 0, S->stream>>>(S->c_bids.p, S->heavy.p, S->b_deg.p, S->b_inc.p, S->d_cnt);
    CB_TRY(timer_stop(S, "k_inc_fill"));
    CB_TRY(timer_start(S, "k_body_reduce"));
    if (S->defer_update_forces)
        k_body_reduce<true><<<blocks_for(nb, 256), 256, 0, S->stream>>>(nb, S->heavy.p, S->b_deg.p, S->b_inc.p,
                                                                        S->ct_force.p, S->ct_torque.p, st.step_size,
                                                                        S->b_force.p, S->hf.p, U, S->d_cnt);
    else
        k_body_reduce<false><<<blocks_for(nb, 256), 256, 0, S->stream>>>(nb, S->heavy.p, S->b_deg.p, S->b_inc.p,
                                                                         S->ct_force.p, S->ct_torque.p, st.step_size,
                                                                         S->b_force.p, S->hf.p, U, S->d_cnt);
    CB_TRY(timer_stop(S, "k_body_reduce"));
    CB_TRY(aux_join(S, 0));
    if (!S->defer_update_forces)  // hf += h (F, T) on top of what k_update_forces wrote
        k_heavy_final<false><<<MAX_HEAVY, 256, 0, S->stream>>>(S->heavy_list.p, S->d_cnt, chunk_cap, S->heavy_part.p,
                                                               st.step_size, S->b_force.p, S->hf.p, U, S->d_cnt);
    CB_TRY(enqueue_bilateral(S));  // joints / motors: hf += D_b gamma_b (RunTimeStep :816-842)
    if (!S->defer_vel_update)
        k_vel_update<<<blocks_for(nb, 256), 256, 0, S->stream>>>(nb, S->bflags.p, S->mass.p, S->inv_inertia.p,
                                                                 S->inv_inertia_off.p, S->hf.p,
                                                                 S->vel.p, S->d_cnt);
    S->launches += S->defer_vel_update ? 5 : 6;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

}  // namespace cb200
