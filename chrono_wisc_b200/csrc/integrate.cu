// integrate.cu -- per-body force gathering and time integration on the device.
//
// Reference path replaced:
//   ChSystemMulticore::UpdateRigidBodies   MC/physics/ChSystemMulticore.cpp:329-377 with ChBody::Update,
//       UpdateForces, ComputeGyro, VariablesFbLoadForces (src/chrono/physics/ChBody.cpp:552-594,423-426,247-257)
//                                                                      -> k_update_forces
//   scatter + ChBody::VariablesQbIncrementPosition  MC/physics/ChSystemMulticore.cpp:134-153,
//       src/chrono/physics/ChBody.cpp:288-310                           -> k_integrate
//
// The reference walks one heap object per body on the host every step; here the state never leaves the device
// and both loops are streaming kernels (one thread per body).
#include "common.cuh"
#include <float.h>

namespace cb200 {

// hf = h * (m g + F_applied ,  T_applied - w x (J w))
__global__ void __launch_bounds__(256) k_update_forces(uint32_t nb, double h, d3 g, const double* __restrict__ mass,
                                                       const double* __restrict__ inertia,
                                                       const double* __restrict__ inertia_off /* may be null */,
                                                       const double* __restrict__ vel,
                                                       const double* __restrict__ fapp /* may be null */,
                                                       double* __restrict__ hf, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb)
        return;
    const double m = mass[b];
    const double* v = vel + 6 * (size_t)b;
    const d3 w = mk3(v[3], v[4], v[5]);
    // ComputeGyro: gyro = w x (J w), J in the body frame (diagonal unless the body was given products of inertia)
    const d3 Jw = sym3_apply(inertia, inertia_off, b, w);
    const d3 gyro = cross3(w, Jw);
    d3 F = g * m;   // UpdateForces: Xforce = G_acc * mass
    d3 T = mk3(0);
    if (fapp) {
        const double* a = fapp + 6 * (size_t)b;
        F += mk3(a[0], a[1], a[2]);
        T += mk3(a[3], a[4], a[5]);
    }
    const d3 hF = h * F;
    const d3 hT = h * (T - gyro);
    double* o = hf + 6 * (size_t)b;
    o[0] = hF.x, o[1] = hF.y, o[2] = hF.z, o[3] = hT.x, o[4] = hT.y, o[5] = hT.z;
}

// x += h v ;  q <- dq(h R w) * q   (ChBody::VariablesQbIncrementPosition), active bodies only
// VEL: the velocity update of the SMC solve (k_vel_update, smc.cu: v += M^-1 hf, same operations) is done here, in
// the same pass over the bodies -- used by the whole-step paths, where nothing reads the velocities in between.
// AABB: whole-step paths again -- the body's own centred sphere gets its world record (obj_data_A_global + radius) and
// its box joins the global bounding box of the NEXT step right here, with the operations of k_aabb<false>
// (broadphase.cu), so that step does not read the positions once more just for that.
struct AabbArgs {
    const int32_t* b_sphere;
    const double* b_rad;
    double* s_wpos;
    double envelope;
    // per-body scratch of the step that ends here, cleared for the next one (common.cuh body_scratch_clean)
    uint32_t* b_deg;
    uint8_t* heavy;
};

template <bool VEL, bool AABB>
__global__ void __launch_bounds__(256, 4) k_integrate(uint32_t nb, double h, const uint8_t* __restrict__ bflags,
                                                   double* __restrict__ vel, double* __restrict__ pos,
                                                   double* __restrict__ rot, const double* __restrict__ mass,
                                                   const double* __restrict__ inv_inertia,
                                                   const double* __restrict__ inv_inertia_off /* may be null */,
                                                   const double* __restrict__ hf, AabbArgs A,
                                                   const double* __restrict__ vlim /* (max_speed, max_wvel) per body or null */,
                                                   DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b == 0)
        cnt->steps_done++;  // this step was applied in full (nothing upstream overflowed)
    const uint32_t fl = b < nb ? bflags[b] : 0u;
    d3 bmn = mk3(DBL_MAX), bmx = mk3(-DBL_MAX);
    bool counted = false;
    if (b < nb) {
        double* p = pos + 3 * (size_t)b;
        d3 pn = mk3(p[0], p[1], p[2]);
        if (fl & 1u) {
            double* v = vel + 6 * (size_t)b;
            if (VEL) {
                const double inv_mass = 1.0 / mass[b];
                const double* f = hf + 6 * (size_t)b;
                v[0] = v[0] + inv_mass * f[0];
                v[1] = v[1] + inv_mass * f[1];
                v[2] = v[2] + inv_mass * f[2];
                const d3 dw = sym3_apply(inv_inertia, inv_inertia_off, b, mk3(f[3], f[4], f[5]));
                v[3] = v[3] + dw.x;
                v[4] = v[4] + dw.y;
                v[5] = v[5] + dw.z;
            }
            pn = mk3(pn.x + v[0] * h, pn.y + v[1] * h, pn.z + v[2] * h);
            p[0] = pn.x, p[1] = pn.y, p[2] = pn.z;

            const double4 qa = reinterpret_cast<const double4*>(rot)[b];
            const double e0 = qa.x, e1 = qa.y, e2 = qa.z, e3 = qa.w;
            // ChMatrix33::SetFromQuaternion (src/chrono/core/ChMatrix33.h:253-274) times the local angular velocity
            const double e0e0 = e0 * e0, e1e1 = e1 * e1, e2e2 = e2 * e2, e3e3 = e3 * e3;
            const double e0e1 = e0 * e1, e0e2 = e0 * e2, e0e3 = e0 * e3, e1e2 = e1 * e2, e1e3 = e1 * e3, e2e3 = e2 * e3;
            const double wx = v[3], wy = v[4], wz = v[5];
            double ax = ((e0e0 + e1e1) * 2 - 1) * wx + ((e1e2 - e0e3) * 2) * wy + ((e1e3 + e0e2) * 2) * wz;
            double ay = ((e1e2 + e0e3) * 2) * wx + ((e0e0 + e2e2) * 2 - 1) * wy + ((e2e3 - e0e1) * 2) * wz;
            double az = ((e1e3 - e0e2) * 2) * wx + ((e2e3 + e0e1) * 2) * wy + ((e0e0 + e3e3) * 2 - 1) * wz;
            const double len = sqrt(ax * ax + ay * ay + az * az);
            const double angle = len * h;
            // ChVector3::Normalize (src/chrono/core/ChVector3.h:873-883)
            if (len < DBL_MIN) {
                ax = 1, ay = 0, az = 0;
            } else {
                const double s = 1 / len;
                ax *= s, ay *= s, az *= s;
            }
            // SetFromAngleAxis (ChQuaternion.h:1052-1059)
            const double half = angle / 2;
            const double sh = sin(half);
            const double d0 = cos(half), d1 = ax * sh, d2 = ay * sh, d3_ = az * sh;
            // mdeltarot * moldrot, ChQuaternion::Cross (ChQuaternion.h:891-903)
            double4 qn;
            qn.x = d0 * e0 - d1 * e1 - d2 * e2 - d3_ * e3;
            qn.y = d0 * e1 + d1 * e0 - d3_ * e2 + d2 * e3;
            qn.z = d0 * e2 + d2 * e0 + d3_ * e1 - d1 * e3;
            qn.w = d0 * e3 + d3_ * e0 - d2 * e1 + d1 * e2;
            reinterpret_cast<double4*>(rot)[b] = qn;
            if (vlim && vlim[2 * b] > 0) {
                // ChBody::ClampSpeed in VariablesQbSetSpeed (src/chrono/physics/ChBody.cpp:268-286,319-329), which the
                // reference calls AFTER VariablesQbIncrementPosition (ChSystemMulticore.cpp:145-146): the pose above
                // moved with the solver's velocity, the stored one is clamped.  The quaternion rate is 0.5 q (0, w),
                // so w = 2 |rot_dt| = |angular velocity|.
                const double w = sqrt(v[3] * v[3] + v[4] * v[4] + v[5] * v[5]);
                if (w > vlim[2 * b + 1]) {
                    const double s = vlim[2 * b + 1] / w;
                    v[3] *= s, v[4] *= s, v[5] *= s;
                }
                const double sp = sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
                if (sp > vlim[2 * b]) {
                    const double s = vlim[2 * b] / sp;
                    v[0] *= s, v[1] *= s, v[2] *= s;
                }
            }
        }
        if (A.b_deg) {  // whole SMC steps only: NSC keeps its CSR by body for the queries that follow a step
            A.b_deg[b] = 0u, A.heavy[b] = 0;
            if (b == 0)
                A.b_deg[nb] = 0u;
        }
        if (AABB) {
            const int32_t si = A.b_sphere[b];
            if (si >= 0) {
                // k_aabb<false>: centred sphere, Rotate(0, q) + p == p
                const double rad = A.b_rad[b];
                reinterpret_cast<double4*>(A.s_wpos)[si] = make_double4(pn.x, pn.y, pn.z, rad);
                if (fl & 2u) {
                    const double r = rad + A.envelope;
                    bmn = pn - r, bmx = pn + r;
                    counted = true;
                }
            }
        }
    }
    if (AABB)
        bbox_block_merge(cnt, bmn, bmx, counted);  // union of the boxes of colliding shapes, as in k_aabb
}

__global__ void k_reset_bbox(DevCounters* cnt) {
    if (threadIdx.x == 0 && !cnt->overflow) {
        cnt->bbox[0] = cnt->bbox[1] = cnt->bbox[2] = ~0ull;
        cnt->bbox[3] = cnt->bbox[4] = cnt->bbox[5] = 0ull;
    }
}

int enqueue_update_forces(System* S) {
    if (S->nb == 0)
        return CB200_OK;
    const d3 g = mk3(S->set.gravity[0], S->set.gravity[1], S->set.gravity[2]);
    k_update_forces<<<blocks_for(S->nb, 256), 256, 0, S->stream>>>(S->nb, S->set.step_size, g, S->mass.p, S->inertia.p,
                                                                   S->inertia_off.p, S->vel.p, S->has_fapp ? S->fapp.p : nullptr,
                                                                   S->hf.p, S->d_cnt);
    S->launches++;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

bool integrate_fuses_aabb(const System* S) {
    return S->fuse_aabb && S->b_sphere.p && S->s_wpos_alt.p && S->ns > 0 && !S->meta_dirty;
}

// The fused integration merges the next step's sphere boxes into the global box: it starts from an empty one.  Anywhere
// after k_bin_count (the only reader of the previous box) will do, so the SMC path runs it on a side branch.
int enqueue_reset_bbox(System* S, cudaStream_t st) {
    if (S->bbox_reset_done)
        return CB200_OK;
    k_reset_bbox<<<1, 32, 0, st>>>(S->d_cnt);
    S->launches++;
    S->bbox_reset_done = true;
    return CB200_OK;
}

int enqueue_integrate(System* S) {
    if (S->nb == 0)
        return CB200_OK;
    const int grid = blocks_for(S->nb, 256);
    const bool fuse = integrate_fuses_aabb(S);
    const bool clear = S->whole_step && S->set.system_type == 1 && S->b_deg.p && S->heavy.p;
    AabbArgs A = {S->b_sphere.p, S->b_rad.p, S->s_wpos_alt.p, S->set.collision_envelope,
                  clear ? S->b_deg.p : nullptr, S->heavy.p};
#define INTEGRATE(VEL, AABB)                                                                                          \
    k_integrate<VEL, AABB><<<grid, 256, 0, S->stream>>>(S->nb, S->set.step_size, S->bflags.p, S->vel.p, S->pos.p,      \
                                                        S->rot.p, S->mass.p, S->inv_inertia.p, S->inv_inertia_off.p, S->hf.p, A,                     \
                                                        S->has_vlim ? S->vlim.p : nullptr, S->d_cnt)
    if (fuse)
        CB_TRY(enqueue_reset_bbox(S, S->stream));
    S->bbox_reset_done = false;
    CB_TRY(timer_start(S, "k_integrate"));
    if (fuse) {
        if (S->defer_vel_update)
            INTEGRATE(true, true);
        else
            INTEGRATE(false, true);
    } else if (S->defer_vel_update) {
        INTEGRATE(true, false);
    } else {
        INTEGRATE(false, false);
    }
#undef INTEGRATE
    CB_TRY(timer_stop(S, "k_integrate"));
    S->sphere_aabb_valid = fuse;
    S->body_scratch_clean = clear;
    S->launches++;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

}  // namespace cb200
