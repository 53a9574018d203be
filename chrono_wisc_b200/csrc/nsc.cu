// nsc.cu -- NSC (DVI) solve: contact frames, matrix-free Schur product, cone projection, APGD / APGDREF.
//
// Reference path replaced (MC = src/chrono_multicore):
//   ChIterativeSolverMulticoreNSC::RunTimeStep / ComputeD / ComputeE / ComputeR / SetR / ComputeImpulses
//                                              MC/solver/ChIterativeSolverMulticoreNSC.cpp:22-139,141-369
//   ChConstraintRigidRigid::Setup / Build_D / Build_b / Build_E / Project
//                                              MC/constraints/ChConstraintRigidRigid.cpp:37-156,180-263,289-452
//   Orthogonalize, Cone_generalized_rigid      MC/constraints/ChConstraintUtils.cpp:22-34,69-92
//   ChSchurProduct::operator()                 MC/solver/ChSchurProduct.cpp:58-173
//   ChSolverMulticoreAPGD::Solve, APGDREF      MC/solver/ChSolverMulticoreAPGD.cpp:53-222, ...APGDREF.cpp:32-210
//   ComputeInvMassMatrix (M_invk)              MC/solver/ChIterativeSolverMulticore.cpp:61-147
//   ChSystemMulticoreNSC::CalculateContactForces  MC/physics/ChSystemMulticoreNSC.cpp:92-109
//
// Design (B200): no matrix is ever built.  The reference assembles D_T (36 non-zeros per sliding contact), its
// transpose and M_inv*D, and streams ~0.9 KB per contact through two single-threaded SpMVs per Schur product.  Here
// a contact keeps an 80-byte frame (normal U, the two world-frame arms s_A, s_B, friction) and the product
//     N x = D^T M^-1 D x + E x
// is two passes: contact -> body (per-body ordered sum over the body's incidence list, no FP atomics) and
// body -> contact.  With F = U x_n + V x_u + W x_v the column block of body A is (-F, R_A^T (F x s_A)) and of body B
// (+F, -R_B^T (F x s_B)); the row block is U.(vp_B - vp_A) with vp = v + w_world x s -- algebraically the rows the
// reference stores ([-U, (R^T U) x sbar], ChConstraintRigidRigid.cpp:410-413), evaluated in the world frame so that
// no quaternion is needed per contact.  The reference-form rows are still produced on request (k_nsc_rows) for the
// row-level parity test.
//
// APGD runs entirely on the device: all scalars (L, t, theta, objective, residual, iteration counter, the
// back-tracking / restart / termination decisions) live in one NscCtl block; every dot product is a fixed-shape
// two-stage reduction (per-CTA partials in a fixed grid, then one CTA) so results are run-to-run reproducible.
// The host enqueues a fixed budget of "slots" (one slot = one APGD iteration or one back-tracking retry); kernels
// of slots past termination return at once.  A budget that proves too small raises OVF_NSC like any other capacity.
#include "common.cuh"
#include <stdlib.h>
#include <float.h>
#include <algorithm>
#include <cstring>

namespace cb200 {

enum { NSC_FRAME = 10, NSC_PART = 8, NSC_THREADS = 256, NSC_NGLOB = 16 /* slot of comm_gsum */ };

struct NscCtl {
    double L, t, theta, theta_new, beta, obj2, residual, objective, g_diff, tol_speed, norm_temp, dot_g_temp;
    double tol_objective;   // settings.solver.tolerance_objective, looked at when test_objective is set
    int32_t iter, max_iter, act, done, redo, accepted, improved, restart, ref_variant, slots, total_iter, hist_cap;
    int32_t test_objective; // settings.solver.test_objective (ChSolverMulticoreAPGD.cpp:184-188)
    int32_t ret_iter;       // sum of the values Solve() returned over the phases (measures.solver.total_iteration)
    uint32_t ticket;        // CTAs of k_nsc_eval that have delivered their partial sums (last one runs the scalar part)
};

struct NscFrame {
    d3 U, sA, sB;
    double mu;
};

__device__ __forceinline__ NscFrame load_frame(const double* __restrict__ frame, uint32_t i) {
    const double2* p = reinterpret_cast<const double2*>(frame + (size_t)NSC_FRAME * i);
    const double2 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3), e = __ldg(p + 4);
    NscFrame F;
    F.U = mk3(a.x, a.y, b.x);
    F.sA = mk3(b.y, c.x, c.y);
    F.sB = mk3(d.x, d.y, e.x);
    F.mu = e.y;
    return F;
}

// MC/constraints/ChConstraintUtils.cpp:22-34 (Vx = U, Vy = V, Vz = W)
__device__ __forceinline__ void orthogonalize(const d3& U, d3& V, d3& W) {
    W = cross3(U, mk3(0, 1, 0));
    double mz = length3(W);
    if (mz < 0.0001) {
        W = cross3(U, mk3(1, 0, 0));
        mz = length3(W);
    }
    W = W / mz;
    V = cross3(W, U);
}

// MC/constraints/ChConstraintUtils.cpp:69-92
__device__ __forceinline__ void cone_generalized(double& gn, double& gu, double& gv, double mu) {
    const double ft = sqrt(gu * gu + gv * gv);
    if (ft < (mu * gn))
        return;
    if ((ft) < -(1 / mu) * gn || (fabs(gn) < 10e-15)) {
        gn = 0, gu = 0, gv = 0;
        return;
    }
    gn = (ft * mu + gn) / (mu * mu + 1);
    const double s = (gn * mu) / ft;
    gu *= s;
    gv *= s;
}

// ChConstraintRigidRigid::func_Project_normal / _sliding / _spinning (ChConstraintRigidRigid.cpp:37-156) on the
// ROWS multipliers of one contact; `act` = rows of the running phase (1 NORMAL, 3 SLIDING, 6 SPINNING).
template <int ROWS>
__device__ __forceinline__ void project_contact(double* g, int act, double mu, double coh, double roll, double spin) {
    if (act == 1 || ROWS == 1) {
        g[0] = (g[0] < -coh) ? -coh : g[0];
#pragma unroll
        for (int k = 1; k < ROWS; k++)
            g[k] = 0;
        return;
    }
    if (ROWS >= 3) {
        if (mu == 0) {
            g[0] = (g[0] < -coh) ? -coh : g[0];
            g[1] = g[2] = 0;
        } else {
            double gn = g[0] + coh, gu = g[1], gv = g[2];
            cone_generalized(gn, gu, gv, mu);
            g[0] = gn - coh, g[1] = gu, g[2] = gv;
        }
    }
    if (ROWS == 6) {
        if (act < 6)
            return;  // func_Project_sliding leaves the spinning multipliers alone
        const double f_n = g[0], t_n = g[3], t_u = g[4], t_v = g[5];
        const double t_tang = sqrt(t_v * t_v + t_u * t_u), t_sp = fabs(t_n);
        if (spin) {
            if (t_sp < spin * f_n) {
            } else if ((t_sp < -(1 / spin) * f_n) || (fabs(f_n) < 10e-15)) {
                g[0] = 0;
                g[3] = 0;
            } else {
                const double fp = (t_sp * spin + f_n) / (spin * spin + 1);
                const double tp = fp * spin, s = tp / t_sp;
                g[0] = fp;
                g[3] = s * t_n;
            }
        }
        if (!roll) {
            g[4] = g[5] = 0;
            if (f_n < 0)
                g[0] = 0;
            return;
        }
        if (t_tang < roll * f_n)
            return;
        if ((t_tang < -(1 / roll) * f_n) || (fabs(f_n) < 10e-15)) {
            g[0] = 0;
            g[4] = g[5] = 0;
            return;
        }
        const double fp = (t_tang * roll + f_n) / (roll * roll + 1);
        const double tp = fp * roll, s = tp / t_tang;
        g[0] = fp;
        g[4] = s * t_u;
        g[5] = s * t_v;
    }
}

// per-contact constants that most scenes do not use (cohesion, compliances, rolling/spinning friction):
// aux = (cohesion, E_n, E_t, mu_roll, mu_spin, E_spin, E_roll, -) per contact, only read when `has_aux`
struct NscAux {
    double coh, En, Et, roll, spin, Es, Er;
};
__device__ __forceinline__ NscAux load_aux(const double* __restrict__ aux, uint32_t i, bool has_aux) {
    NscAux a;
    a.coh = a.En = a.Et = a.roll = a.spin = a.Es = a.Er = 0;
    if (has_aux) {
        const double* p = aux + 8 * (size_t)i;
        a.coh = p[0], a.En = p[1], a.Et = p[2], a.roll = p[3], a.spin = p[4], a.Es = p[5], a.Er = p[6];
    }
    return a;
}
template <int ROWS>
__device__ __forceinline__ double aux_E(const NscAux& a, int k) {
    return k == 0 ? a.En : (k < 3 ? a.Et : (k == 3 ? a.Es : a.Er));
}

template <int ROWS>
__device__ __forceinline__ void load_vec(const double* __restrict__ v, uint32_t i, double* o) {
#pragma unroll
    for (int k = 0; k < ROWS; k++)
        o[k] = v[(size_t)ROWS * i + k];
}
template <int ROWS>
__device__ __forceinline__ void store_vec(double* __restrict__ v, uint32_t i, const double* o) {
#pragma unroll
    for (int k = 0; k < ROWS; k++)
        v[(size_t)ROWS * i + k] = o[k];
}

// rows of D^T applied to the body-space vector w (6 per body: v, world angular velocity)
template <int ROWS>
__device__ __forceinline__ void rows_dot(const NscFrame& F, int act, const double* __restrict__ w, uint32_t bA,
                                         uint32_t bB, double* out) {
    const double2* pa = reinterpret_cast<const double2*>(w + 6 * (size_t)bA);
    const double2* pb = reinterpret_cast<const double2*>(w + 6 * (size_t)bB);
    const double2 a0 = pa[0], a1 = pa[1], a2 = pa[2], b0 = pb[0], b1 = pb[1], b2 = pb[2];
    const d3 vA = mk3(a0.x, a0.y, a1.x), wA = mk3(a1.y, a2.x, a2.y);
    const d3 vB = mk3(b0.x, b0.y, b1.x), wB = mk3(b1.y, b2.x, b2.y);
    const d3 d = (vB + cross3(wB, F.sB)) - (vA + cross3(wA, F.sA));
    out[0] = dot3(F.U, d);
    if (ROWS >= 3) {
        out[1] = out[2] = 0;
        if (ROWS == 6)
            out[3] = out[4] = out[5] = 0;
        if (act >= 3) {
            d3 V, W;
            orthogonalize(F.U, V, W);
            out[1] = dot3(V, d);
            out[2] = dot3(W, d);
            if (ROWS == 6 && act == 6) {
                const d3 dw = wB - wA;
                out[3] = dot3(F.U, dw);
                out[4] = dot3(V, dw);
                out[5] = dot3(W, dw);
            }
        }
    }
}

// ... applied to the difference wa - wb of two body-space vectors (subtracted per component first)
template <int ROWS>
__device__ __forceinline__ void rows_dot_diff(const NscFrame& F, int act, const double* __restrict__ wa,
                                              const double* __restrict__ wb, uint32_t bA, uint32_t bB, double* out) {
    const double2* pa = reinterpret_cast<const double2*>(wa + 6 * (size_t)bA);
    const double2* pb = reinterpret_cast<const double2*>(wa + 6 * (size_t)bB);
    const double2* qa = reinterpret_cast<const double2*>(wb + 6 * (size_t)bA);
    const double2* qb = reinterpret_cast<const double2*>(wb + 6 * (size_t)bB);
    const double2 a0 = pa[0], a1 = pa[1], a2 = pa[2], b0 = pb[0], b1 = pb[1], b2 = pb[2];
    const double2 c0 = qa[0], c1 = qa[1], c2 = qa[2], e0 = qb[0], e1 = qb[1], e2 = qb[2];
    const d3 vA = mk3(a0.x - c0.x, a0.y - c0.y, a1.x - c1.x), wA = mk3(a1.y - c1.y, a2.x - c2.x, a2.y - c2.y);
    const d3 vB = mk3(b0.x - e0.x, b0.y - e0.y, b1.x - e1.x), wB = mk3(b1.y - e1.y, b2.x - e2.x, b2.y - e2.y);
    const d3 d = (vB + cross3(wB, F.sB)) - (vA + cross3(wA, F.sA));
    out[0] = dot3(F.U, d);
    if (ROWS >= 3) {
        out[1] = out[2] = 0;
        if (ROWS == 6)
            out[3] = out[4] = out[5] = 0;
        if (act >= 3) {
            d3 V, W;
            orthogonalize(F.U, V, W);
            out[1] = dot3(V, d);
            out[2] = dot3(W, d);
            if (ROWS == 6 && act == 6) {
                const d3 dw = wB - wA;
                out[3] = dot3(F.U, dw);
                out[4] = dot3(V, dw);
                out[5] = dot3(W, dw);
            }
        }
    }
}

// fixed-shape CTA reduction of NP partial sums -> part[blockIdx.x][0..NP)
template <int NP>
__device__ __forceinline__ void block_partials(double* acc, double* __restrict__ part) {
    __shared__ double sm[NSC_THREADS / 32][NP];
#pragma unroll
    for (int q = 0; q < NP; q++)
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
            acc[q] += __shfl_down_sync(0xffffffffu, acc[q], d);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0)
#pragma unroll
        for (int q = 0; q < NP; q++)
            sm[wid][q] = acc[q];
    __syncthreads();
    if (threadIdx.x < NP) {
        double s = sm[0][threadIdx.x];
#pragma unroll
        for (int w = 1; w < NSC_THREADS / 32; w++)
            s += sm[w][threadIdx.x];
        part[(size_t)blockIdx.x * NSC_PART + threadIdx.x] = s;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// setup
// ---------------------------------------------------------------------------------------------------------------
// M_invk = v + M_inv hf (rows of inactive bodies do not exist: M_invk = v), kept in the body frame for the final
// velocity update (mk) and with the angular part rotated to the world frame for the row products (w0).
__global__ void __launch_bounds__(256) k_nsc_body_prep(uint32_t nb, const uint8_t* __restrict__ bflags,
                                                       const double* __restrict__ mass,
                                                       const double* __restrict__ inv_inertia,
                                                       const double* __restrict__ inv_off /* may be null */,
                                                       const double* __restrict__ rot, const double* __restrict__ vel,
                                                       const double* __restrict__ hf, double* __restrict__ mk,
                                                       double* __restrict__ w0, uint32_t* __restrict__ deg,
                                                       const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb)
        return;
    deg[b] = 0;
    if (b == 0)
        deg[nb] = 0;
    const double* v = vel + 6 * (size_t)b;
    const double* f = hf + 6 * (size_t)b;
    double m[6];
    if (bflags[b] & 1u) {
        const double im = 1.0 / mass[b];
        m[0] = v[0] + im * f[0], m[1] = v[1] + im * f[1], m[2] = v[2] + im * f[2];
        const d3 dw = sym3_apply(inv_inertia, inv_off, b, mk3(f[3], f[4], f[5]));
        m[3] = v[3] + dw.x;
        m[4] = v[4] + dw.y;
        m[5] = v[5] + dw.z;
    } else {
#pragma unroll
        for (int k = 0; k < 6; k++)
            m[k] = v[k];
    }
    const double4 q = reinterpret_cast<const double4*>(rot)[b];
    const d3 ww = rotate(mk3(m[3], m[4], m[5]), mkq(q.x, q.y, q.z, q.w));
    double* o = mk + 6 * (size_t)b;
    double* p = w0 + 6 * (size_t)b;
#pragma unroll
    for (int k = 0; k < 6; k++)
        o[k] = m[k];
    p[0] = m[0], p[1] = m[1], p[2] = m[2], p[3] = ww.x, p[4] = ww.y, p[5] = ww.z;
}

struct NscSetupParams {
    double inv_h, inv_hpa, inv_hhpa, alpha, recovery_speed;
    int mode, nm, has_aux;
    const double* mat_ovr;  // per-contact composite records set through an AddContactCallback (api.cu), else null
};

// Per contact: frame, composite material, b (Build_b), E (Build_E), R_full = -b - D^T M_invk, gamma = 0, and the
// body -> contact incidence counts of active bodies.
template <int ROWS>
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_setup(NscSetupParams P, const int32_t* __restrict__ bids,
                                                           const uint32_t* __restrict__ c_mats,
                                                           const double* __restrict__ c_norm,
                                                           const double* __restrict__ c_pta,
                                                           const double* __restrict__ c_ptb,
                                                           const double* __restrict__ c_depth,
                                                           const double* __restrict__ pos,
                                                           const uint8_t* __restrict__ bflags,
                                                           const double* __restrict__ mat,
                                                           const double* __restrict__ w0, double* __restrict__ frame,
                                                           double* __restrict__ aux, double* __restrict__ rfull,
                                                           double* __restrict__ gamma, uint32_t* __restrict__ deg,
                                                           const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nc = cnt->n_contacts;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        const int2 bb = reinterpret_cast<const int2*>(bids)[i];
        const uint32_t bA = (uint32_t)bb.x, bB = (uint32_t)bb.y;
        NscFrame F;
        F.U = mk3(c_norm[3 * i], c_norm[3 * i + 1], c_norm[3 * i + 2]);
        F.sA = mk3(c_pta[3 * i], c_pta[3 * i + 1], c_pta[3 * i + 2]) - mk3(pos[3 * bA], pos[3 * bA + 1], pos[3 * bA + 2]);
        F.sB = mk3(c_ptb[3 * i], c_ptb[3 * i + 1], c_ptb[3 * i + 2]) - mk3(pos[3 * bB], pos[3 * bB + 1], pos[3 * bB + 2]);
        const uint32_t mm = c_mats[i];
        const double* M = P.mat_ovr ? P.mat_ovr + (size_t)i * NSC_MAT_STRIDE
                                    : mat + ((size_t)(mm & 0xffff) * P.nm + (mm >> 16)) * NSC_MAT_STRIDE;
        F.mu = M[0];
        double* fo = frame + (size_t)NSC_FRAME * i;
        fo[0] = F.U.x, fo[1] = F.U.y, fo[2] = F.U.z, fo[3] = F.sA.x, fo[4] = F.sA.y, fo[5] = F.sA.z;
        fo[6] = F.sB.x, fo[7] = F.sB.y, fo[8] = F.sB.z, fo[9] = F.mu;
        if (P.has_aux) {
            double* a = aux + 8 * (size_t)i;
            a[0] = M[3];
            a[1] = P.inv_hhpa * M[4];
            a[2] = P.mode == 1 ? P.inv_hhpa * M[5] : 0.0;  // sliding compliance only in SLIDING mode (Build_E quirk)
            a[3] = M[1], a[4] = M[2];
            a[5] = P.mode == 2 ? P.inv_hhpa * M[7] : 0.0;
            a[6] = P.mode == 2 ? P.inv_hhpa * M[6] : 0.0;
            a[7] = 0;
        }
        const double depth = c_depth[i];
        double bi;
        if (P.alpha > 0)
            bi = P.inv_hpa * depth;
        else if (P.recovery_speed < 0)
            bi = P.inv_h * depth;
        else
            bi = fmax(P.inv_h * depth, -P.recovery_speed);
        double r[ROWS];
        rows_dot<ROWS>(F, ROWS, w0, bA, bB, r);
        r[0] = -bi - r[0];
#pragma unroll
        for (int k = 1; k < ROWS; k++)
            r[k] = -0.0 - r[k];
        store_vec<ROWS>(rfull, i, r);
        double z[ROWS];
#pragma unroll
        for (int k = 0; k < ROWS; k++)
            z[k] = 0;
        store_vec<ROWS>(gamma, i, z);
        if (bflags[bA] & 1u)
            atomicAdd(&deg[bA], 1u);
        if (bflags[bB] & 1u)
            atomicAdd(&deg[bB], 1u);
    }
}

__global__ void __launch_bounds__(256) k_nsc_inc_fill(const int32_t* __restrict__ bids,
                                                      const uint8_t* __restrict__ bflags,
                                                      const uint32_t* __restrict__ off, uint32_t* __restrict__ cursor,
                                                      uint32_t* __restrict__ inc, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nc = cnt->n_contacts;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        const uint32_t b1 = (uint32_t)bids[2 * i], b2 = (uint32_t)bids[2 * i + 1];
        if (bflags[b1] & 1u)
            inc[off[b1] + atomicAdd(&cursor[b1], 1u)] = i << 1;
        if (bflags[b2] & 1u)
            inc[off[b2] + atomicAdd(&cursor[b2], 1u)] = (i << 1) | 1u;
    }
}

// ascending contact order inside every body's list: the summation order of all later passes (run-to-run fixed)
__global__ void __launch_bounds__(256) k_nsc_inc_sort(uint32_t nb, const uint32_t* __restrict__ off,
                                                      uint32_t* __restrict__ inc, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb)
        return;
    const uint32_t s = off[b], e = off[b + 1];
    for (uint32_t i = s + 1; i < e; i++) {
        const uint32_t key = inc[i];
        uint32_t j = i;
        while (j > s && inc[j - 1] > key) {
            inc[j] = inc[j - 1];
            j--;
        }
        inc[j] = key;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// contact -> body pass:  w = M^-1 D x   (one thread per active body, ordered sum over its incidence list)
//   OUT = 0: w for the row products (v, world angular velocity)
//   OUT = 1: ComputeImpulses  v = M_invk + M^-1 D gamma (body frame) and Fc = D gamma / h (CalculateContactForces)
// ---------------------------------------------------------------------------------------------------------------
template <int ROWS, int OUT>
__global__ void __launch_bounds__(256) k_nsc_scatter(uint32_t nb, const NscCtl* __restrict__ ctl, int skip_on_redo,
                                                     const uint8_t* __restrict__ bflags,
                                                     const double* __restrict__ mass,
                                                     const double* __restrict__ inv_inertia,
                                                     const double* __restrict__ inv_off /* may be null */,
                                                     const double* __restrict__ rot, const uint32_t* __restrict__ off,
                                                     const uint32_t* __restrict__ inc,
                                                     const double* __restrict__ frame, const double* __restrict__ x,
                                                     double* __restrict__ w, const double* __restrict__ mk,
                                                     double* __restrict__ vel, double* __restrict__ b_force,
                                                     double inv_h, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    if (OUT == 0 && (ctl->done || (skip_on_redo && ctl->redo)))
        return;
    const int act = OUT == 0 ? ctl->act : ROWS;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb)
        return;
    d3 lin = mk3(0), ang = mk3(0);
    const bool active = bflags[b] & 1u;
    if (active) {
        const uint32_t s = off[b], e = off[b + 1];
        for (uint32_t k = s; k < e; k++) {
            const uint32_t code = inc[k], i = code >> 1, side = code & 1u;
            const NscFrame F = load_frame(frame, i);
            double g[ROWS];
            load_vec<ROWS>(x, i, g);
            d3 f = F.U * g[0];
            d3 tau = mk3(0);
            if (ROWS >= 3 && act >= 3) {
                d3 V, W;
                orthogonalize(F.U, V, W);
                f = f + V * g[1] + W * g[2];
                if (ROWS == 6 && act == 6)
                    tau = F.U * g[3] + V * g[4] + W * g[5];
            }
            if (side == 0) {
                lin = lin - f;
                ang = ang + cross3(f, F.sA) - tau;
            } else {
                lin = lin + f;
                ang = ang - cross3(f, F.sB) + tau;
            }
        }
    }
    const double4 qq = reinterpret_cast<const double4*>(rot)[b];
    const q4 q = mkq(qq.x, qq.y, qq.z, qq.w);
    const d3 ang_l = rotate_t(ang, q);  // body frame
    if (OUT == 0) {
        double* o = w + 6 * (size_t)b;
        if (active) {
            const double im = 1.0 / mass[b];
            const d3 wl = sym3_apply(inv_inertia, inv_off, b, ang_l);
            const d3 ww = rotate(wl, q);
            o[0] = im * lin.x, o[1] = im * lin.y, o[2] = im * lin.z, o[3] = ww.x, o[4] = ww.y, o[5] = ww.z;
        } else {
            o[0] = o[1] = o[2] = o[3] = o[4] = o[5] = 0;
        }
    } else {
        double* fo = b_force + 6 * (size_t)b;
        fo[0] = lin.x * inv_h, fo[1] = lin.y * inv_h, fo[2] = lin.z * inv_h;
        fo[3] = ang_l.x * inv_h, fo[4] = ang_l.y * inv_h, fo[5] = ang_l.z * inv_h;
        const double* m = mk + 6 * (size_t)b;
        double* v = vel + 6 * (size_t)b;
        if (active) {
            const double im = 1.0 / mass[b];
            v[0] = m[0] + im * lin.x, v[1] = m[1] + im * lin.y, v[2] = m[2] + im * lin.z;
            const d3 dw = sym3_apply(inv_inertia, inv_off, b, ang_l);
            v[3] = m[3] + dw.x;
            v[4] = m[4] + dw.y;
            v[5] = m[5] + dw.z;
        } else {
#pragma unroll
            for (int k = 0; k < 6; k++)
                v[k] = m[k];
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// APGD, ChSolverMulticoreAPGD::Solve (MC/solver/ChSolverMulticoreAPGD.cpp:53-222); line numbers below are its.
// Vectors hold ROWS doubles per contact, interleaved (n, u, v[, tn, tu, tv]); rows outside the running phase are
// kept at zero (they are zero in the reference too: R, the restricted product and the projection zero them).
// ---------------------------------------------------------------------------------------------------------------

// temp = gamma - 1 over all `size` rows (:85), ||temp||^2 -> part[.][0]
template <int ROWS>
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_init_temp(const double* __restrict__ gamma,
                                                               double* __restrict__ temp, double* __restrict__ part,
                                                               const uint8_t* __restrict__ own,
                                                               const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nc = cnt->n_contacts;
    double acc[1] = {0};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        double g[ROWS];
        load_vec<ROWS>(gamma, i, g);
        const bool mine = !own || own[i];  // decomposed systems: a contact enters the global sums on one rank only
#pragma unroll
        for (int k = 0; k < ROWS; k++) {
            g[k] = g[k] - 1.0;
            if (mine)
                acc[0] += g[k] * g[k];
        }
        store_vec<ROWS>(temp, i, g);
    }
    block_partials<1>(acc, part);
}

// out = D^T w + E x on the rows of the phase; ||out||^2 -> part[.][1]     (SchurProduct(temp, temp), :114)
template <int ROWS>
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_gather_norm(const NscCtl* __restrict__ ctl,
                                                                 const int32_t* __restrict__ bids,
                                                                 const double* __restrict__ frame,
                                                                 const double* __restrict__ aux, int has_aux,
                                                                 const double* __restrict__ w,
                                                                 const double* __restrict__ x, double* __restrict__ out,
                                                                 double* __restrict__ part,
                                                                 const uint8_t* __restrict__ own,
                                                                 const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nc = cnt->n_contacts;
    const int act = ctl->act;
    double acc[2] = {0, 0};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        const int2 bb = reinterpret_cast<const int2*>(bids)[i];
        const NscFrame F = load_frame(frame, i);
        double o[ROWS];
        rows_dot<ROWS>(F, act, w, (uint32_t)bb.x, (uint32_t)bb.y, o);
        if (has_aux) {
            const NscAux A = load_aux(aux, i, true);
            double xv[ROWS];
            load_vec<ROWS>(x, i, xv);
#pragma unroll
            for (int k = 0; k < ROWS; k++)
                if (k < act)
                    o[k] += aux_E<ROWS>(A, k) * xv[k];
        }
        if (!own || own[i])
#pragma unroll
            for (int k = 0; k < ROWS; k++)
                acc[1] += o[k] * o[k];
        if (out)
            store_vec<ROWS>(out, i, o);
    }
    block_partials<2>(acc, part);
}

// fixed-order sum of the per-CTA partials (one CTA)
template <int NP>
__device__ __forceinline__ void reduce_partials(const double* __restrict__ part, int nblocks, double* s) {
    __shared__ double sm[NSC_THREADS / 32][NP];
    double acc[NP];
#pragma unroll
    for (int q = 0; q < NP; q++)
        acc[q] = 0;
    for (int b = threadIdx.x; b < nblocks; b += NSC_THREADS)
#pragma unroll
        for (int q = 0; q < NP; q++)
            acc[q] += __ldcg(part + (size_t)b * NSC_PART + q);  // possibly written by other CTAs of this very launch
#pragma unroll
    for (int q = 0; q < NP; q++)
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
            acc[q] += __shfl_down_sync(0xffffffffu, acc[q], d);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0)
#pragma unroll
        for (int q = 0; q < NP; q++)
            sm[wid][q] = acc[q];
    __syncthreads();
#pragma unroll
    for (int q = 0; q < NP; q++) {
        double t = sm[0][q];
#pragma unroll
        for (int w = 1; w < NSC_THREADS / 32; w++)
            t += sm[w][q];
        s[q] = t;
    }
}

// start of a phase (one Solve call): scalars of :74-83
// nglob: decomposed systems -- the number of contacts of the whole system (every contact counted on one rank)
__global__ void k_nsc_phase_begin(NscCtl* __restrict__ ctl, int act, int max_iter, int rows, int ref_variant,
                                  double tol_speed, int test_objective, double tol_objective, int first_phase,
                                  int hist_cap, const double* __restrict__ nglob,
                                  const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const double ncg = nglob ? *nglob : (double)cnt->n_contacts;
    const double size = (double)rows * ncg;
    ctl->act = act;
    ctl->max_iter = max_iter;
    ctl->iter = 0;
    ctl->done = (max_iter <= 0 || ncg == 0) ? 1 : 0;
    ctl->redo = 0;
    ctl->accepted = 0;
    ctl->improved = 0;
    ctl->restart = 0;
    ctl->ref_variant = ref_variant;
    ctl->residual = 10e30;
    ctl->objective = 0;  // the reference keeps the last value of measures.solver.objective_value; reported only
    ctl->g_diff = 1.0 / pow(size, 2.0);
    ctl->t = ctl->L = 1.0;
    ctl->theta = ctl->theta_new = 1;
    ctl->beta = 0;
    ctl->obj2 = 0;
    ctl->dot_g_temp = 0;
    ctl->tol_speed = tol_speed;
    ctl->test_objective = test_objective;
    ctl->tol_objective = tol_objective;
    ctl->hist_cap = hist_cap;
    if (first_phase) {
        ctl->total_iter = 0;
        ctl->ret_iter = 0;
        ctl->slots = 0;
    }
}

// L = ||N temp|| / ||temp||, t = 1/L (:102-131)
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_init_L(NscCtl* __restrict__ ctl, const double* __restrict__ part0,
                                                            const double* __restrict__ part1, int nblocks,
                                                            const DevCounters* __restrict__ cnt) {
    if (cnt->overflow || ctl->done)
        return;
    double s0[1], s1[2];
    reduce_partials<1>(part0, nblocks, s0);
    __syncthreads();
    reduce_partials<2>(part1, nblocks, s1);
    if (threadIdx.x == 0) {
        const double norm_temp = sqrt(s0[0]);
        double L;
        if (norm_temp == 0)
            L = 1.0;
        else
            L = sqrt(s1[1]) / norm_temp;
        if (L == 0)
            L = 1;
        ctl->norm_temp = norm_temp;
        ctl->L = L;
        ctl->t = 1.0 / L;
    }
}

// y = gamma, gamma_hat = gamma (:133-137); APGDREF starts gamma_hat at ones (APGDREF.cpp:67)
template <int ROWS>
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_start(const NscCtl* __restrict__ ctl,
                                                           const double* __restrict__ gamma, double* __restrict__ y,
                                                           double* __restrict__ gamma_hat,
                                                           const DevCounters* __restrict__ cnt) {
    if (cnt->overflow || ctl->done)
        return;
    const uint32_t nc = cnt->n_contacts;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        double g[ROWS];
        load_vec<ROWS>(gamma, i, g);
        store_vec<ROWS>(y, i, g);
        store_vec<ROWS>(gamma_hat, i, g);
    }
}

// slot kernel B: temp = N y (rows via w_y), g = temp - r, gamma_new = Project(y - t g)        (:140-143)
// partials: [0] y.temp  [1] y.r.   On a back-tracking retry (:148-154) only the projection is redone.
template <int ROWS>
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_grad(const NscCtl* __restrict__ ctl,
                                                          const int32_t* __restrict__ bids,
                                                          const double* __restrict__ frame,
                                                          const double* __restrict__ aux, int has_aux,
                                                          const double* __restrict__ w, const double* __restrict__ y,
                                                          const double* __restrict__ rfull, double* __restrict__ g,
                                                          double* __restrict__ gn, double* __restrict__ part,
                                                          const uint8_t* __restrict__ own,
                                                          const DevCounters* __restrict__ cnt) {
    if (cnt->overflow || ctl->done)
        return;
    const uint32_t nc = cnt->n_contacts;
    const int act = ctl->act;
    const bool redo = ctl->redo != 0;
    const double t = ctl->t;
    double acc[2] = {0, 0};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        const NscFrame F = load_frame(frame, i);
        const NscAux A = load_aux(aux, i, has_aux != 0);
        double yv[ROWS], gv[ROWS], o[ROWS];
        load_vec<ROWS>(y, i, yv);
        if (!redo) {
            const int2 bb = reinterpret_cast<const int2*>(bids)[i];
            double tmp[ROWS], rv[ROWS];
            rows_dot<ROWS>(F, act, w, (uint32_t)bb.x, (uint32_t)bb.y, tmp);
            load_vec<ROWS>(rfull, i, rv);
            const double mine = (!own || own[i]) ? 1.0 : 0.0;
#pragma unroll
            for (int k = 0; k < ROWS; k++) {
                if (k < act) {
                    if (has_aux)
                        tmp[k] += aux_E<ROWS>(A, k) * yv[k];
                    acc[0] += mine * (yv[k] * tmp[k]);
                    acc[1] += mine * (yv[k] * rv[k]);
                    gv[k] = tmp[k] - rv[k];
                } else {
                    gv[k] = 0;
                }
            }
            store_vec<ROWS>(g, i, gv);
        } else {
            load_vec<ROWS>(g, i, gv);
        }
#pragma unroll
        for (int k = 0; k < ROWS; k++)
            o[k] = yv[k] - t * gv[k];
        project_contact<ROWS>(o, act, F.mu, A.coh, A.roll, A.spin);
        store_vec<ROWS>(gn, i, o);
    }
    if (!redo)
        block_partials<2>(acc, part);
}

__device__ __forceinline__ void nsc_scalar_body(NscCtl* __restrict__ ctl, const double* __restrict__ partG,
                                                const double* __restrict__ partE, int nblocks,
                                                double* __restrict__ hist);

// slot kernel D: N gamma_new and every dot product of :144-147, :159-173
// partials: [0] gn.Ngn  [1] gn.r  [2] (gn-y).g  [3] |gn-y|^2  [4] g.(gn-gamma)  [5] |res|^2  [6] (gn-y).N(gn-y)
// With d = gn - y:  N gn = N y + N d = (g + r) + N d, and N d comes from the body-space difference w_gn - w_y.  The
// last partial is what the back-tracking test needs (k_nsc_scalar).
template <int ROWS>
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_eval(NscCtl* ctl, const double* partG /* null: no scalar part */,
                                                          double* __restrict__ hist,
                                                          const int32_t* __restrict__ bids,
                                                          const double* __restrict__ frame,
                                                          const double* __restrict__ aux, int has_aux,
                                                          const double* __restrict__ w, const double* __restrict__ wy,
                                                          const double* __restrict__ y,
                                                          const double* __restrict__ rfull,
                                                          const double* __restrict__ g, const double* __restrict__ gn,
                                                          const double* __restrict__ gamma, double* __restrict__ part,
                                                          const uint8_t* __restrict__ own,
                                                          const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    if (ctl->done) {
        if (partG && blockIdx.x == 0 && threadIdx.x == 0)
            ctl->accepted = 0;  // what k_nsc_scalar does for a finished phase: the commit kernel has nothing to do
        return;
    }
    const uint32_t nc = cnt->n_contacts;
    const int act = ctl->act;
    const double g_diff = ctl->g_diff, inv_g_diff = 1.0 / g_diff;
    double acc[7] = {0, 0, 0, 0, 0, 0, 0};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        const int2 bb = reinterpret_cast<const int2*>(bids)[i];
        const NscFrame F = load_frame(frame, i);
        const NscAux A = load_aux(aux, i, has_aux != 0);
        double n[ROWS], nd[ROWS], gnv[ROWS], yv[ROWS], gv[ROWS], gam[ROWS], rv[ROWS], tmp[ROWS];
        rows_dot_diff<ROWS>(F, act, w, wy, (uint32_t)bb.x, (uint32_t)bb.y, nd);
        load_vec<ROWS>(gn, i, gnv);
        load_vec<ROWS>(y, i, yv);
        load_vec<ROWS>(g, i, gv);
        load_vec<ROWS>(gamma, i, gam);
        load_vec<ROWS>(rfull, i, rv);
        const double mine = (!own || own[i]) ? 1.0 : 0.0;
#pragma unroll
        for (int k = 0; k < ROWS; k++) {
            const double d = gnv[k] - yv[k];
            if (k >= act) {
                n[k] = 0, rv[k] = 0, nd[k] = 0;
            } else {
                if (has_aux)
                    nd[k] += aux_E<ROWS>(A, k) * d;
                n[k] = (gv[k] + rv[k]) + nd[k];
            }
            acc[0] += mine * (gnv[k] * n[k]);
            acc[1] += mine * (gnv[k] * rv[k]);
            acc[6] += mine * (d * nd[k]);
            acc[2] += mine * (d * gv[k]);
            acc[3] += mine * (d * d);
            acc[4] += mine * (gv[k] * (gnv[k] - gam[k]));
            tmp[k] = gnv[k] - g_diff * (n[k] - rv[k]);
        }
        project_contact<ROWS>(tmp, act, F.mu, A.coh, A.roll, A.spin);
#pragma unroll
        for (int k = 0; k < ROWS; k++) {
            const double e = inv_g_diff * (gnv[k] - tmp[k]);
            acc[5] += mine * (e * e);
        }
    }
    block_partials<7>(acc, part);
    if (!partG)
        return;
    // single-system solve: the CTA that delivers last also runs the scalar part of the slot (k_nsc_scalar's work),
    // which saves a launch of one CTA per slot on the critical path
    __shared__ uint32_t s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
        s_last = atomicAdd(&ctl->ticket, 1u) == gridDim.x - 1 ? 1u : 0u;
    __syncthreads();
    if (!s_last)
        return;
    if (threadIdx.x == 0)
        ctl->ticket = 0;
    __threadfence();
    nsc_scalar_body(ctl, partG, part, (int)gridDim.x, hist);
}

// slot kernel E (one CTA): the scalar part of an iteration -- back-tracking test (:147), theta/beta (:155-156),
// residual and best-iterate bookkeeping (:173-180), termination (:184-192), restart (:194-197), L <- 0.9 L (:199).
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_scalar(NscCtl* __restrict__ ctl, const double* __restrict__ partG,
                                                            const double* __restrict__ partE, int nblocks,
                                                            double* __restrict__ hist,
                                                            const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    if (ctl->done) {
        if (threadIdx.x == 0)
            ctl->accepted = 0;
        return;
    }
    nsc_scalar_body(ctl, partG, partE, nblocks, hist);
}

// (one whole CTA; the caller has checked that the phase is not done)
__device__ __forceinline__ void nsc_scalar_body(NscCtl* __restrict__ ctl, const double* __restrict__ partG,
                                                const double* __restrict__ partE, int nblocks,
                                                double* __restrict__ hist) {
    double sg[2], se[7];
    const bool redo = ctl->redo != 0;
    if (!redo)
        reduce_partials<2>(partG, nblocks, sg);
    __syncthreads();
    reduce_partials<7>(partE, nblocks, se);
    if (threadIdx.x != 0)
        return;
    ctl->slots++;
    if (!redo)
        ctl->obj2 = 0.5 * sg[0] - sg[1];
    double L = ctl->L;
    const double lhs = 0.5 * se[0] - se[1];
    // The reference's test (:147)  obj(gn) > obj(y) + d.g + L/2 |d|^2,  d = gn - y, g = N y - r.  With N symmetric
    // obj(gn) - obj(y) - d.g = 1/2 d.N d exactly, so the test is  d.N d > L |d|^2.  It is evaluated in that form: the
    // two objectives agree to ~13 digits near convergence and their difference is then rounding noise, which the
    // reference only survives because both are summed by the same code in the same order -- not so here, where N y is
    // carried along by linearity (k_nsc_commit) while N gn is summed afresh.
    if (ctl->ref_variant ? (se[6] >= L * se[3]) : (se[6] > L * se[3])) {
        L = 2.0 * L;
        ctl->L = L;
        ctl->t = 1.0 / L;
        ctl->redo = 1;
        ctl->accepted = 0;
        return;
    }
    ctl->redo = 0;
    ctl->accepted = 1;
    const double theta = ctl->theta;
    double theta_new = (-pow(theta, 2.0) + theta * sqrt(pow(theta, 2.0) + 4.0)) / 2.0;
    ctl->beta = theta * (1.0 - theta) / (pow(theta, 2.0) + theta_new);
    ctl->dot_g_temp = se[4];
    const double res = sqrt(se[5]);
    ctl->improved = 0;
    if (res < ctl->residual) {
        ctl->residual = res;
        ctl->improved = 1;
        ctl->objective = lhs;
    }
    if (ctl->ref_variant)
        ctl->objective = lhs;
    const int gi = ctl->total_iter;
    if (hist && gi < ctl->hist_cap) {
        hist[2 * gi] = ctl->residual;
        hist[2 * gi + 1] = ctl->objective;
    }
    ctl->total_iter = gi + 1;
    // :184-192: the objective (of the best iterate so far) against tolerance_objective, or the residual against tol_speed
    if ((ctl->test_objective && !ctl->ref_variant) ? ctl->objective <= ctl->tol_objective
                                                   : ctl->residual < ctl->tol_speed) {  // (APGDREF.cpp:174: residual only)
        ctl->done = 1;  // break: the iteration counter is not advanced (:189-191)
        ctl->restart = 0;
        ctl->ret_iter += ctl->iter;
        return;
    }
    ctl->restart = se[4] > 0 ? 1 : 0;
    if (ctl->restart)
        theta_new = 1.0;
    L = 0.9 * L;
    ctl->L = L;
    ctl->t = 1.0 / L;
    ctl->theta = theta_new;
    ctl->theta_new = theta_new;
    ctl->iter++;
    if (ctl->iter >= ctl->max_iter) {
        ctl->done = 1;
        ctl->ret_iter += ctl->iter;
    }
}

// slot kernel F: commit an accepted iteration: gamma_hat (:175), y (:159,195), gamma (:203)
template <int ROWS>
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_commit(const NscCtl* __restrict__ ctl,
                                                            const double* __restrict__ gn, double* __restrict__ gamma,
                                                            double* __restrict__ y, double* __restrict__ gamma_hat,
                                                            uint32_t nw /* 6 nb */, const double* __restrict__ w_new,
                                                            double* __restrict__ w_y, double* __restrict__ w_g,
                                                            const DevCounters* __restrict__ cnt) {
    if (cnt->overflow || !ctl->accepted)
        return;
    const uint32_t nc = cnt->n_contacts;
    const bool improved = ctl->improved != 0, restart = ctl->restart != 0;
    const bool stop_tol = ctl->done && ctl->iter < ctl->max_iter;  // left by the tolerance break: nothing else moves
    const double beta = ctl->beta;
    // The body-space images follow their vectors by linearity: M^-1 D y' = M^-1 D (gamma' + beta (gamma' - gamma)) is
    // formed from M^-1 D gamma' (what pass C left in w_new) and M^-1 D gamma, instead of a pass over all contacts at
    // the start of the next iteration (the reference evaluates N y afresh, ChSolverMulticoreAPGD.cpp:136).
    if (!stop_tol)
        for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < nw; j += gridDim.x * blockDim.x) {
            const double a = w_new[j];
            w_y[j] = restart ? a : beta * (a - w_g[j]) + a;
            w_g[j] = a;
        }
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        double a[ROWS], o[ROWS];
        load_vec<ROWS>(gn, i, a);
        if (improved)
            store_vec<ROWS>(gamma_hat, i, a);
        if (stop_tol)
            continue;
        if (restart) {
            store_vec<ROWS>(y, i, a);
        } else {
            load_vec<ROWS>(gamma, i, o);
#pragma unroll
            for (int k = 0; k < ROWS; k++)
                o[k] = beta * (a[k] - o[k]) + a[k];
            store_vec<ROWS>(y, i, o);
        }
        store_vec<ROWS>(gamma, i, a);
    }
}

// end of a phase: gamma = gamma_hat (:219); a phase that ran out of slots before terminating raises OVF_NSC
template <int ROWS>
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_phase_end(const NscCtl* __restrict__ ctl,
                                                               const double* __restrict__ gamma_hat,
                                                               double* __restrict__ gamma, int had_iterations,
                                                               DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    if (!ctl->done) {
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            cnt->need_nsc_slots = (uint32_t)(ctl->slots + (ctl->max_iter - ctl->iter) * 2 + 16);
            atomicOr(&cnt->overflow, (uint32_t)OVF_NSC);
        }
        return;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        cnt->nsc_iters = (uint32_t)ctl->ret_iter;
        cnt->nsc_slots = (uint32_t)ctl->slots;
        cnt->nsc_hist_count = (uint32_t)ctl->total_iter;
        cnt->nsc_residual = ctl->residual;
        cnt->nsc_objective = ctl->objective;
    }
    if (!had_iterations)
        return;
    const uint32_t nc = cnt->n_contacts;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        double a[ROWS];
        load_vec<ROWS>(gamma_hat, i, a);
        store_vec<ROWS>(gamma, i, a);
    }
}

// D_T rows in the reference's arithmetic (Build_D, ChConstraintRigidRigid.cpp:377-452): per contact and row
// k < ROWS twelve values, body A block (-dir, (R_A^T dir) x sbar_A) then body B block (dir, -(R_B^T dir) x sbar_B);
// spinning rows (0, -R_A^T dir), (0, R_B^T dir).  Parity/diagnostic path only.
template <int ROWS>
__global__ void __launch_bounds__(256) k_nsc_rows(const int32_t* __restrict__ bids, const double* __restrict__ c_norm,
                                                  const double* __restrict__ c_pta, const double* __restrict__ c_ptb,
                                                  const double* __restrict__ pos, const double* __restrict__ rot,
                                                  double* __restrict__ out, const DevCounters* __restrict__ cnt) {
    const uint32_t nc = cnt->n_contacts;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        const uint32_t bA = (uint32_t)bids[2 * i], bB = (uint32_t)bids[2 * i + 1];
        const double4 ra = reinterpret_cast<const double4*>(rot)[bA], rb = reinterpret_cast<const double4*>(rot)[bB];
        const q4 qa = conj(mkq(ra.x, ra.y, ra.z, ra.w)), qb = conj(mkq(rb.x, rb.y, rb.z, rb.w));
        const d3 sa = rotate(mk3(c_pta[3 * i], c_pta[3 * i + 1], c_pta[3 * i + 2]) -
                                 mk3(pos[3 * bA], pos[3 * bA + 1], pos[3 * bA + 2]), qa);
        const d3 sb = rotate(mk3(c_ptb[3 * i], c_ptb[3 * i + 1], c_ptb[3 * i + 2]) -
                                 mk3(pos[3 * bB], pos[3 * bB + 1], pos[3 * bB + 2]), qb);
        d3 dir[3];
        dir[0] = mk3(c_norm[3 * i], c_norm[3 * i + 1], c_norm[3 * i + 2]);
        orthogonalize(dir[0], dir[1], dir[2]);
        for (int k = 0; k < (ROWS >= 3 ? 3 : 1); k++) {
            const d3 dA = rotate(dir[k], qa), dB = rotate(dir[k], qb);
            const d3 TA = cross3(dA, sa), TB = cross3(dB, sb);
            double* o = out + ((size_t)ROWS * i + k) * 12;
            o[0] = -dir[k].x, o[1] = -dir[k].y, o[2] = -dir[k].z, o[3] = TA.x, o[4] = TA.y, o[5] = TA.z;
            o[6] = dir[k].x, o[7] = dir[k].y, o[8] = dir[k].z, o[9] = -TB.x, o[10] = -TB.y, o[11] = -TB.z;
            if (ROWS == 6) {
                double* s = out + ((size_t)ROWS * i + 3 + k) * 12;
                s[0] = s[1] = s[2] = 0, s[3] = -dA.x, s[4] = -dA.y, s[5] = -dA.z;
                s[6] = s[7] = s[8] = 0, s[9] = dB.x, s[10] = dB.y, s[11] = dB.z;
            }
        }
    }
}

// Project(x) on a caller-provided vector (ChProjectConstraints::operator(), MC/solver/ChSolverMulticore.cpp:25-30)
template <int ROWS>
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_project(int act, const double* __restrict__ frame,
                                                             const double* __restrict__ aux, int has_aux,
                                                             double* __restrict__ x, const DevCounters* __restrict__ cnt) {
    const uint32_t nc = cnt->n_contacts;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        const NscFrame F = load_frame(frame, i);
        const NscAux A = load_aux(aux, i, has_aux != 0);
        double g[ROWS];
        load_vec<ROWS>(x, i, g);
        project_contact<ROWS>(g, act, F.mu, A.coh, A.roll, A.spin);
        store_vec<ROWS>(x, i, g);
    }
}

__global__ void k_nsc_set_act(NscCtl* ctl, int act) {
    ctl->act = act;
    ctl->done = 0;
    ctl->redo = 0;
}

// CalculateContactForces for bodies without rows in M_inv (inactive: container walls): their share of D gamma / h is
// not needed by the solve, so it is accumulated only when somebody asks (the reference's Fc_current protocol,
// MC/physics/ChSystemMulticoreNSC.cpp:92-109).  FP64 atomics: reporting path only.
template <int ROWS>
__global__ void __launch_bounds__(256) k_nsc_fc_inactive(const int32_t* __restrict__ bids,
                                                         const uint8_t* __restrict__ bflags,
                                                         const double* __restrict__ frame,
                                                         const double* __restrict__ gamma, double* __restrict__ acc,
                                                         const DevCounters* __restrict__ cnt) {
    const uint32_t nc = cnt->n_contacts;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        const uint32_t bA = (uint32_t)bids[2 * i], bB = (uint32_t)bids[2 * i + 1];
        const bool inA = !(bflags[bA] & 1u), inB = !(bflags[bB] & 1u);
        if (!inA && !inB)
            continue;
        const NscFrame F = load_frame(frame, i);
        double g[ROWS];
        load_vec<ROWS>(gamma, i, g);
        d3 f = F.U * g[0], tau = mk3(0);
        if (ROWS >= 3) {
            d3 V, W;
            orthogonalize(F.U, V, W);
            f = f + V * g[1] + W * g[2];
            if (ROWS == 6)
                tau = F.U * g[3] + V * g[4] + W * g[5];
        }
        if (inA) {
            const d3 t = cross3(f, F.sA) - tau;
            double* o = acc + 6 * (size_t)bA;
            atomicAdd(o, -f.x), atomicAdd(o + 1, -f.y), atomicAdd(o + 2, -f.z);
            atomicAdd(o + 3, t.x), atomicAdd(o + 4, t.y), atomicAdd(o + 5, t.z);
        }
        if (inB) {
            const d3 t = tau - cross3(f, F.sB);
            double* o = acc + 6 * (size_t)bB;
            atomicAdd(o, f.x), atomicAdd(o + 1, f.y), atomicAdd(o + 2, f.z);
            atomicAdd(o + 3, t.x), atomicAdd(o + 4, t.y), atomicAdd(o + 5, t.z);
        }
    }
}

__global__ void __launch_bounds__(256) k_nsc_fc_finish(uint32_t nb, const uint8_t* __restrict__ bflags,
                                                       const double* __restrict__ rot, const double* __restrict__ acc,
                                                       double inv_h, double* __restrict__ b_force) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb || (bflags[b] & 1u))
        return;
    const double* a = acc + 6 * (size_t)b;
    const double4 qq = reinterpret_cast<const double4*>(rot)[b];
    const d3 tl = rotate_t(mk3(a[3], a[4], a[5]), mkq(qq.x, qq.y, qq.z, qq.w));
    double* o = b_force + 6 * (size_t)b;
    o[0] = a[0] * inv_h, o[1] = a[1] * inv_h, o[2] = a[2] * inv_h;
    o[3] = tl.x * inv_h, o[4] = tl.y * inv_h, o[5] = tl.z * inv_h;
}

// ---------------------------------------------------------------------------------------------------------------
// host side: enqueue only
// ---------------------------------------------------------------------------------------------------------------
static int nsc_rows_of(const System* S) { return S->set.solver_mode == 0 ? 1 : (S->set.solver_mode == 1 ? 3 : 6); }
static int nsc_grid(const System* S) { return S->num_sms * 4; }

static int nsc_reserve(System* S) {
    const int rows = nsc_rows_of(S);
    const size_t c = S->cap_contacts, nb = S->nb;
    if (!S->n_ctl) {
        CB_CUDA(cudaMalloc(&S->n_ctl, sizeof(NscCtl)));
        CB_CUDA(cudaMemsetAsync(S->n_ctl, 0, sizeof(NscCtl), S->stream));
    }
    CB_TRY(S->n_frame.reserve((size_t)NSC_FRAME * c));
    if (S->nsc_has_aux)
        CB_TRY(S->n_aux.reserve(8 * c));
    for (DevBuf<double>* b : {&S->gamma, &S->n_gamma_hat, &S->n_y, &S->n_gn, &S->n_g, &S->n_rfull})
        CB_TRY(b->reserve((size_t)rows * c));
    S->nsc_rows_alloc = rows;
    CB_TRY(S->n_w.reserve(6 * nb + 6));
    CB_TRY(S->n_w0.reserve(6 * nb + 6));
    CB_TRY(S->n_wy.reserve(6 * nb + 6));
    CB_TRY(S->n_wg.reserve(6 * nb + 6));
    CB_TRY(S->n_mk.reserve(6 * nb + 6));
    CB_TRY(S->n_part.reserve((size_t)3 * nsc_grid(S) * NSC_PART));
    const int iters = std::max(0, S->set.max_iteration_normal) + std::max(0, S->set.max_iteration_sliding) +
                      std::max(0, S->set.max_iteration_spinning);
    if (S->nsc_hist_cap < iters + 1) {
        CB_TRY(S->n_hist.reserve(2 * (size_t)(iters + 1)));
        S->nsc_hist_cap = iters + 1;
    }
    return CB200_OK;
}

// ---- decomposed systems: which contacts this rank sums, and sums over all ranks -----------------------------------
// A contact of a slab system exists on every rank that sees both bodies (same state, same arithmetic, so its multipliers
// stay identical everywhere); it enters the dot products on the rank that OWNS (integrates) its dynamic body with the
// smaller GLOBAL id -- a rule every rank evaluates the same way, whatever its local slot order.
__global__ void __launch_bounds__(256) k_nsc_own(const int32_t* __restrict__ bids, const uint8_t* __restrict__ bflags,
                                                 const int64_t* __restrict__ gid, uint32_t nb_static,
                                                 uint8_t* __restrict__ own, double* __restrict__ gsum,
                                                 const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nc = cnt->n_contacts;
    uint32_t mine = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        const int2 bb = reinterpret_cast<const int2*>(bids)[i];
        const uint32_t a = (uint32_t)bb.x, b = (uint32_t)bb.y;
        uint32_t ref = a >= nb_static ? a : b;
        if (a >= nb_static && b >= nb_static && gid[b] < gid[a])
            ref = b;
        const uint8_t o = (ref >= nb_static && (bflags[ref] & 1u)) ? 1 : 0;
        own[i] = o;
        mine += o;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
        mine += __shfl_xor_sync(0xffffffffu, mine, d);
    if ((threadIdx.x & 31) == 0 && mine)
        atomicAdd(&gsum[0], (double)mine);  // whole numbers: exact whatever the order
}

// the per-CTA partials of two reductions -> gsum[0..8), gsum[8..16) (fixed order), ready for the sum over the ranks
__global__ void __launch_bounds__(NSC_THREADS) k_nsc_fold(const NscCtl* __restrict__ ctl, const double* __restrict__ partA,
                                                          const double* __restrict__ partB, int nblocks,
                                                          double* __restrict__ gsum, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow || ctl->done)
        return;
    double a[NSC_PART], b[NSC_PART];
    reduce_partials<NSC_PART>(partA, nblocks, a);
    __syncthreads();
    reduce_partials<NSC_PART>(partB, nblocks, b);
    if (threadIdx.x < NSC_PART) {
        gsum[threadIdx.x] = a[threadIdx.x];
        gsum[NSC_PART + threadIdx.x] = b[threadIdx.x];
    }
}

template <int ROWS>
static int nsc_schur_pass(Group& g, int xsel /* 0: gamma_new -> n_w, 1: y -> n_wy */, int skip_on_redo) {
    for (System* S : g.sys) {
        NscCtl* ctl = (NscCtl*)S->n_ctl;
        const double* x = xsel == 0 ? S->n_gn.p : S->n_y.p;
        k_nsc_scatter<ROWS, 0><<<blocks_for(S->nb, 256), 256, 0, S->stream>>>(
            S->nb, ctl, skip_on_redo, S->bflags.p, S->mass.p, S->inv_inertia.p, S->inv_inertia_off.p, S->rot.p, S->b_deg.p,
            S->b_inc.p,
            S->n_frame.p, x, xsel == 0 ? S->n_w.p : S->n_wy.p, nullptr, nullptr, nullptr, 0.0, S->d_cnt);
        S->launches++;
    }
    // a ghost body's w is only complete on the rank that owns it: owners -> ghost copies
    if (comm_multi(g.sys[0]))
        CB_TRY(comm_halo6(g, xsel == 0 ? 0 : 2));
    return CB200_OK;
}

template <int ROWS>
static int nsc_setup(Group& g) {
    const bool multi = comm_multi(g.sys[0]);
    for (System* S : g.sys) {
        const uint32_t nb = S->nb;
        k_nsc_body_prep<<<blocks_for(nb, 256), 256, 0, S->stream>>>(nb, S->bflags.p, S->mass.p, S->inv_inertia.p,
                                                                    S->inv_inertia_off.p,
                                                                   S->rot.p, S->vel.p, S->hf.p, S->n_mk.p, S->n_w0.p,
                                                                   S->b_deg.p, S->d_cnt);
        CB_CUDA(cudaMemsetAsync(S->b_cursor.p, 0, (size_t)nb * 4, S->stream));
        S->launches++;
    }
    // M_invk = v + M^-1 hf of a ghost body is the owner's (here the ghost is inactive: no M^-1)
    if (multi)
        CB_TRY(comm_halo6(g, 1));
    for (System* S : g.sys) {
        const cb200_settings& st = S->set;
        const uint32_t nb = S->nb;
        const int G = nsc_grid(S);
        NscSetupParams P;
        P.inv_h = 1 / st.step_size;
        P.inv_hpa = 1 / (st.step_size + st.alpha);
        P.inv_hhpa = P.inv_h * P.inv_hpa;
        P.alpha = st.alpha;
        P.recovery_speed = st.contact_recovery_speed;
        P.mode = st.solver_mode;
        P.nm = S->nm_nsc;
        P.has_aux = S->nsc_has_aux;
        P.mat_ovr = S->mat_ovr_valid ? S->mat_ovr.p : nullptr;
        k_nsc_setup<ROWS><<<G, NSC_THREADS, 0, S->stream>>>(P, S->c_bids.p, S->c_mats.p, S->c_norm.p, S->c_pta.p,
                                                            S->c_ptb.p, S->c_depth.p, S->pos.p, S->bflags.p,
                                                            S->mat_nsc.p, S->n_w0.p, S->n_frame.p, S->n_aux.p,
                                                            S->n_rfull.p, S->gamma.p, S->b_deg.p, S->d_cnt);
        S->launches++;
        CB_TRY(scan_u32(S, S->b_deg.p, S->b_deg.p, &S->d_cnt->n_bodies, nb));
        k_nsc_inc_fill<<<G, 256, 0, S->stream>>>(S->c_bids.p, S->bflags.p, S->b_deg.p, S->b_cursor.p, S->b_inc.p,
                                                 S->d_cnt);
        k_nsc_inc_sort<<<blocks_for(nb, 256), 256, 0, S->stream>>>(nb, S->b_deg.p, S->b_inc.p, S->d_cnt);
        S->launches += 2;
        if (multi) {
            CB_CUDA(cudaMemsetAsync(comm_gsum(S) + NSC_NGLOB, 0, sizeof(double), S->stream));
            k_nsc_own<<<G, 256, 0, S->stream>>>(S->c_bids.p, S->bflags.p, S->b_gid.p, S->slab_nb_static, S->c_own.p,
                                                comm_gsum(S) + NSC_NGLOB, S->d_cnt);
            S->launches++;
        }
        CB_CUDA(cudaGetLastError());
    }
    if (multi)
        CB_TRY(comm_allreduce(g, NSC_NGLOB, 1));  // contacts of the whole system
    return CB200_OK;
}

// one Solve call of the reference (one phase): ChSolverMulticoreAPGD::Solve with local_solver_mode = phase
template <int ROWS>
static int nsc_phase(Group& g, int act, int max_iter, bool first_phase) {
    const bool multi = comm_multi(g.sys[0]);
    struct Ptr {
        double *part0, *part1, *part2, *gs;
        const double* ng;
        const uint8_t* own;
    };
    auto ptrs = [&](System* S) {
        const int G = nsc_grid(S);
        Ptr p;
        p.part0 = S->n_part.p, p.part1 = S->n_part.p + (size_t)G * NSC_PART, p.part2 = S->n_part.p + (size_t)2 * G * NSC_PART;
        p.gs = multi ? comm_gsum(S) : nullptr;
        p.ng = multi ? comm_gsum(S) + NSC_NGLOB : nullptr;
        p.own = multi ? S->c_own.p : nullptr;
        return p;
    };
    for (System* S : g.sys) {
        const cb200_settings& st = S->set;
        const Ptr P = ptrs(S);
        k_nsc_phase_begin<<<1, 1, 0, S->stream>>>((NscCtl*)S->n_ctl, act, max_iter, ROWS, st.solver_type == 1 ? 1 : 0,
                                                  st.step_size * st.tolerance, st.test_objective,
                                                  st.tolerance_objective, first_phase ? 1 : 0, S->nsc_hist_cap,
                                                  P.ng, S->d_cnt);
        // initial step length: temp = gamma - 1, L = |N temp| / |temp|
        k_nsc_init_temp<ROWS><<<nsc_grid(S), NSC_THREADS, 0, S->stream>>>(S->gamma.p, S->n_gn.p, P.part0, P.own, S->d_cnt);
        S->launches += 2;
    }
    CB_TRY(nsc_schur_pass<ROWS>(g, 0, 0));
    for (System* S : g.sys) {
        const Ptr P = ptrs(S);
        const int G = nsc_grid(S);
        k_nsc_gather_norm<ROWS><<<G, NSC_THREADS, 0, S->stream>>>((NscCtl*)S->n_ctl, S->c_bids.p, S->n_frame.p, S->n_aux.p,
                                                                  S->nsc_has_aux, S->n_w.p, S->n_gn.p, nullptr, P.part1,
                                                                  P.own, S->d_cnt);
        if (multi)
            k_nsc_fold<<<1, NSC_THREADS, 0, S->stream>>>((NscCtl*)S->n_ctl, P.part0, P.part1, G, P.gs, S->d_cnt);
        S->launches += multi ? 2 : 1;
    }
    if (multi)
        CB_TRY(comm_allreduce(g, 0, 2 * NSC_PART));
    for (System* S : g.sys) {
        const Ptr P = ptrs(S);
        NscCtl* ctl = (NscCtl*)S->n_ctl;
        if (multi)
            k_nsc_init_L<<<1, NSC_THREADS, 0, S->stream>>>(ctl, P.gs, P.gs + NSC_PART, 1, S->d_cnt);
        else
            k_nsc_init_L<<<1, NSC_THREADS, 0, S->stream>>>(ctl, P.part0, P.part1, nsc_grid(S), S->d_cnt);
        k_nsc_start<ROWS><<<nsc_grid(S), NSC_THREADS, 0, S->stream>>>(ctl, S->gamma.p, S->n_y.p, S->n_gamma_hat.p, S->d_cnt);
        S->launches += 2;
    }
    // A, once per phase: w_y = w_g = M^-1 D y for y = gamma_0; afterwards k_nsc_commit carries both along
    CB_TRY(nsc_schur_pass<ROWS>(g, 1, 0));
    for (System* S : g.sys)
        CB_CUDA(cudaMemcpyAsync(S->n_wg.p, S->n_wy.p, 6 * (size_t)S->nb * sizeof(double), cudaMemcpyDeviceToDevice,
                                S->stream));
    const int slots = max_iter + max_iter / 2 + 16 + g.sys[0]->nsc_extra_slots;
    static const bool direct_wy = getenv("CB200_NSC_DIRECT_WY") != nullptr;  // debugging aid: N y by a pass of its own
    for (int s = 0; s < slots; s++) {
        if (direct_wy)
            CB_TRY(nsc_schur_pass<ROWS>(g, 1, 1));
        for (System* S : g.sys) {
            const Ptr P = ptrs(S);
            k_nsc_grad<ROWS><<<nsc_grid(S), NSC_THREADS, 0, S->stream>>>((NscCtl*)S->n_ctl, S->c_bids.p, S->n_frame.p,
                                                                         S->n_aux.p, S->nsc_has_aux, S->n_wy.p, S->n_y.p,
                                                                         S->n_rfull.p, S->n_g.p, S->n_gn.p, P.part1, P.own,
                                                                         S->d_cnt);  // B
            S->launches++;
        }
        CB_TRY(nsc_schur_pass<ROWS>(g, 0, 0));  // C: w = M^-1 D gamma_new
        for (System* S : g.sys) {
            const Ptr P = ptrs(S);
            const int G = nsc_grid(S);
            k_nsc_eval<ROWS><<<G, NSC_THREADS, 0, S->stream>>>((NscCtl*)S->n_ctl, multi ? nullptr : P.part1, S->n_hist.p,
                                                               S->c_bids.p, S->n_frame.p, S->n_aux.p,
                                                               S->nsc_has_aux, S->n_w.p, S->n_wy.p, S->n_y.p, S->n_rfull.p,
                                                               S->n_g.p, S->n_gn.p, S->gamma.p, P.part2, P.own, S->d_cnt);  // D
            if (multi)
                k_nsc_fold<<<1, NSC_THREADS, 0, S->stream>>>((NscCtl*)S->n_ctl, P.part1, P.part2, G, P.gs, S->d_cnt);
            S->launches += multi ? 2 : 1;
        }
        // the one synchronisation point of a slot: {y.Ny, y.r} and {g'.Ng', g'.r, d.g, |d|^2, g.dgamma, |res|^2} over
        // all ranks (ChSolverMulticoreAPGD.cpp:145,147,160,173 in one message)
        if (multi)
            CB_TRY(comm_allreduce(g, 0, 2 * NSC_PART));
        for (System* S : g.sys) {
            const Ptr P = ptrs(S);
            NscCtl* ctl = (NscCtl*)S->n_ctl;
            if (multi)  // (a single system ran this part in the last CTA of k_nsc_eval)
                k_nsc_scalar<<<1, NSC_THREADS, 0, S->stream>>>(ctl, P.gs, P.gs + NSC_PART, 1, S->n_hist.p, S->d_cnt);  // E
            k_nsc_commit<ROWS><<<nsc_grid(S), NSC_THREADS, 0, S->stream>>>(ctl, S->n_gn.p, S->gamma.p, S->n_y.p,
                                                                           S->n_gamma_hat.p, 6 * S->nb, S->n_w.p,
                                                                           S->n_wy.p, S->n_wg.p, S->d_cnt);  // F
            S->launches += multi ? 2 : 1;
        }
    }
    for (System* S : g.sys) {
        k_nsc_phase_end<ROWS><<<nsc_grid(S), NSC_THREADS, 0, S->stream>>>((NscCtl*)S->n_ctl, S->n_gamma_hat.p, S->gamma.p, 1,
                                                                          S->d_cnt);
        S->launches++;
        CB_CUDA(cudaGetLastError());
    }
    return CB200_OK;
}

template <int ROWS>
static int nsc_run(Group& g) {
    const cb200_settings& st = g.sys[0]->set;
    CB_TRY(nsc_setup<ROWS>(g));
    const int iters[3] = {st.max_iteration_normal, st.max_iteration_sliding, st.max_iteration_spinning};
    const int acts[3] = {1, 3, 6};
    bool first = true;
    for (int ph = 0; ph <= st.solver_mode; ph++) {
        if (iters[ph] <= 0)
            continue;
        CB_TRY((nsc_phase<ROWS>(g, acts[ph], iters[ph], first)));
        first = false;
    }
    for (System* S : g.sys) {
        if (first) {  // no phase ran: counters of an empty solve
            k_nsc_phase_begin<<<1, 1, 0, S->stream>>>((NscCtl*)S->n_ctl, ROWS, 0, ROWS, 0, 0.0, 0, 0.0, 1, S->nsc_hist_cap, nullptr,
                                                      S->d_cnt);
            k_nsc_phase_end<ROWS><<<1, NSC_THREADS, 0, S->stream>>>((NscCtl*)S->n_ctl, S->n_gamma_hat.p, S->gamma.p, 0,
                                                                    S->d_cnt);
            S->launches += 2;
        }
        // ComputeImpulses + CalculateContactForces
        k_nsc_scatter<ROWS, 1><<<blocks_for(S->nb, 256), 256, 0, S->stream>>>(
            S->nb, (NscCtl*)S->n_ctl, 0, S->bflags.p, S->mass.p, S->inv_inertia.p, S->inv_inertia_off.p, S->rot.p, S->b_deg.p,
            S->b_inc.p,
            S->n_frame.p, S->gamma.p, nullptr, S->n_mk.p, S->vel.p, S->b_force.p, 1 / st.step_size, S->d_cnt);
        S->launches++;
        CB_CUDA(cudaGetLastError());
    }
    return CB200_OK;
}

int nsc_contact_forces_inactive(System* S) {
    if (S->nsc_fc_current || S->nb == 0 || !S->n_ctl)
        return CB200_OK;
    CB_CUDA(cudaMemsetAsync(S->n_w.p, 0, 6 * (size_t)S->nb * 8, S->stream));
    const int G = nsc_grid(S);
    switch (S->nsc_rows_alloc) {
        case 1: k_nsc_fc_inactive<1><<<G, 256, 0, S->stream>>>(S->c_bids.p, S->bflags.p, S->n_frame.p, S->gamma.p, S->n_w.p, S->d_cnt); break;
        case 3: k_nsc_fc_inactive<3><<<G, 256, 0, S->stream>>>(S->c_bids.p, S->bflags.p, S->n_frame.p, S->gamma.p, S->n_w.p, S->d_cnt); break;
        default: k_nsc_fc_inactive<6><<<G, 256, 0, S->stream>>>(S->c_bids.p, S->bflags.p, S->n_frame.p, S->gamma.p, S->n_w.p, S->d_cnt); break;
    }
    k_nsc_fc_finish<<<blocks_for(S->nb, 256), 256, 0, S->stream>>>(S->nb, S->bflags.p, S->rot.p, S->n_w.p,
                                                                  1 / S->set.step_size, S->b_force.p);
    S->launches += 2;
    CB_CUDA(cudaGetLastError());
    S->nsc_fc_current = true;
    return CB200_OK;
}

int enqueue_nsc_group(Group& g) {
    for (System* S : g.sys) {
        if (S->nb == 0)
            return CB200_OK;
        const cb200_settings& st = S->set;
        if (S->ns > 0 && S->nm_nsc == 0)
            return set_error(CB200_ERR_STATE, "cb200_set_materials_nsc was not called");
        if (st.solver_type != 0 && st.solver_type != 1)
            return set_error(CB200_ERR_UNSUPPORTED, "only the APGD and APGDREF solvers are implemented");
        if (st.solver_mode < 0 || st.solver_mode > 2)
            return set_error(CB200_ERR_ARG, "solver_mode must be NORMAL, SLIDING or SPINNING");
        CB_TRY(nsc_reserve(S));
        if (comm_multi(S))
            CB_TRY(S->c_own.reserve(S->cap_contacts));
        S->nsc_fc_current = false;
    }
    switch (nsc_rows_of(g.sys[0])) {
        case 1: return nsc_run<1>(g);
        case 3: return nsc_run<3>(g);
        default: return nsc_run<6>(g);
    }
}

int enqueue_nsc(System* S) {
    Group g;
    g.sys.push_back(S);
    return enqueue_nsc_group(g);
}

}  // namespace cb200

// ---------------------------------------------------------------------------------------------------------------
// C ABI: NSC read-backs and the functor seams of the inner solver (SURVEY 8b "inner NSC solver")
// ---------------------------------------------------------------------------------------------------------------
using namespace cb200;

namespace {
System* SYSN(cb200_system* s) { return reinterpret_cast<System*>(s); }

// index of (contact i, row k) in the reference layout [n | u v | tn tu tv], MC/ChDataManager.h:307-317
inline size_t ref_index(size_t nc, size_t i, int k) {
    if (k == 0)
        return i;
    if (k < 3)
        return nc + 2 * i + (size_t)(k - 1);
    return 3 * nc + 3 * i + (size_t)(k - 3);
}

int download_rows_vector(System* S, const double* dev, double* out_ref_layout, int capacity) {
    CB_TRY(fetch_counters(S));
    const size_t nc = S->h_cnt->n_contacts;
    const int rows = S->nsc_rows_alloc;
    if (nc == 0 || !out_ref_layout)
        return CB200_OK;
    if (!dev || rows == 0)
        return set_error(CB200_ERR_STATE, "no NSC solve has run yet");
    if (capacity >= 0 && (size_t)capacity < rows * nc)
        return set_error(CB200_ERR_ARG, "output capacity smaller than rows * num_rigid_contacts");
    std::vector<double> h(rows * nc);
    CB_CUDA(cudaMemcpyAsync(h.data(), dev, h.size() * 8, cudaMemcpyDeviceToHost, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    for (size_t i = 0; i < nc; i++)
        for (int k = 0; k < rows; k++)
            out_ref_layout[ref_index(nc, i, k)] = h[rows * i + k];
    return CB200_OK;
}

int upload_rows_vector(System* S, const double* in_ref_layout, DevBuf<double>& buf) {
    const size_t nc = S->h_cnt->n_contacts;
    const int rows = S->nsc_rows_alloc;
    std::vector<double> h(rows * nc);
    for (size_t i = 0; i < nc; i++)
        for (int k = 0; k < rows; k++)
            h[rows * i + k] = in_ref_layout[ref_index(nc, i, k)];
    CB_TRY(buf.reserve(h.size()));
    CB_CUDA(cudaMemcpyAsync(buf.p, h.data(), h.size() * 8, cudaMemcpyHostToDevice, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}
}  // namespace

extern "C" int cb200_download_gamma(cb200_system* sys, double* gamma, int capacity) {
    System* S = SYSN(sys);
    if (!S)
        return set_error(CB200_ERR_ARG, "null system");
    CB_CUDA(cudaSetDevice(S->device));
    return download_rows_vector(S, S->gamma.p, gamma, capacity);
}

extern "C" int cb200_download_nsc_rhs(cb200_system* sys, double* r_full) {
    System* S = SYSN(sys);
    if (!S)
        return set_error(CB200_ERR_ARG, "null system");
    CB_CUDA(cudaSetDevice(S->device));
    return download_rows_vector(S, S->n_rfull.p, r_full, -1);
}

extern "C" int cb200_download_nsc_history(cb200_system* sys, double* residual, double* objective, int capacity,
                                          int* count) {
    System* S = SYSN(sys);
    if (!S || !count)
        return set_error(CB200_ERR_ARG, "null argument");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(fetch_counters(S));
    const int n = std::min<int>((int)S->h_cnt->nsc_hist_count, S->nsc_hist_cap);
    *count = n;
    if (n == 0 || capacity <= 0)
        return CB200_OK;
    std::vector<double> h(2 * (size_t)n);
    CB_CUDA(cudaMemcpyAsync(h.data(), S->n_hist.p, h.size() * 8, cudaMemcpyDeviceToHost, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    for (int i = 0; i < n && i < capacity; i++) {
        if (residual)
            residual[i] = h[2 * i];
        if (objective)
            objective[i] = h[2 * i + 1];
    }
    return CB200_OK;
}

extern "C" int cb200_download_nsc_jacobian(cb200_system* sys, double* rows12, int32_t* body_ids) {
    System* S = SYSN(sys);
    if (!S || !rows12)
        return set_error(CB200_ERR_ARG, "null argument");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(fetch_counters(S));
    const size_t nc = S->h_cnt->n_contacts;
    if (nc == 0)
        return CB200_OK;
    const int rows = nsc_rows_of(S);
    CB_TRY(S->n_rows.reserve(12 * (size_t)rows * nc));
    const int G = nsc_grid(S);
    switch (rows) {
        case 1: k_nsc_rows<1><<<G, 256, 0, S->stream>>>(S->c_bids.p, S->c_norm.p, S->c_pta.p, S->c_ptb.p, S->pos.p, S->rot.p, S->n_rows.p, S->d_cnt); break;
        case 3: k_nsc_rows<3><<<G, 256, 0, S->stream>>>(S->c_bids.p, S->c_norm.p, S->c_pta.p, S->c_ptb.p, S->pos.p, S->rot.p, S->n_rows.p, S->d_cnt); break;
        default: k_nsc_rows<6><<<G, 256, 0, S->stream>>>(S->c_bids.p, S->c_norm.p, S->c_pta.p, S->c_ptb.p, S->pos.p, S->rot.p, S->n_rows.p, S->d_cnt); break;
    }
    S->launches++;
    std::vector<double> h(12 * (size_t)rows * nc);
    std::vector<int32_t> bids(2 * nc);
    CB_CUDA(cudaMemcpyAsync(h.data(), S->n_rows.p, h.size() * 8, cudaMemcpyDeviceToHost, S->stream));
    CB_CUDA(cudaMemcpyAsync(bids.data(), S->c_bids.p, bids.size() * 4, cudaMemcpyDeviceToHost, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    for (size_t i = 0; i < nc; i++)
        for (int k = 0; k < rows; k++) {
            const size_t r = ref_index(nc, i, k);
            memcpy(rows12 + 12 * r, h.data() + 12 * (rows * i + k), 12 * sizeof(double));
            if (body_ids)
                body_ids[2 * r] = bids[2 * i], body_ids[2 * r + 1] = bids[2 * i + 1];
        }
    return CB200_OK;
}

template <int ROWS>
static int schur_host(System* S, int act, const double* x, double* out) {
    NscCtl* ctl = (NscCtl*)S->n_ctl;
    CB_TRY(upload_rows_vector(S, x, S->n_gn));
    k_nsc_set_act<<<1, 1, 0, S->stream>>>(ctl, act);
    Group g;
    g.sys.push_back(S);
    CB_TRY(nsc_schur_pass<ROWS>(g, 0, 0));
    k_nsc_gather_norm<ROWS><<<nsc_grid(S), NSC_THREADS, 0, S->stream>>>(ctl, S->c_bids.p, S->n_frame.p, S->n_aux.p,
                                                                        S->nsc_has_aux, S->n_w.p, S->n_gn.p, S->n_g.p,
                                                                        S->n_part.p, nullptr, S->d_cnt);
    S->launches += 2;
    CB_CUDA(cudaGetLastError());
    return download_rows_vector(S, S->n_g.p, out, -1);
}

// ChSchurProduct::operator()(x, out) with local_solver_mode = local_mode (0 NORMAL, 1 SLIDING, 2 SPINNING), on the
// system assembled by the last cb200_run_solver_nsc / step.  x, out: host vectors in the reference layout.
extern "C" int cb200_nsc_schur_product(cb200_system* sys, int local_mode, const double* x, double* out) {
    System* S = SYSN(sys);
    if (!S || !x || !out)
        return set_error(CB200_ERR_ARG, "null argument");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(fetch_counters(S));
    if (S->h_cnt->n_contacts == 0)
        return CB200_OK;
    if (!S->n_ctl || S->nsc_rows_alloc != nsc_rows_of(S))
        return set_error(CB200_ERR_STATE, "no NSC solve has run yet");
    if (local_mode < 0 || local_mode > S->set.solver_mode)
        return set_error(CB200_ERR_ARG, "local_mode must not exceed solver_mode");
    const int act = local_mode == 0 ? 1 : (local_mode == 1 ? 3 : 6);
    switch (S->nsc_rows_alloc) {
        case 1: return schur_host<1>(S, act, x, out);
        case 3: return schur_host<3>(S, act, x, out);
        default: return schur_host<6>(S, act, x, out);
    }
}

// ChProjectConstraints::operator()(gamma) with local_solver_mode = local_mode; in place on a host vector
extern "C" int cb200_nsc_project(cb200_system* sys, int local_mode, double* gamma) {
    System* S = SYSN(sys);
    if (!S || !gamma)
        return set_error(CB200_ERR_ARG, "null argument");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(fetch_counters(S));
    if (S->h_cnt->n_contacts == 0)
        return CB200_OK;
    if (!S->n_ctl || S->nsc_rows_alloc != nsc_rows_of(S))
        return set_error(CB200_ERR_STATE, "no NSC solve has run yet");
    if (local_mode < 0 || local_mode > S->set.solver_mode)
        return set_error(CB200_ERR_ARG, "local_mode must not exceed solver_mode");
    const int act = local_mode == 0 ? 1 : (local_mode == 1 ? 3 : 6);
    CB_TRY(upload_rows_vector(S, gamma, S->n_gn));
    const int G = nsc_grid(S);
    switch (S->nsc_rows_alloc) {
        case 1: k_nsc_project<1><<<G, NSC_THREADS, 0, S->stream>>>(act, S->n_frame.p, S->n_aux.p, S->nsc_has_aux, S->n_gn.p, S->d_cnt); break;
        case 3: k_nsc_project<3><<<G, NSC_THREADS, 0, S->stream>>>(act, S->n_frame.p, S->n_aux.p, S->nsc_has_aux, S->n_gn.p, S->d_cnt); break;
        default: k_nsc_project<6><<<G, NSC_THREADS, 0, S->stream>>>(act, S->n_frame.p, S->n_aux.p, S->nsc_has_aux, S->n_gn.p, S->d_cnt); break;
    }
    S->launches++;
    CB_CUDA(cudaGetLastError());
    return download_rows_vector(S, S->n_gn.p, gamma, -1);
}
