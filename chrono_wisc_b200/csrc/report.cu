// report.cu -- contact reporting: the arguments of ReportContactCallback::OnReportContact for every contact, batched.
//
// Reference path replaced (MC = src/chrono_multicore):
//   ChContactContainerMulticore::ReportAllContacts  MC/collision/ChContactContainerMulticore.cpp:65-149
//   ChMatrix33::SetFromAxisX / ChVector3::GetDirectionAxesAsX  src/chrono/core/ChMatrix33.h:389-395, ChVector3.h:899-921
//
// The reference walks the contact list on the host and calls the user's callback once per contact.  Here one kernel
// evaluates the contact plane and the force / torque in that plane for all contacts; the host side (system.py
// ReportAllContacts, the C++ shim) downloads the batch and runs the callback loop over it.
#include "common.cuh"
#include <float.h>

namespace cb200 {

constexpr int REPORT_STRIDE = 21;  // pA 3, pB 3, plane 9 (row-major, columns = X Y Z axes), force 3, torque 3

// ChVector3::Normalize (src/chrono/core/ChVector3.h:873-883)
__device__ __forceinline__ bool normalize_ch(d3& v) {
    const double len = sqrt(v.x * v.x + v.y * v.y + v.z * v.z);
    if (len < DBL_MIN) {
        v = mk3(1, 0, 0);
        return false;
    }
    const double s = 1.0 / len;
    v = mk3(v.x * s, v.y * s, v.z * s);
    return true;
}

// system_type 1 (SMC): ct = the 96-byte force records of k_smc_force; 0 (NSC): gamma, rows per contact interleaved
__global__ void __launch_bounds__(128) k_report_contacts(int system_type, int rows, double h,
                                                        const double* __restrict__ c_norm,
                                                        const double* __restrict__ c_pta,
                                                        const double* __restrict__ c_ptb,
                                                        const int32_t* __restrict__ bids,
                                                        const double* __restrict__ pos, const double* __restrict__ rot,
                                                        const double* __restrict__ ct, const double* __restrict__ gamma,
                                                        double* __restrict__ out, const DevCounters* __restrict__ cnt) {
    const uint32_t nc = cnt->n_contacts;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += gridDim.x * blockDim.x) {
        const d3 n = mk3(c_norm[3 * i], c_norm[3 * i + 1], c_norm[3 * i + 2]);
        const d3 pA = mk3(c_pta[3 * i], c_pta[3 * i + 1], c_pta[3 * i + 2]);
        const d3 pB = mk3(c_ptb[3 * i], c_ptb[3 * i + 1], c_ptb[3 * i + 2]);
        // contact plane: X along the normal, Y suggested by VECT_Y (GetDirectionAxesAsX)
        d3 X = n;
        if (!normalize_ch(X))
            X = mk3(1, 0, 0);
        d3 sugg = mk3(0, 1, 0);
        d3 Z = cross3(X, sugg);
        bool ok = normalize_ch(Z);
        for (int idx = 0; !ok && idx < 3; idx++) {
            if (idx == 0)
                sugg.x += 1.0;
            else if (idx == 1)
                sugg.y += 1.0;
            else
                sugg.z += 1.0;
            Z = cross3(X, sugg);
            ok = normalize_ch(Z);
        }
        const d3 Y = cross3(Z, X);
        d3 F = mk3(0), T = mk3(0);
        if (system_type == 0) {
            // :108-131: multipliers over the step, already expressed in the contact frame
            const double* g = gamma + (size_t)rows * i;
            F.x = g[0] / h;
            if (rows >= 3)
                F.y = g[1] / h, F.z = g[2] / h;
            if (rows >= 6)
                T = mk3(g[3] / h, g[4] / h, g[5] / h);
        } else {
            // :132-145: force and torque on body B; the torque is moved from the body origin to the contact point
            const double* r = ct + CT_STRIDE * (size_t)i;
            const d3 f_abs = mk3(r[0], r[1], r[2]);
            const d3 t_loc = mk3(r[8], r[9], r[10]);
            const uint32_t b = (uint32_t)bids[2 * i + 1];
            const double4 qa = reinterpret_cast<const double4*>(rot)[b];
            const q4 q = mkq(qa.x, qa.y, qa.z, qa.w);
            const d3 p = mk3(pos[3 * b], pos[3 * b + 1], pos[3 * b + 2]);
            const d3 f_loc = rotate_t(f_abs, q);
            const d3 pB_loc = rotate_t(pB - p, q);
            const d3 t_loc1 = t_loc - cross3(pB_loc, f_loc);
            const d3 t_abs = rotate(t_loc1, q);
            F = mk3(dot3(X, f_abs), dot3(Y, f_abs), dot3(Z, f_abs));
            T = mk3(dot3(X, t_abs), dot3(Y, t_abs), dot3(Z, t_abs));
        }
        double* o = out + (size_t)REPORT_STRIDE * i;
        o[0] = pA.x, o[1] = pA.y, o[2] = pA.z, o[3] = pB.x, o[4] = pB.y, o[5] = pB.z;
        o[6] = X.x, o[7] = Y.x, o[8] = Z.x, o[9] = X.y, o[10] = Y.y, o[11] = Z.y, o[12] = X.z, o[13] = Y.z, o[14] = Z.z;
        o[15] = F.x, o[16] = F.y, o[17] = F.z, o[18] = T.x, o[19] = T.y, o[20] = T.z;
    }
}

// Fills S->report (REPORT_STRIDE doubles per contact) for the current contact list; the caller has fetched the counters.
int enqueue_report(System* S) {
    const size_t nc = S->h_cnt->n_contacts;
    if (nc == 0)
        return CB200_OK;
    CB_TRY(ensure_contact_geometry(S));
    CB_TRY(S->report.reserve(REPORT_STRIDE * nc));
    const bool smc = S->set.system_type == 1;
    if (smc && S->ct_force.cap < CT_STRIDE * nc)
        return set_error(CB200_ERR_STATE, "no force computation has run on this contact list yet");
    const int rows = smc ? 0 : S->nsc_rows_alloc;
    if (!smc && (rows == 0 || S->gamma.cap < (size_t)rows * nc))
        return set_error(CB200_ERR_STATE, "no NSC solve has run on this contact list yet");
    k_report_contacts<<<grid_for(S, nc, 128), 128, 0, S->stream>>>(
        smc ? 1 : 0, rows, S->set.step_size, S->c_norm.p, S->c_pta.p, S->c_ptb.p, S->c_bids.p, S->pos.p, S->rot.p,
        S->ct_force.p, S->gamma.p, S->report.p, S->d_cnt);
    S->launches++;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

}  // namespace cb200
