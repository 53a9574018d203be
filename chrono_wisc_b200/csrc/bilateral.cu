// bilateral.cu -- bilateral (joint / motor) constraint rows between bodies, solved at velocity level with MINRES.
//
// Reference path replaced (MC = src/chrono_multicore):
//   ChConstraintBilateral::Build_b / Build_D (BODY_BODY rows)   MC/constraints/ChConstraintBilateral.cpp:34-78
//   ChIterativeSolverMulticoreSMC::ComputeR, RunTimeStep        MC/solver/ChIterativeSolverMulticoreSMC.cpp:772-852
//   ChSchurProductBilateral (N = D_b^T M^-1 D_b)                MC/solver/ChSchurProduct.cpp:175-185
//   ChSolverMulticoreMinRes::Solve                              MC/solver/ChSolverMulticoreMINRES.cpp:19-112
//   PerformStabilization, ComputeImpulses                       MC/solver/ChIterativeSolverMulticore.cpp:230-252,
//                                                               ...SMC.cpp:854-868
//
// The rows themselves (Cq_a, Cq_b, right-hand side) are produced by Chrono's link classes on the host every step --
// that is not hot-path work -- and handed over with cb200_set_bilaterals.  The systems are tiny next to the contact
// problem (6 rows for the mixer's motor, ~100 for a vehicle), so one CTA runs the Krylov loop; what matters is that
// the solve sits on the device, between the force accumulation and the integration, without a host round trip.
// Shaft rows (SHAFT_SHAFT, SHAFT_BODY, ...) are not implemented.
#include "common.cuh"
#include <math.h>

namespace cb200 {

constexpr int BIL_THREADS = 256;

__device__ __forceinline__ double bil_minv(const double* __restrict__ mass, const double* __restrict__ inv_inertia,
                                           uint32_t b, int k) {
    return k < 3 ? 1.0 / mass[b] : inv_inertia[3 * b + (k - 3)];
}

// R_b = -b - D_b^T M_invk, M_invk = v + M^-1 hf (rows of inactive bodies do not exist: their M_invk is v)
__global__ void __launch_bounds__(BIL_THREADS) k_bil_setup(uint32_t n, const int32_t* __restrict__ ids,
                                                           const double* __restrict__ cq, const double* __restrict__ rhs,
                                                           const uint8_t* __restrict__ bflags,
                                                           const double* __restrict__ mass,
                                                           const double* __restrict__ inv_inertia,
                                                           const double* __restrict__ vel, const double* __restrict__ hf,
                                                           double* __restrict__ R, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n)
        return;
    double r = -rhs[i];
    for (int s = 0; s < 2; s++) {
        const uint32_t b = (uint32_t)ids[2 * i + s];
        const bool act = bflags[b] & 1u;
        for (int k = 0; k < 6; k++) {
            double mik = vel[6 * (size_t)b + k];
            if (act)
                mik = mik + bil_minv(mass, inv_inertia, b, k) * hf[6 * (size_t)b + k];
            r -= cq[12 * (size_t)i + 6 * s + k] * mik;
        }
    }
    R[i] = r;
}

// N_ij = sum over the bodies rows i and j share of D_i M^-1 D_j (active bodies only)
__global__ void __launch_bounds__(BIL_THREADS) k_bil_matrix(uint32_t n, const int32_t* __restrict__ ids,
                                                            const double* __restrict__ cq,
                                                            const uint8_t* __restrict__ bflags,
                                                            const double* __restrict__ mass,
                                                            const double* __restrict__ inv_inertia,
                                                            double* __restrict__ N, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < (size_t)n * n; t += (size_t)gridDim.x * blockDim.x) {
        const uint32_t i = (uint32_t)(t / n), j = (uint32_t)(t % n);
        double a = 0;
        for (int s = 0; s < 2; s++) {
            const uint32_t b = (uint32_t)ids[2 * i + s];
            if (!(bflags[b] & 1u))
                continue;
            for (int u = 0; u < 2; u++) {
                if ((uint32_t)ids[2 * j + u] != b)
                    continue;
                for (int k = 0; k < 6; k++)
                    a += cq[12 * (size_t)i + 6 * s + k] * bil_minv(mass, inv_inertia, b, k) * cq[12 * (size_t)j + 6 * u + k];
            }
        }
        N[t] = a;
    }
}

// fixed-shape block sum, the result broadcast to every thread
__device__ __forceinline__ double bil_block_sum(double v, double* sm) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
        v += __shfl_xor_sync(0xffffffffu, v, d);
    __syncthreads();
    if ((threadIdx.x & 31) == 0)
        sm[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0;
#pragma unroll
    for (int w = 0; w < BIL_THREADS / 32; w++)
        t += sm[w];
    return t;
}

// ChSolverMulticoreMinRes::Solve with x0 = 0 (RunTimeStep zeroes gamma first); vec = 7 work vectors of n doubles.
// info: [0] iterations run, [1] residual
__global__ void __launch_bounds__(BIL_THREADS) k_bil_minres(uint32_t n, int max_iter, double tol,
                                                            const double* __restrict__ N, const double* __restrict__ mb,
                                                            double* __restrict__ x, double* __restrict__ vec,
                                                            double* __restrict__ info, const DevCounters* __restrict__ cnt) {
    __shared__ double sm[BIL_THREADS / 32];
    if (cnt->overflow)
        return;
    double *v = vec, *v_old = vec + n, *v_hat = vec + 2 * (size_t)n, *w = vec + 3 * (size_t)n, *w_old = vec + 4 * (size_t)n,
           *w_oold = vec + 5 * (size_t)n, *Av = vec + 6 * (size_t)n;
    const uint32_t tid = threadIdx.x;
    double part = 0;
    for (uint32_t i = tid; i < n; i += BIL_THREADS) {
        x[i] = 0;
        v_hat[i] = mb[i];  // mb - N x
        v[i] = 0, w[i] = 0, w_old[i] = 0, w_oold[i] = 0, v_old[i] = 0;
        part += mb[i] * mb[i];
    }
    double beta = sqrt(bil_block_sum(part, sm));
    double c = 1, eta = beta, norm_rMR = beta, c_old = 1, s_old = 0, s = 0;
    const double norm_r0 = beta;
    int it = 0;
    double residual = beta == 0 ? 0.0 : 1.0;
    if (!(beta == 0 || norm_rMR / norm_r0 < tol)) {
        for (it = 0; it < max_iter; it++) {
            // Lanczos
            const double ib = 1.0 / beta;
            for (uint32_t i = tid; i < n; i += BIL_THREADS) {
                v_old[i] = v[i];
                v[i] = ib * v_hat[i];
            }
            __syncthreads();
            part = 0;
            for (uint32_t i = tid; i < n; i += BIL_THREADS) {
                double a = 0;
                for (uint32_t j = 0; j < n; j++)
                    a += N[(size_t)i * n + j] * v[j];
                Av[i] = a;
                part += v[i] * a;
            }
            const double alpha = bil_block_sum(part, sm);
            part = 0;
            for (uint32_t i = tid; i < n; i += BIL_THREADS) {
                double h = Av[i] - alpha * v[i];
                h -= beta * v_old[i];
                v_hat[i] = h;
                part += h * h;
            }
            const double beta_old = beta;
            beta = sqrt(bil_block_sum(part, sm));
            // QR factorisation, Givens rotation
            const double c_oold = c_old;
            c_old = c;
            const double s_oold = s_old;
            s_old = s;
            const double r1_hat = c_old * alpha - c_oold * s_old * beta_old;
            const double r1 = 1 / sqrt(r1_hat * r1_hat + beta * beta);
            const double r2 = s_old * alpha + c_oold * c_old * beta_old;
            const double r3 = s_oold * beta_old;
            c = r1_hat * r1;
            s = beta * r1;
            for (uint32_t i = tid; i < n; i += BIL_THREADS) {
                const double woo = w_old[i], wo = w[i];
                w_oold[i] = woo;
                w_old[i] = wo;
                double wn = v[i] - r3 * woo;
                wn -= r2 * wo;
                wn *= r1;
                w[i] = wn;
                x[i] = x[i] + c * eta * wn;
            }
            norm_rMR = norm_rMR * fabs(s);
            eta = -s * eta;
            residual = norm_rMR / norm_r0;
            __syncthreads();
            if (residual < tol)
                break;
        }
    }
    if (tid == 0) {
        info[0] = (double)it;
        info[1] = residual;
    }
}

// hf += D_b gamma_b for active bodies: with it v = v + M^-1 hf of the velocity update is M_invk + M_invD gamma_b
// (ComputeImpulses).  Thread k owns component k and walks the rows in order: fixed summation order, no atomics.
__global__ void k_bil_apply(uint32_t n, const int32_t* __restrict__ ids, const double* __restrict__ cq,
                            const double* __restrict__ gamma, const uint8_t* __restrict__ bflags, double* hf,
                            const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const int k = threadIdx.x;
    if (k >= 6)
        return;
    for (uint32_t i = 0; i < n; i++) {
        const double g = gamma[i];
        for (int s = 0; s < 2; s++) {
            const uint32_t b = (uint32_t)ids[2 * i + s];
            if (bflags[b] & 1u)
                hf[6 * (size_t)b + k] += cq[12 * (size_t)i + 6 * s + k] * g;
        }
    }
}

// After the contact forces are in hf and before the velocity update (enqueue_smc).
int enqueue_bilateral(System* S) {
    const uint32_t n = S->n_bil;
    if (n == 0)
        return CB200_OK;
    const int g = blocks_for(n, BIL_THREADS);
    k_bil_setup<<<g, BIL_THREADS, 0, S->stream>>>(n, S->bil_ids.p, S->bil_cq.p, S->bil_b.p, S->bflags.p, S->mass.p,
                                                  S->inv_inertia.p, S->vel.p, S->hf.p, S->bil_R.p, S->d_cnt);
    k_bil_matrix<<<grid_for(S, (size_t)n * n, BIL_THREADS), BIL_THREADS, 0, S->stream>>>(
        n, S->bil_ids.p, S->bil_cq.p, S->bflags.p, S->mass.p, S->inv_inertia.p, S->bil_N.p, S->d_cnt);
    k_bil_minres<<<1, BIL_THREADS, 0, S->stream>>>(n, S->bil_max_iter, S->bil_tol, S->bil_N.p, S->bil_R.p,
                                                   S->bil_gamma.p, S->bil_vec.p, S->bil_info.p, S->d_cnt);
    k_bil_apply<<<1, 32, 0, S->stream>>>(n, S->bil_ids.p, S->bil_cq.p, S->bil_gamma.p, S->bflags.p, S->hf.p, S->d_cnt);
    S->launches += 4;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

}  // namespace cb200

using namespace cb200;

// ChSystemMulticore::UpdateBilaterals + ChConstraintBilateral::Build_D / Build_b, BODY_BODY rows: row i couples bodies
// body_ids[2i], body_ids[2i+1] with Jacobian blocks cq_a[6i..], cq_b[6i..] (linear part in the absolute frame, angular
// part in the body frame, as ChConstraintTwoBodies holds them) and right-hand side rhs[i]:  cq_a . v_a + cq_b . v_b +
// rhs = 0.  n = 0 removes all rows.  The rows stay until replaced (a caller with moving joints replaces them every step).
extern "C" int cb200_set_bilaterals(cb200_system* sys, int n, const int32_t* body_ids, const double* cq_a,
                                    const double* cq_b, const double* rhs, int max_iteration_bilateral,
                                    double tol_speed) {
    System* S = reinterpret_cast<System*>(sys);
    if (!S || n < 0 || (n && (!body_ids || !cq_a || !cq_b || !rhs)))
        return set_error(CB200_ERR_ARG, "bad argument to cb200_set_bilaterals");
    if (n && S->set.system_type != 1)
        return set_error(CB200_ERR_UNSUPPORTED, "bilateral rows are implemented for SMC systems only");
    CB_CUDA(cudaSetDevice(S->device));
    for (int i = 0; i < 2 * n; i++)
        if (body_ids[i] < 0 || (uint32_t)body_ids[i] >= S->nb)
            return set_error(CB200_ERR_ARG, "bilateral row refers to a body that does not exist");
    if (n > 0 && S->inertia_off.p)
        return set_error(CB200_ERR_UNSUPPORTED, "bilateral rows use the diagonal of the inverse inertia only "
                                                "(cb200_set_inertia_full was called)");
    if ((uint32_t)n != S->n_bil)
        invalidate_step_graph(S);  // the launch geometry of the bilateral kernels is part of the step graph
    S->n_bil = (uint32_t)n;
    S->bil_max_iter = max_iteration_bilateral;
    S->bil_tol = tol_speed;
    if (n == 0)
        return CB200_OK;
    const size_t cap0 = S->bil_cq.cap;
    CB_TRY(S->bil_ids.reserve(2 * (size_t)n));
    CB_TRY(S->bil_cq.reserve(12 * (size_t)n));
    CB_TRY(S->bil_b.reserve(n));
    CB_TRY(S->bil_R.reserve(n));
    CB_TRY(S->bil_gamma.reserve(n));
    CB_TRY(S->bil_N.reserve((size_t)n * n));
    CB_TRY(S->bil_vec.reserve(7 * (size_t)n));
    CB_TRY(S->bil_info.reserve(2));
    if (S->bil_cq.cap != cap0)
        invalidate_step_graph(S);
    std::vector<double> cq(12 * (size_t)n);
    for (int i = 0; i < n; i++)
        for (int k = 0; k < 6; k++) {
            cq[12 * (size_t)i + k] = cq_a[6 * (size_t)i + k];
            cq[12 * (size_t)i + 6 + k] = cq_b[6 * (size_t)i + k];
        }
    CB_CUDA(cudaMemcpyAsync(S->bil_ids.p, body_ids, 2 * (size_t)n * 4, cudaMemcpyHostToDevice, S->stream));
    CB_CUDA(cudaMemcpyAsync(S->bil_cq.p, cq.data(), cq.size() * 8, cudaMemcpyHostToDevice, S->stream));
    CB_CUDA(cudaMemcpyAsync(S->bil_b.p, rhs, (size_t)n * 8, cudaMemcpyHostToDevice, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}

// gamma_b of the last solve (reaction impulses; ConstraintsFetch_react divides by the step), iterations and residual
extern "C" int cb200_download_bilateral_gamma(cb200_system* sys, double* gamma, int* iterations, double* residual) {
    System* S = reinterpret_cast<System*>(sys);
    if (!S)
        return set_error(CB200_ERR_ARG, "null system");
    CB_CUDA(cudaSetDevice(S->device));
    double info[2] = {0, 0};
    if (S->n_bil) {
        if (gamma)
            CB_CUDA(cudaMemcpyAsync(gamma, S->bil_gamma.p, (size_t)S->n_bil * 8, cudaMemcpyDeviceToHost, S->stream));
        CB_CUDA(cudaMemcpyAsync(info, S->bil_info.p, 16, cudaMemcpyDeviceToHost, S->stream));
    }
    CB_CUDA(cudaStreamSynchronize(S->stream));
    if (iterations)
        *iterations = (int)info[0];
    if (residual)
        *residual = info[1];
    return CB200_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free variant of the MINRES solve.  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
// that forbid it).  The kernel is k_bil_minres itself; what differs is where its 7 work vectors live: instead of
// System::bil_vec (grown with cudaMalloc by cb200_set_bilaterals) they sit in the file-scope array below, which every
// GPU that loads the library holds until the process exits (7 x 2048 doubles = 112 KiB).
//
// In host code `g_minres_scratch` is only a stand-in for the symbol, not a device address: the address a kernel can
// use is looked up once per device, with that device current, by cudaGetSymbolAddress, and cached.
// ---------------------------------------------------------------------------------------------------------------
#include <mutex>

#define CB200_NOALLOC_MINRES_MAX_ROWS 2048
static __device__ __align__(256) double g_minres_scratch[7 * CB200_NOALLOC_MINRES_MAX_ROWS];

static std::mutex g_minres_mu;
static double* g_minres_scratch_ptr[64] = {};  // per device ordinal

static int minres_scratch_for(int device, double** out) {
    if (device < 0 || device >= 64)
        return set_error(CB200_ERR_ARG, "device ordinal out of range");
    std::lock_guard<std::mutex> lock(g_minres_mu);
    if (!g_minres_scratch_ptr[device]) {
        CB_CUDA(cudaSetDevice(device));
        double* p = nullptr;
        CB_CUDA(cudaGetSymbolAddress((void**)&p, g_minres_scratch));
        if (!p)
            return set_error(CB200_ERR_CUDA, "cudaGetSymbolAddress returned a null scratch address");
        g_minres_scratch_ptr[device] = p;
    }
    *out = g_minres_scratch_ptr[device];
    return CB200_OK;
}

// Solves N x = b (N symmetric n x n, row-major; x0 = 0) exactly as the bilateral stage of a step does, on the system's
// stream.  N_dev, b_dev, x_dev (n doubles out) and info_dev (2 doubles out: iterations run, residual) are DEVICE
// pointers owned by the caller.  Calls no cudaMalloc / cudaFree.  All calls on one GPU share the scratch array: two
// calls that overlap on the GPU (different streams, different systems on one device) race; calls on one stream are
// ordered and safe.  The scratch keeps the previous call's contents, so the part this call uses is cleared first.
extern "C" int cb200_minres_noalloc(cb200_system* sys, int n, const double* N_dev, const double* b_dev, double* x_dev,
                                    int max_iter, double tol, double* info_dev) {
    System* S = reinterpret_cast<System*>(sys);
    if (!S || n < 0 || max_iter < 0 || (n && (!N_dev || !b_dev || !x_dev)) || !info_dev)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_minres_noalloc");
    if (n > CB200_NOALLOC_MINRES_MAX_ROWS)
        return set_error(CB200_ERR_ARG, "cb200_minres_noalloc: " + std::to_string(n) + " rows exceed the fixed scratch (" +
                                            std::to_string(CB200_NOALLOC_MINRES_MAX_ROWS) +
                                            " rows); use cb200_set_bilaterals for larger systems");
    double* scratch = nullptr;
    CB_TRY(minres_scratch_for(S->device, &scratch));
    CB_CUDA(cudaSetDevice(S->device));
    if (n)
        CB_CUDA(cudaMemsetAsync(scratch, 0, 7 * (size_t)n * sizeof(double), S->stream));
    k_bil_minres<<<1, BIL_THREADS, 0, S->stream>>>((uint32_t)n, max_iter, tol, N_dev, b_dev, x_dev, scratch, info_dev,
                                                   S->d_cnt);
    S->launches++;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

extern "C" int cb200_minres_noalloc_max_rows(void) { return CB200_NOALLOC_MINRES_MAX_ROWS; }
