// comm.cu -- one system decomposed into slabs across GPUs: the per-step data plane, in the library.
//
// The reference is single-node OpenMP; BASELINE.json's north_star asks for 1-D slabs, ghost-layer exchange each step
// and, for NSC, an all-reduce of the APGD scalars.  One process per GPU; the transfers are NCCL
// send / recv / all-reduce on the system's own stream (no host synchronisation, no Python in between), resolved with
// dlsym from the NCCL library the host program already has loaded (torch's), so there is one NCCL in the process.
//
// What moves:
//   ghost records   after every step: bodies within the ghost width of a cut plane, packed by k_slab_pack (slab.cu),
//                   sent to that neighbour, unpacked into its ghost region (arrivals adopted)
//   NSC halo        inside every Schur product: the owner's body-space vector (6 doubles per ghost body, M^-1 D x;
//                   once per solve also M_invk) to the neighbours' ghost copies.  Contacts in the overlap exist on
//                   both ranks with identical arithmetic, so only this broadcast is needed -- no halo sum
//   NSC scalars     one all-reduce of 16 doubles per APGD slot (the dot products of ChSolverMulticoreAPGD.cpp:145,147,
//                   160,173, each contact summed on exactly one rank)
//
// The same code drives several slabs living in ONE process on one stream ("local group": the transfers become
// device-to-device copies and the all-reduce a small kernel).  That is how the decomposition is tested on a single
// GPU; nothing waits on another launch.
#include "common.cuh"
#include "slab.cuh"
#include <dlfcn.h>
#include <nccl.h>
#include <algorithm>

namespace cb200 {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

static int nccl_load(NcclApi& A, const char* path) {
    if (A.lib)
        return CB200_OK;
    A.lib = dlopen(path && path[0] ? path : "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!A.lib)
        return set_error(CB200_ERR_STATE, std::string("cannot load NCCL: ") + dlerror());
#define SYM(field, name)                                                            \
    *(void**)(&A.field) = dlsym(A.lib, name);                                       \
    if (!A.field)                                                                   \
        return set_error(CB200_ERR_STATE, std::string("NCCL symbol missing: ") + name)
    SYM(GetUniqueId, "ncclGetUniqueId");
    SYM(CommInitRank, "ncclCommInitRank");
    SYM(CommDestroy, "ncclCommDestroy");
    SYM(GroupStart, "ncclGroupStart");
    SYM(GroupEnd, "ncclGroupEnd");
    SYM(Send, "ncclSend");
    SYM(Recv, "ncclRecv");
    SYM(AllReduce, "ncclAllReduce");
    SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    return CB200_OK;
}

#define CB_NCCL(C, call)                                                                                        \
    do {                                                                                                        \
        ncclResult_t r__ = (call);                                                                              \
        if (r__ != ncclSuccess)                                                                                 \
            return set_error(CB200_ERR_CUDA, std::string(#call) + ": " + (C)->api.GetErrorString(r__));         \
    } while (0)

// What one slab knows about its place in the decomposed system.
struct Comm {
    int rank = 0, world = 1;
    NcclApi api;
    ncclComm_t nccl = nullptr;      // multi-process
    std::vector<System*> local;     // single-process emulation: all slabs, in rank order (one stream)
    DevBuf<double> send[2], recv[2];    // ghost records, left / right neighbour (slab.cu layout)
    DevBuf<double> hsend[2], hrecv[2];  // NSC halo: 6 doubles per record of the last ghost exchange
    DevBuf<double> gsum;                // 24 doubles: sums over all ranks
    uint32_t n_send[2] = {0, 0}, n_recv[2] = {0, 0};  // records of the last ghost exchange (host copies)
    bool counts_valid = false;
    double* h_counts = nullptr;         // pinned, 4 doubles
};

bool comm_multi(const System* S) { return S->comm != nullptr && S->comm->world > 1; }
double* comm_gsum(System* S) { return S->comm->gsum.p; }

static int comm_alloc(System* S) {
    Comm* C = S->comm;
    const size_t nrec = (size_t)cb200_slab_buffer_doubles(reinterpret_cast<cb200_system*>(S));
    for (int side = 0; side < 2; side++) {
        CB_TRY(C->send[side].reserve(nrec));
        CB_TRY(C->recv[side].reserve(nrec));
        CB_TRY(C->hsend[side].reserve(6 * (size_t)S->slab_ghost_cap + 8));
        CB_TRY(C->hrecv[side].reserve(6 * (size_t)S->slab_ghost_cap + 8));
        CB_CUDA(cudaMemsetAsync(C->send[side].p, 0, nrec * 8, S->stream));
        CB_CUDA(cudaMemsetAsync(C->recv[side].p, 0, nrec * 8, S->stream));
    }
    CB_TRY(C->gsum.reserve(24));
    CB_CUDA(cudaMemsetAsync(C->gsum.p, 0, 24 * 8, S->stream));
    if (!C->h_counts)
        CB_CUDA(cudaMallocHost(&C->h_counts, 4 * sizeof(double)));
    return CB200_OK;
}

// ---- transfers: NCCL between processes, device copies inside a local group ---------------------------------------
enum { BUF_RECORDS = 0, BUF_HALO = 1 };

// `doubles`: message length when it is the same for everybody (capacity); 0: the record counts of the last ghost
// exchange times `per_record`.
static int comm_sendrecv(Group& g, int which, size_t doubles, int per_record) {
    System* S0 = g.sys[0];
    Comm* C0 = S0->comm;
    if (C0->nccl) {
        DevBuf<double>* snd = which == BUF_RECORDS ? C0->send : C0->hsend;
        DevBuf<double>* rcv = which == BUF_RECORDS ? C0->recv : C0->hrecv;
        const int peer[2] = {C0->rank - 1, C0->rank + 1};
        const int has[2] = {S0->slab_has_left, S0->slab_has_right};
        CB_NCCL(C0, C0->api.GroupStart());
        for (int side = 0; side < 2; side++) {
            if (!has[side])
                continue;
            const size_t ns = doubles ? doubles : (size_t)per_record * C0->n_send[side];
            const size_t nr = doubles ? doubles : (size_t)per_record * C0->n_recv[side];
            if (ns)
                CB_NCCL(C0, C0->api.Send(snd[side].p, ns, ncclDouble, peer[side], C0->nccl, S0->stream));
            if (nr)
                CB_NCCL(C0, C0->api.Recv(rcv[side].p, nr, ncclDouble, peer[side], C0->nccl, S0->stream));
        }
        CB_NCCL(C0, C0->api.GroupEnd());
        return CB200_OK;
    }
    // local group: slab i's right neighbour is slab i + 1
    for (size_t i = 0; i + 1 < g.sys.size(); i++) {
        Comm *A = g.sys[i]->comm, *B = g.sys[i + 1]->comm;
        DevBuf<double>* as = which == BUF_RECORDS ? A->send : A->hsend;
        DevBuf<double>* ar = which == BUF_RECORDS ? A->recv : A->hrecv;
        DevBuf<double>* bs = which == BUF_RECORDS ? B->send : B->hsend;
        DevBuf<double>* br = which == BUF_RECORDS ? B->recv : B->hrecv;
        const size_t n_ab = doubles ? doubles : (size_t)per_record * A->n_send[1];
        const size_t n_ba = doubles ? doubles : (size_t)per_record * B->n_send[0];
        if (n_ab)
            CB_CUDA(cudaMemcpyAsync(br[0].p, as[1].p, n_ab * 8, cudaMemcpyDeviceToDevice, g.sys[0]->stream));
        if (n_ba)
            CB_CUDA(cudaMemcpyAsync(ar[1].p, bs[0].p, n_ba * 8, cudaMemcpyDeviceToDevice, g.sys[0]->stream));
    }
    return CB200_OK;
}

struct SumPtrs {
    double* p[16];
    int n;
};
__global__ void k_local_allreduce(SumPtrs P, int first, int count) {
    const int k = threadIdx.x;
    if (k >= count)
        return;
    double s = 0;
    for (int r = 0; r < P.n; r++)  // rank order: every slab gets the same bits
        s += P.p[r][first + k];
    for (int r = 0; r < P.n; r++)
        P.p[r][first + k] = s;
}

int comm_allreduce(Group& g, int first, int n) {
    System* S0 = g.sys[0];
    Comm* C0 = S0->comm;
    if (C0->nccl) {
        CB_NCCL(C0, C0->api.AllReduce(C0->gsum.p + first, C0->gsum.p + first, (size_t)n, ncclDouble, ncclSum, C0->nccl,
                                      S0->stream));
        return CB200_OK;
    }
    if (g.sys.size() > 16 || n > 32)
        return set_error(CB200_ERR_UNSUPPORTED, "local groups hold at most 16 slabs");
    SumPtrs P;
    P.n = (int)g.sys.size();
    for (int r = 0; r < P.n; r++)
        P.p[r] = g.sys[r]->comm->gsum.p;
    k_local_allreduce<<<1, 32, 0, S0->stream>>>(P, first, n);
    S0->launches++;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

// ---- NSC halo: the owner's 6 doubles per body -> the ghost copies on the neighbours -------------------------------
// The records of the last ghost exchange define the mapping: the j-th record a rank sent to a side is the j-th ghost
// slot of that neighbour's region for this rank (slab.cu); the flags and CTA offsets of that pack are still there
// (slab.cuh).
__global__ void __launch_bounds__(SLAB_CTA) k_halo_pack(uint32_t owned_cap, uint32_t nb_static,
                                                        const uint8_t* __restrict__ flags,
                                                        const uint32_t* __restrict__ cta_off,
                                                        const double* __restrict__ arr, double* __restrict__ send_left,
                                                        double* __restrict__ send_right,
                                                        const DevCounters* __restrict__ cnt) {
    __shared__ uint32_t wtot[3][8];
    if (cnt->overflow)
        return;
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per owned slot, as in the pack pass
    const uint32_t f = s < owned_cap ? flags[s] : 0u;
    uint32_t rank[3];
    slab_ranks(f, cta_off, wtot, rank);
    if (s >= owned_cap)
        return;
    const double* a = arr + 6 * (size_t)(nb_static + s);
    if ((f & SLAB_F_LEFT) && send_left) {
        double* d = send_left + 6 * (size_t)rank[0];
        for (int k = 0; k < 6; k++)
            d[k] = a[k];
    }
    if ((f & SLAB_F_RIGHT) && send_right) {
        double* d = send_right + 6 * (size_t)rank[1];
        for (int k = 0; k < 6; k++)
            d[k] = a[k];
    }
}

__global__ void __launch_bounds__(256) k_halo_unpack(uint32_t ghost_cap, uint32_t ghost_base,
                                                     const double* __restrict__ records /* header: count */,
                                                     const double* __restrict__ halo, const uint8_t* __restrict__ bflags,
                                                     double* __restrict__ arr, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t n = min((uint32_t)records[0], ghost_cap);
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        const uint32_t b = ghost_base + j;
        if (bflags[b] & 4u)
            continue;  // the record was an arrival: the body lives in an owned slot now, its vector is computed here
        for (int k = 0; k < 6; k++)
            arr[6 * (size_t)b + k] = halo[6 * (size_t)j + k];
    }
}

int comm_halo6(Group& g, int which) {
    for (System* S : g.sys) {
        Comm* C = S->comm;
        const double* arr = which == 0 ? S->n_w.p : (which == 1 ? S->n_w0.p : S->n_wy.p);
        k_halo_pack<<<(S->slab_owned_cap + SLAB_CTA - 1) / SLAB_CTA, SLAB_CTA, 0, S->stream>>>(
            S->slab_owned_cap, S->slab_nb_static, reinterpret_cast<const uint8_t*>(S->slab_flags.p), S->slab_cta.p, arr,
            S->slab_has_left ? C->hsend[0].p : nullptr, S->slab_has_right ? C->hsend[1].p : nullptr, S->d_cnt);
        S->launches++;
    }
    const bool exact = g.sys[0]->comm->counts_valid;
    CB_TRY(comm_sendrecv(g, BUF_HALO, exact ? 0 : 6 * (size_t)g.sys[0]->slab_ghost_cap, 6));
    for (System* S : g.sys) {
        Comm* C = S->comm;
        double* arr = which == 0 ? S->n_w.p : (which == 1 ? S->n_w0.p : S->n_wy.p);
        const uint32_t gl = S->slab_nb_static + S->slab_owned_cap, gr = gl + S->slab_ghost_cap;
        const int grid = grid_for(S, S->slab_ghost_cap, 256);
        if (S->slab_has_left)
            k_halo_unpack<<<grid, 256, 0, S->stream>>>(S->slab_ghost_cap, gl, C->recv[0].p, C->hrecv[0].p, S->bflags.p, arr,
                                                       S->d_cnt);
        if (S->slab_has_right)
            k_halo_unpack<<<grid, 256, 0, S->stream>>>(S->slab_ghost_cap, gr, C->recv[1].p, C->hrecv[1].p, S->bflags.p, arr,
                                                       S->d_cnt);
        S->launches += 2;
        CB_CUDA(cudaGetLastError());
    }
    return CB200_OK;
}

// ---- ghost exchange and stepping ---------------------------------------------------------------------------------
static int comm_exchange_ghosts(Group& g) {
    for (System* S : g.sys) {
        Comm* C = S->comm;
        CB_TRY(cb200_slab_pack_dev(reinterpret_cast<cb200_system*>(S), S->slab_has_left ? C->send[0].p : nullptr,
                                   S->slab_has_right ? C->send[1].p : nullptr));
        C->counts_valid = false;
    }
    // at capacity: the counts live on the device and a step must not wait for the host (SMC steps take ~1 ms)
    CB_TRY(comm_sendrecv(g, BUF_RECORDS, (size_t)cb200_slab_buffer_doubles(reinterpret_cast<cb200_system*>(g.sys[0])), 0));
    for (System* S : g.sys) {
        Comm* C = S->comm;
        CB_TRY(cb200_slab_unpack_dev(reinterpret_cast<cb200_system*>(S), C->recv[0].p, C->recv[1].p));
    }
    return CB200_OK;
}

// record counts of the last exchange -> host (one synchronisation): sizes of the NSC halo messages of the next solve
static int comm_read_counts(Group& g) {
    for (System* S : g.sys) {
        Comm* C = S->comm;
        for (int k = 0; k < 4; k++) {
            const double* src = (k < 2 ? C->send[k] : C->recv[k - 2]).p;
            CB_CUDA(cudaMemcpyAsync(C->h_counts + k, src, sizeof(double), cudaMemcpyDeviceToHost, S->stream));
        }
    }
    for (System* S : g.sys)
        CB_CUDA(cudaStreamSynchronize(S->stream));
    for (System* S : g.sys) {
        Comm* C = S->comm;
        C->n_send[0] = S->slab_has_left ? (uint32_t)C->h_counts[0] : 0u;
        C->n_send[1] = S->slab_has_right ? (uint32_t)C->h_counts[1] : 0u;
        C->n_recv[0] = S->slab_has_left ? (uint32_t)C->h_counts[2] : 0u;
        C->n_recv[1] = S->slab_has_right ? (uint32_t)C->h_counts[3] : 0u;
        C->counts_valid = true;
    }
    return CB200_OK;
}

// overflow mask of any rank / slab != 0 (one synchronisation; the flag itself travels as a double through the all-reduce)
__global__ void k_overflow_flag(const DevCounters* __restrict__ cnt, double* __restrict__ out) {
    *out = cnt->overflow ? 1.0 : 0.0;
}
static int comm_any_overflow(Group& g, bool& any) {
    for (System* S : g.sys) {
        k_overflow_flag<<<1, 1, 0, S->stream>>>(S->d_cnt, S->comm->gsum.p + 20);
        S->launches++;
    }
    CB_TRY(comm_allreduce(g, 20, 1));
    for (System* S : g.sys) {
        CB_CUDA(cudaMemcpyAsync(S->comm->h_counts, S->comm->gsum.p + 20, sizeof(double), cudaMemcpyDeviceToHost, S->stream));
        CB_TRY(fetch_counters(S));  // (synchronises the stream)
    }
    any = false;
    for (System* S : g.sys)
        any = any || S->comm->h_counts[0] != 0.0;
    return CB200_OK;
}

static int comm_group_of(cb200_system** systems, int n, Group& g) {
    if (!systems || n < 1)
        return set_error(CB200_ERR_ARG, "empty system list");
    for (int i = 0; i < n; i++) {
        System* S = reinterpret_cast<System*>(systems[i]);
        if (!S || !S->slab_on || !S->comm)
            return set_error(CB200_ERR_STATE, "the system is not a slab with a communicator (cb200_slab_configure, "
                                              "cb200_comm_init_nccl / cb200_comm_init_local)");
        if (i && S->stream != reinterpret_cast<System*>(systems[0])->stream)
            return set_error(CB200_ERR_STATE, "the slabs of a local group must share one stream (cb200_set_stream)");
        g.sys.push_back(S);
    }
    if (n > 1 && g.sys[0]->comm->nccl)
        return set_error(CB200_ERR_ARG, "one system per process when the communicator is NCCL");
    return CB200_OK;
}

void comm_release(System* S) {
    Comm* C = S->comm;
    if (!C)
        return;
    if (C->nccl)
        C->api.CommDestroy(C->nccl);
    for (int s = 0; s < 2; s++)
        C->send[s].release(), C->recv[s].release(), C->hsend[s].release(), C->hrecv[s].release();
    C->gsum.release();
    if (C->h_counts)
        cudaFreeHost(C->h_counts);
    delete C;
    S->comm = nullptr;
}

}  // namespace cb200

using namespace cb200;

extern "C" {

// ncclGetUniqueId of the NCCL library at `nccl_path` (rank 0 calls it and hands the 128 bytes to the other ranks).
int cb200_comm_unique_id(const char* nccl_path, char* out128) {
    static NcclApi api;
    CB_TRY(nccl_load(api, nccl_path));
    ncclUniqueId id;
    if (api.GetUniqueId(&id) != ncclSuccess)
        return set_error(CB200_ERR_CUDA, "ncclGetUniqueId failed");
    memcpy(out128, id.internal, NCCL_UNIQUE_ID_BYTES);
    return CB200_OK;
}

// This process's slab joins an NCCL communicator of `world` ranks (slab i is rank i; neighbours are rank +- 1).
int cb200_comm_init_nccl(cb200_system* sys, const char* nccl_path, int rank, int world, const char* id128) {
    System* S = reinterpret_cast<System*>(sys);
    if (!S || !S->slab_on || !id128 || rank < 0 || rank >= world)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_comm_init_nccl (configure the slab first)");
    CB_CUDA(cudaSetDevice(S->device));
    comm_release(S);
    Comm* C = new Comm();
    S->comm = C;
    C->rank = rank, C->world = world;
    CB_TRY(nccl_load(C->api, nccl_path));
    ncclUniqueId id;
    memcpy(id.internal, id128, NCCL_UNIQUE_ID_BYTES);
    CB_NCCL(C, C->api.CommInitRank(&C->nccl, world, id, rank));
    invalidate_step_graph(S);
    return comm_alloc(S);
}

// All slabs of a decomposed system in ONE process (rank order), sharing one stream: tests and single-GPU emulation.
int cb200_comm_init_local(cb200_system** systems, int n) {
    if (!systems || n < 1)
        return set_error(CB200_ERR_ARG, "empty system list");
    std::vector<System*> all;
    for (int i = 0; i < n; i++) {
        System* S = reinterpret_cast<System*>(systems[i]);
        if (!S || !S->slab_on)
            return set_error(CB200_ERR_STATE, "cb200_slab_configure was not called");
        all.push_back(S);
    }
    for (int i = 0; i < n; i++) {
        System* S = all[i];
        CB_CUDA(cudaSetDevice(S->device));
        comm_release(S);
        Comm* C = new Comm();
        S->comm = C;
        C->rank = i, C->world = n;
        C->local = all;
        invalidate_step_graph(S);
        CB_TRY(comm_alloc(S));
    }
    return CB200_OK;
}

// Ghost-layer exchange of the listed slabs (one per process over NCCL; all of a local group).  Asynchronous for SMC;
// NSC systems also fetch the record counts (one synchronisation) and run the exchange twice, so that a body handed
// over in the first round is already in its new owner's send list (its halo values must reach the old owner).
int cb200_comm_exchange(cb200_system** systems, int n) {
    Group g;
    CB_TRY(comm_group_of(systems, n, g));
    CB_CUDA(cudaSetDevice(g.sys[0]->device));
    CB_TRY(comm_exchange_ghosts(g));
    if (g.sys[0]->set.system_type != 1) {
        CB_TRY(comm_exchange_ghosts(g));
        CB_TRY(comm_read_counts(g));
    }
    return CB200_OK;
}

// nsteps x (DoStepDynamics of every listed slab + ghost exchange), enqueued from here: no Python between the step, the
// pack / unpack kernels and the transfers.  SMC: fully asynchronous (cb200_check reports capacity overflows).  NSC:
// the solve runs slot by slot over the group with the halo and all-reduce transfers in between.
int cb200_comm_steps(cb200_system** systems, int n, int nsteps) {
    Group g;
    CB_TRY(comm_group_of(systems, n, g));
    CB_CUDA(cudaSetDevice(g.sys[0]->device));
    const bool smc = g.sys[0]->set.system_type == 1;
    for (System* S : g.sys)
        CB_TRY(ensure_capacity(S));
    for (int step = 0; step < nsteps; step++) {
        if (smc) {
            for (System* S : g.sys)
                CB_TRY(cb200_enqueue_steps(reinterpret_cast<cb200_system*>(S), 1));
            CB_TRY(comm_exchange_ghosts(g));
            continue;
        }
        if (!g.sys[0]->comm->counts_valid)
            CB_TRY(comm_read_counts(g));
        for (int attempt = 0;; attempt++) {
            for (System* S : g.sys) {
                CB_TRY(timer_start(S, "step"));
                CB_TRY(enqueue_update_forces(S));
                CB_TRY(timer_start(S, "collision"));
                CB_TRY(enqueue_aabb(S));
                CB_TRY(enqueue_bins(S));
                CB_TRY(enqueue_pairs_only(S, false, true, false));
                CB_TRY(timer_stop(S, "collision"));
                CB_TRY(timer_start(S, "advance"));
            }
            CB_TRY(enqueue_nsc_group(g));
            for (System* S : g.sys) {
                CB_TRY(timer_stop(S, "advance"));
                CB_TRY(enqueue_integrate(S));
                CB_TRY(timer_stop(S, "step"));
            }
            // Did a capacity (contacts, entries, APGD slot budget ...) overflow anywhere?  Every kernel after the overflow
            // was a no-op on that rank, the transfers still ran, nothing was integrated: all ranks redo the step together.
            bool any = false;
            CB_TRY(comm_any_overflow(g, any));
            if (!any)
                break;
            if (attempt >= 8)
                return set_error(CB200_ERR_STATE, "capacity growth did not converge");
            int extra = 0;
            for (System* S : g.sys) {
                if (S->h_cnt->overflow)
                    CB_TRY(grow_after_overflow(S));
                extra = std::max(extra, S->nsc_extra_slots);
            }
            for (System* S : g.sys)
                S->nsc_extra_slots = extra;
        }
        for (System* S : g.sys)
            S->timed_steps++;
        CB_TRY(comm_exchange_ghosts(g));
        CB_TRY(comm_exchange_ghosts(g));
        CB_TRY(comm_read_counts(g));  // (the one host synchronisation of an NSC slab step)
    }
    return CB200_OK;
}

}  // extern "C"
