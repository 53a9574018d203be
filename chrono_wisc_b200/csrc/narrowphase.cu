// narrowphase.cu -- per-pair contact generation and contact-list build.
//
// Reference path replaced (CMC = src/chrono/collision/multicore):
//   PreprocessLocalToParent   CMC/ChNarrowphase.cpp:173-196   (fused into k_aabb, broadphase.cu)
//   PreprocessCount           CMC/ChNarrowphase.cpp:94-155    -> k_np_slots (+ scan) only when some pair can give >1 contact
//   DispatchPRIMS / Hybrid    CMC/ChNarrowphase.cpp:264-320 + PRIMSCollision (ChNarrowphasePRIMS.cpp:1406-1607)
//                                                            -> k_np_dispatch
//   count + remove_if         CMC/ChNarrowphase.cpp:366-385   -> scan of per-pair contact counts + k_np_compact
//                                                               (stable, so contact order == reference order)
#include "common.cuh"
#include "prims.cuh"

namespace cb200 {

struct ShapeView {
    int type;
    d3 pos;
    q4 rot;
    double d0, d1, d2, d3_;
};

__device__ __forceinline__ ShapeView load_shape(int type, uint32_t s, const double* __restrict__ s_wpos,
                                                const double* __restrict__ s_wrot, const double* __restrict__ s_dims) {
    ShapeView v;
    v.type = type;
    const double4 a = reinterpret_cast<const double4*>(s_wpos)[s];
    v.pos = mk3(a.x, a.y, a.z);
    v.d0 = a.w;
    v.d1 = v.d2 = v.d3_ = 0;
    v.rot = mkq(1, 0, 0, 0);
    if (type != CB200_SPHERE) {
        const double4 r = reinterpret_cast<const double4*>(s_wrot)[s];
        v.rot = mkq(r.x, r.y, r.z, r.w);
        const double4 d = reinterpret_cast<const double4*>(s_dims)[s];
        v.d1 = d.y, v.d2 = d.z, v.d3_ = d.w;
    }
    return v;
}

// max contacts per pair, CMC/ChNarrowphase.cpp:121-134 (MPR: always 1, :99-101)
__device__ __forceinline__ uint32_t max_contacts(int t1, int t2, int algorithm) {
    if (algorithm == 0)
        return 1;
    if (t1 == CB200_SPHERE || t2 == CB200_SPHERE)
        return 1;
    if (t1 == CB200_CAPSULE || t2 == CB200_CAPSULE)
        return 2;
    if (t1 == CB200_CYLSHELL || t2 == CB200_CYLSHELL)
        return 8;
    if (t1 == CB200_BOX && t2 == CB200_BOX)
        return 8;
    if ((t1 == CB200_BOX && t2 == CB200_TRIANGLE) || (t1 == CB200_TRIANGLE && t2 == CB200_BOX))
        return 6;
    return 1;
}

__global__ void __launch_bounds__(256) k_np_slots(uint32_t np, const PairAux* __restrict__ aux, int algorithm,
                                                  uint32_t* __restrict__ slot_index) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < np) {
        const uint32_t t = aux[i].types;
        slot_index[i] = max_contacts(t & 0xff, (t >> 8) & 0xff, algorithm);
    }
}

// Analytic dispatcher for pairs with at most one contact (everything that involves a sphere).
// Returns 1/0 contacts, or -1 when the combination is not analytic (HYBRID then falls back to MPR).
__device__ __forceinline__ int prims_single(const ShapeView& A, const ShapeView& B, double sep, Contact& c) {
    bool hit;
    if (A.type == CB200_SPHERE && B.type == CB200_SPHERE)
        return sphere_sphere(A.pos, A.d0, B.pos, B.d0, sep, c) ? 1 : 0;
    if (B.type == CB200_SPHERE) {
        switch (A.type) {
            case CB200_CAPSULE:
                return capsule_sphere(A.pos, A.rot, A.d0, A.d1, B.pos, B.d0, sep, c) ? 1 : 0;
            case CB200_CYLINDER:
                return cylinder_sphere(A.pos, A.rot, A.d0, A.d1, B.pos, B.d0, sep, c) ? 1 : 0;
            case CB200_ROUNDEDCYL:
                return roundedcyl_sphere(A.pos, A.rot, A.d0, A.d1, A.d3_, B.pos, B.d0, sep, c) ? 1 : 0;
            case CB200_BOX:
                return box_sphere(A.pos, A.rot, mk3(A.d0, A.d1, A.d2), B.pos, B.d0, sep, c) ? 1 : 0;
            case CB200_ROUNDEDBOX:
                return roundedbox_sphere(A.pos, A.rot, mk3(A.d0, A.d1, A.d2), A.d3_, B.pos, B.d0, sep, c) ? 1 : 0;
            default:
                return -1;
        }
    }
    if (A.type == CB200_SPHERE) {
        // swapped call: the shape function sees (B, A); points are exchanged and the normal negated
        // (ChNarrowphasePRIMS.cpp:1435-1443 and siblings)
        switch (B.type) {
            case CB200_CAPSULE:
                hit = capsule_sphere(B.pos, B.rot, B.d0, B.d1, A.pos, A.d0, sep, c);
                break;
            case CB200_CYLINDER:
                hit = cylinder_sphere(B.pos, B.rot, B.d0, B.d1, A.pos, A.d0, sep, c);
                break;
            case CB200_ROUNDEDCYL:
                hit = roundedcyl_sphere(B.pos, B.rot, B.d0, B.d1, B.d3_, A.pos, A.d0, sep, c);
                break;
            case CB200_BOX:
                hit = box_sphere(B.pos, B.rot, mk3(B.d0, B.d1, B.d2), A.pos, A.d0, sep, c);
                break;
            case CB200_ROUNDEDBOX:
                hit = roundedbox_sphere(B.pos, B.rot, mk3(B.d0, B.d1, B.d2), B.d3_, A.pos, A.d0, sep, c);
                break;
            default:
                return -1;
        }
        if (hit) {
            const d3 t = c.p1;
            c.p1 = c.p2;
            c.p2 = t;
            c.n = -c.n;
            return 1;
        }
        return 0;
    }
    return -1;
}

__global__ void __launch_bounds__(256) k_np_dispatch(uint32_t np, const int64_t* __restrict__ pairs,
                                                     const PairAux* __restrict__ aux,
                                                     const uint32_t* __restrict__ slot_index /* null: identity */,
                                                     const double* __restrict__ s_wpos,
                                                     const double* __restrict__ s_wrot,
                                                     const double* __restrict__ s_dims, double envelope, int algorithm,
                                                     double* __restrict__ t_norm, double* __restrict__ t_pta,
                                                     double* __restrict__ t_ptb, double* __restrict__ t_depth,
                                                     double* __restrict__ t_erad, uint32_t* __restrict__ pair_nc,
                                                     DevCounters* __restrict__ cnt) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np)
        return;
    const int64_t key = pairs[i];
    const uint32_t sA = (uint32_t)(key >> 32), sB = (uint32_t)(key & 0xffffffff);
    const uint32_t t = aux[i].types;
    const ShapeView A = load_shape(t & 0xff, sA, s_wpos, s_wrot, s_dims);
    const ShapeView B = load_shape((t >> 8) & 0xff, sB, s_wpos, s_wrot, s_dims);
    const uint32_t slot = slot_index ? slot_index[i] : i;
    Contact c;
    int nC = -1;
    if (algorithm != 0)  // PRIMS or HYBRID
        nC = prims_single(A, B, 2 * envelope, c);
    if (nC < 0) {
        // MPR fallback / multi-contact analytic pairs are not implemented yet in this backend
        atomicAdd(&cnt->unsupported, 1u);
        nC = 0;
    }
    if (nC == 1) {
        t_norm[3 * slot] = c.n.x, t_norm[3 * slot + 1] = c.n.y, t_norm[3 * slot + 2] = c.n.z;
        t_pta[3 * slot] = c.p1.x, t_pta[3 * slot + 1] = c.p1.y, t_pta[3 * slot + 2] = c.p1.z;
        t_ptb[3 * slot] = c.p2.x, t_ptb[3 * slot + 1] = c.p2.y, t_ptb[3 * slot + 2] = c.p2.z;
        t_depth[slot] = c.depth;
        t_erad[slot] = c.erad;
    }
    pair_nc[i] = (uint32_t)nC;
}

__global__ void __launch_bounds__(256) k_np_compact(uint32_t np, const int64_t* __restrict__ pairs,
                                                    const PairAux* __restrict__ aux,
                                                    const uint32_t* __restrict__ slot_index,
                                                    const uint32_t* __restrict__ pair_off /* scanned pair_nc, np+1 */,
                                                    const double* __restrict__ t_norm, const double* __restrict__ t_pta,
                                                    const double* __restrict__ t_ptb, const double* __restrict__ t_depth,
                                                    const double* __restrict__ t_erad, double* __restrict__ c_norm,
                                                    double* __restrict__ c_pta, double* __restrict__ c_ptb,
                                                    double* __restrict__ c_depth, double* __restrict__ c_erad,
                                                    int32_t* __restrict__ c_bids, int64_t* __restrict__ c_shapes,
                                                    uint32_t* __restrict__ c_mats) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= np)
        return;
    const uint32_t o0 = pair_off[i], n = pair_off[i + 1] - o0;
    if (n == 0)
        return;
    const uint32_t slot0 = slot_index ? slot_index[i] : i;
    const PairAux a = aux[i];
    const int64_t key = pairs[i];
    for (uint32_t k = 0; k < n; k++) {
        const uint32_t s = slot0 + k, o = o0 + k;
        c_norm[3 * o] = t_norm[3 * s], c_norm[3 * o + 1] = t_norm[3 * s + 1], c_norm[3 * o + 2] = t_norm[3 * s + 2];
        c_pta[3 * o] = t_pta[3 * s], c_pta[3 * o + 1] = t_pta[3 * s + 1], c_pta[3 * o + 2] = t_pta[3 * s + 2];
        c_ptb[3 * o] = t_ptb[3 * s], c_ptb[3 * o + 1] = t_ptb[3 * s + 1], c_ptb[3 * o + 2] = t_ptb[3 * s + 2];
        c_depth[o] = t_depth[s];
        c_erad[o] = t_erad[s];
        c_bids[2 * o] = (int32_t)a.bodyA;
        c_bids[2 * o + 1] = (int32_t)a.bodyB;
        c_shapes[o] = key;
        c_mats[o] = a.mats;
    }
}

int stage_narrowphase(System* S) {
    S->n_contacts = 0;
    S->n_slots = 0;
    const uint32_t np = S->n_pairs;
    if (np == 0)
        return CB200_OK;
    const int algorithm = S->set.narrowphase_algorithm;
    // multi_slot was raised by the pair store pass; read it (sync point shared with the caller's counters fetch)
    CB_TRY(fetch_counters(S));
    const bool multi = S->h_cnt->multi_slot != 0 && algorithm != 0;
    uint32_t nslots = np;
    const uint32_t* slot_index = nullptr;
    if (multi) {
        CB_TRY(S->slot_index.reserve(np + 1));
        k_np_slots<<<blocks_for(np, 256), 256, 0, S->stream>>>(np, S->pair_aux.p, algorithm, S->slot_index.p);
        S->launches++;
        CB_TRY(scan_u32(S, S->slot_index.p, S->slot_index.p, np));
        CB_CUDA(cudaMemcpyAsync(&nslots, S->slot_index.p + np, 4, cudaMemcpyDeviceToHost, S->stream));
        CB_CUDA(cudaStreamSynchronize(S->stream));
        slot_index = S->slot_index.p;
    }
    S->n_slots = nslots;
    CB_TRY(S->t_norm.reserve(3 * (size_t)nslots));
    CB_TRY(S->t_pta.reserve(3 * (size_t)nslots));
    CB_TRY(S->t_ptb.reserve(3 * (size_t)nslots));
    CB_TRY(S->t_depth.reserve(nslots));
    CB_TRY(S->t_erad.reserve(nslots));
    CB_TRY(S->pair_nc.reserve(np + 1));
    k_np_dispatch<<<blocks_for(np, 256), 256, 0, S->stream>>>(np, S->pairs.p, S->pair_aux.p, slot_index, S->s_wpos.p,
                                                              S->s_wrot.p, S->s_dims.p, S->set.collision_envelope,
                                                              algorithm, S->t_norm.p, S->t_pta.p, S->t_ptb.p,
                                                              S->t_depth.p, S->t_erad.p, S->pair_nc.p, S->d_cnt);
    S->launches++;
    CB_TRY(scan_u32(S, S->pair_nc.p, S->pair_nc.p, np));
    uint32_t nc = 0;
    CB_CUDA(cudaMemcpyAsync(&nc, S->pair_nc.p + np, 4, cudaMemcpyDeviceToHost, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    S->n_contacts = nc;
    if (nc == 0)
        return CB200_OK;
    CB_TRY(S->c_norm.reserve(3 * (size_t)nc));
    CB_TRY(S->c_pta.reserve(3 * (size_t)nc));
    CB_TRY(S->c_ptb.reserve(3 * (size_t)nc));
    CB_TRY(S->c_depth.reserve(nc));
    CB_TRY(S->c_erad.reserve(nc));
    CB_TRY(S->c_bids.reserve(2 * (size_t)nc));
    CB_TRY(S->c_shapes.reserve(nc));
    CB_TRY(S->c_mats.reserve(nc));
    k_np_compact<<<blocks_for(np, 256), 256, 0, S->stream>>>(np, S->pairs.p, S->pair_aux.p, slot_index, S->pair_nc.p,
                                                             S->t_norm.p, S->t_pta.p, S->t_ptb.p, S->t_depth.p,
                                                             S->t_erad.p, S->c_norm.p, S->c_pta.p, S->c_ptb.p,
                                                             S->c_depth.p, S->c_erad.p, S->c_bids.p, S->c_shapes.p,
                                                             S->c_mats.p);
    S->launches++;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

}  // namespace cb200
