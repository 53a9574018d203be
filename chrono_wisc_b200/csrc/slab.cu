// slab.cu -- spatial slab decomposition of one system across GPUs: device-side pack / unpack of ghost-layer bodies.
//
// The reference is single-node OpenMP; BASELINE.json's north_star asks for 1-D slabs with a ghost-layer exchange
// each step.  One process per GPU owns the bodies whose centre lies in [lo, hi) along `axis`.
//
// Slot model (no array ever changes size, so a step stays free of host synchronisation):
//   bodies [0, nb_static)                       static geometry replicated on every rank (container walls)
//   bodies [nb_static, +owned_cap)              owned region: bodies this rank integrates; vacant slots in between
//   then ghost_cap slots for the left neighbour's ghost layer, ghost_cap for the right neighbour's
// Every dynamic slot b owns exactly one shape, ns_static + (b - nb_static).  A vacant slot has bflags bit 2 set and
// is invisible to collision (k_build_meta turns its shape into a removed one).  Ghost bodies are inactive
// (not integrated, M_inv = 0) but collide: a contact between an owned body and a ghost is evaluated on both ranks
// from the same state and applied to the owned body only, so no force is exchanged.
//
// Per step, after integration:
//   pack    one pass over the owned region + one packed 64-bit scan: bodies within `width` of a cut plane (or past
//           it) are written, in slot order, to that neighbour's send buffer; bodies past the plane are handed over
//           (the neighbour adopts them; here the slot lingers one step as a ghost, then becomes vacant); the list
//           of vacant owned slots is rebuilt in slot order.
//   (the caller moves the two buffers: NCCL send/recv between neighbours, chrono_wisc_b200/slab.py)
//   unpack  the ghost regions are rebuilt from scratch from the received records; a record whose centre lies in
//           [lo, hi) is an arrival and takes the next vacant owned slot (in record order: deterministic).
// Record = 33 doubles: gid, pos 3, rot 4, vel 6, mass, inertia 3, inv_inertia 3, shape dims 4, shape local pos 3,
// local rot 4, packed (type | material << 8 | family << 32).  Buffer = [count, 0, records...].
#include "common.cuh"
#include "slab.cuh"
#include <float.h>
#include <limits.h>
#include <vector>

namespace cb200 {

enum { SLAB_REC = 33, SLAB_HDR = 2 };

struct SlabArrays {
    double *pos, *rot, *vel, *mass, *inertia, *inv_inertia;
    int64_t* gid;
    uint8_t* bflags;
    double *s_dims, *s_lpos, *s_lrot;
    int32_t *s_type, *s_mat;
    uint32_t* s_fam;
    ShapeMeta* meta;  // kept current slot by slot (what k_build_meta would write), so an exchange does not force a rebuild
    uint32_t nb_static, ns_static;
};

// The integer record of the shape of dynamic slot b, as k_build_meta (broadphase.cu) forms it for a slab system
__device__ __forceinline__ void slab_write_meta(const SlabArrays& A, uint32_t b, uint32_t fl) {
    const uint32_t s = A.ns_static + (b - A.nb_static);
    ShapeMeta m;
    m.body = (fl & 4u) ? UINT_MAX : b;
    m.fam = A.s_fam[s];
    m.pad = 0;
    const uint32_t type = (uint32_t)A.s_type[s] & 0xff;
    m.info = type << 8 | ((uint32_t)A.s_mat[s] & 0xffff) << 16 | (fl & 3u);
    if (type == CB200_SPHERE && A.s_lpos[3 * (size_t)s] == 0 && A.s_lpos[3 * (size_t)s + 1] == 0 &&
        A.s_lpos[3 * (size_t)s + 2] == 0)
        m.info |= 4u;
    A.meta[s] = m;
}

// Both directions fetch the whole record into registers before the first store: with a store after every load the
// in-order issue waits for each load in turn (the arrays may alias as far as the compiler knows), 33 round trips in a row
// for the one or two flagged slots of a CTA -- which was most of the duration of the pack and unpack kernels.
__device__ __forceinline__ void slab_write_record(const SlabArrays& A, uint32_t b, double* __restrict__ r) {
    const uint32_t s = A.ns_static + (b - A.nb_static);
    double v[SLAB_REC];
    v[0] = (double)A.gid[b];
#pragma unroll
    for (int k = 0; k < 3; k++)
        v[1 + k] = A.pos[3 * (size_t)b + k];
#pragma unroll
    for (int k = 0; k < 4; k++)
        v[4 + k] = A.rot[4 * (size_t)b + k];
#pragma unroll
    for (int k = 0; k < 6; k++)
        v[8 + k] = A.vel[6 * (size_t)b + k];
    v[14] = A.mass[b];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        v[15 + k] = A.inertia[3 * (size_t)b + k];
        v[18 + k] = A.inv_inertia[3 * (size_t)b + k];
        v[25 + k] = A.s_lpos[3 * (size_t)s + k];
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        v[21 + k] = A.s_dims[4 * (size_t)s + k];
        v[28 + k] = A.s_lrot[4 * (size_t)s + k];
    }
    const unsigned long long bits = ((unsigned long long)A.s_type[s] & 0xffull) |
                                    (((unsigned long long)A.s_mat[s] & 0xffffull) << 8) |
                                    ((unsigned long long)A.s_fam[s] << 32);
    v[32] = __longlong_as_double((long long)bits);
#pragma unroll
    for (int k = 0; k < SLAB_REC; k++)
        r[k] = v[k];
}

// The same record written by a whole warp: lane k fetches word k (ten short divergent paths, every load in flight at
// once, no record held in registers) -- the pack kernel keeps the occupancy its 5000 CTAs need.
__device__ __forceinline__ void slab_write_record_warp(const SlabArrays& A, uint32_t b, double* __restrict__ r,
                                                       int lane) {
    const size_t s = A.ns_static + (b - A.nb_static);
    double v;
    if (lane == 0)
        v = (double)A.gid[b];
    else if (lane < 4)
        v = A.pos[3 * (size_t)b + (lane - 1)];
    else if (lane < 8)
        v = A.rot[4 * (size_t)b + (lane - 4)];
    else if (lane < 14)
        v = A.vel[6 * (size_t)b + (lane - 8)];
    else if (lane == 14)
        v = A.mass[b];
    else if (lane < 18)
        v = A.inertia[3 * (size_t)b + (lane - 15)];
    else if (lane < 21)
        v = A.inv_inertia[3 * (size_t)b + (lane - 18)];
    else if (lane < 25)
        v = A.s_dims[4 * s + (lane - 21)];
    else if (lane < 28)
        v = A.s_lpos[3 * s + (lane - 25)];
    else
        v = A.s_lrot[4 * s + (lane - 28)];
    r[lane] = v;
    if (lane == 0) {
        const unsigned long long bits = ((unsigned long long)A.s_type[s] & 0xffull) |
                                        (((unsigned long long)A.s_mat[s] & 0xffffull) << 8) |
                                        ((unsigned long long)A.s_fam[s] << 32);
        r[32] = __longlong_as_double((long long)bits);
    }
}

__device__ __forceinline__ void slab_read_record(const SlabArrays& A, uint32_t b, const double* __restrict__ r,
                                                 uint8_t flags) {
    const uint32_t s = A.ns_static + (b - A.nb_static);
    double v[SLAB_REC];
#pragma unroll
    for (int k = 0; k < SLAB_REC; k++)
        v[k] = r[k];
    A.gid[b] = (int64_t)v[0];
#pragma unroll
    for (int k = 0; k < 3; k++)
        A.pos[3 * (size_t)b + k] = v[1 + k];
#pragma unroll
    for (int k = 0; k < 4; k++)
        A.rot[4 * (size_t)b + k] = v[4 + k];
#pragma unroll
    for (int k = 0; k < 6; k++)
        A.vel[6 * (size_t)b + k] = v[8 + k];
    A.mass[b] = v[14];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        A.inertia[3 * (size_t)b + k] = v[15 + k];
        A.inv_inertia[3 * (size_t)b + k] = v[18 + k];
        A.s_lpos[3 * (size_t)s + k] = v[25 + k];
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        A.s_dims[4 * (size_t)s + k] = v[21 + k];
        A.s_lrot[4 * (size_t)s + k] = v[28 + k];
    }
    const unsigned long long bits = (unsigned long long)__double_as_longlong(v[32]);
    const uint32_t type = (uint32_t)(bits & 0xff), mat = (uint32_t)((bits >> 8) & 0xffff), fam = (uint32_t)(bits >> 32);
    A.s_type[s] = (int32_t)type;
    A.s_mat[s] = (int32_t)mat;
    A.s_fam[s] = fam;
    A.bflags[b] = flags;
    // (slab_write_meta without reading back what was just stored)
    ShapeMeta m;
    m.body = (flags & 4u) ? UINT_MAX : b;
    m.fam = fam;
    m.pad = 0;
    m.info = type << 8 | mat << 16 | (flags & 3u);
    if (type == CB200_SPHERE && v[25] == 0 && v[26] == 0 && v[27] == 0)
        m.info |= 4u;
    A.meta[s] = m;
}

struct SlabGeom {
    int axis, has_left, has_right;
    double lo, hi, w;
    uint32_t owned_cap, ghost_cap;
};

// Pack, pass 1: flags of one owned slot (slab.cuh) and the three counts of every CTA of 256 slots.  One thread per slot.
__global__ void __launch_bounds__(SLAB_CTA) k_slab_flags(SlabGeom G, SlabArrays A, uint8_t* __restrict__ flags,
                                                         uint32_t* __restrict__ cta_cnt) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t f = 0;
    if (s < G.owned_cap) {
        const uint32_t b = A.nb_static + s;
        if (A.bflags[b] & (4u | 8u)) {
            f = SLAB_F_VACANT;  // vacant, or left during the previous exchange (kept one step as a ghost, see k_slab_pack)
        } else {
            const double p = A.pos[3 * (size_t)b + G.axis];
            if (G.has_left && p < G.lo + G.w)
                f |= SLAB_F_LEFT;
            if (G.has_right && p >= G.hi - G.w)
                f |= SLAB_F_RIGHT;
        }
        flags[s] = (uint8_t)f;
    }
    const int cl = __syncthreads_count(f & 1u), cr = __syncthreads_count(f & 2u), cv = __syncthreads_count(f & 4u);
    if (threadIdx.x == 0) {
        cta_cnt[3 * blockIdx.x] = (uint32_t)cl;
        cta_cnt[3 * blockIdx.x + 1] = (uint32_t)cr;
        cta_cnt[3 * blockIdx.x + 2] = (uint32_t)cv;
    }
}

// Pass 2: exclusive scan of the per-CTA counts (a few thousand triples: one CTA), totals behind the last triple
__global__ void __launch_bounds__(1024) k_slab_cta_scan(uint32_t ncta, uint32_t* __restrict__ cta_cnt) {
    __shared__ uint32_t sm[3][32];
    __shared__ uint32_t carry[3];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x < 3)
        carry[threadIdx.x] = 0;
    __syncthreads();
    // chunks of 1024 triples, a triple per thread (consecutive threads, consecutive words), the running totals carried
    for (uint32_t c0 = 0; c0 < ncta; c0 += blockDim.x) {
        const uint32_t c = c0 + threadIdx.x;
        uint32_t t[3] = {0, 0, 0}, in[3];
        if (c < ncta)
            for (int k = 0; k < 3; k++)
                t[k] = cta_cnt[3 * c + k];
        for (int k = 0; k < 3; k++) {
            uint32_t v = t[k];
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xffffffffu, v, d);
                if (lane >= d)
                    v += o;
            }
            if (lane == 31)
                sm[k][wid] = v;
            in[k] = v;
        }
        __syncthreads();
        if (wid == 0) {
            for (int k = 0; k < 3; k++) {
                const uint32_t w = sm[k][lane];
                uint32_t v = w;
                for (int d = 1; d < 32; d <<= 1) {
                    const uint32_t o = __shfl_up_sync(0xffffffffu, v, d);
                    if (lane >= d)
                        v += o;
                }
                sm[k][lane] = v - w;  // exclusive over the warps
            }
        }
        __syncthreads();
        uint32_t tot[3];
        for (int k = 0; k < 3; k++) {
            const uint32_t base = carry[k];
            if (c < ncta)
                cta_cnt[3 * c + k] = base + sm[k][wid] + in[k] - t[k];
            tot[k] = base + sm[k][wid] + in[k];  // inclusive: the last thread's is the new carry
        }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1)
            for (int k = 0; k < 3; k++)
                carry[k] = tot[k];
        __syncthreads();
    }
    if (threadIdx.x < 3)
        cta_cnt[3 * ncta + threadIdx.x] = carry[threadIdx.x];  // totals
}

// Pass 3: records into the send buffers, vacant list, departures; the slot order is kept (CTA offset + rank in CTA)
__global__ void __launch_bounds__(SLAB_CTA) k_slab_pack(SlabGeom G, SlabArrays A, const uint8_t* __restrict__ flags,
                                                        const uint32_t* __restrict__ cta_off, uint32_t ncta,
                                                        double* __restrict__ send_left, double* __restrict__ send_right,
                                                        uint32_t* __restrict__ vacant, uint32_t* __restrict__ scnt,
                                                        DevCounters* __restrict__ cnt) {
    __shared__ uint32_t wtot[3][8];
    if (cnt->overflow)
        return;
    const uint32_t nL = cta_off[3 * ncta], nR = cta_off[3 * ncta + 1], nV = cta_off[3 * ncta + 2];
    if (nL > G.ghost_cap || nR > G.ghost_cap) {
        if (blockIdx.x == 0 && threadIdx.x == 0)
            atomicOr(&cnt->overflow, (uint32_t)OVF_SLAB);
        return;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (send_left)
            send_left[0] = (double)nL, send_left[1] = 0;
        if (send_right)
            send_right[0] = (double)nR, send_right[1] = 0;
        scnt[1] = nV;
    }
    // nothing flagged among this CTA's 256 slots (all but the CTAs of the two ghost bands and of vacant slots): done.
    // (The totals sit behind the last triple, so the difference works for the last CTA too.)
    const uint32_t* co = cta_off + 3 * (size_t)blockIdx.x;
    if (co[3] == co[0] && co[4] == co[1] && co[5] == co[2])
        return;
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t f = s < G.owned_cap ? flags[s] : 0u;
    uint32_t rank[3];
    slab_ranks(f, cta_off, wtot, rank);
    const uint32_t b = A.nb_static + s;
    {
        // the records of the warp's flagged slots, one after the other, each by all 32 lanes
        const int lane = threadIdx.x & 31;
        for (int side = 0; side < 2; side++) {
            double* buf = side ? send_right : send_left;
            for (unsigned m = __ballot_sync(0xffffffffu, f & (side ? SLAB_F_RIGHT : SLAB_F_LEFT)); m; m &= m - 1u) {
                const int src = __ffs((int)m) - 1;
                const uint32_t bs = __shfl_sync(0xffffffffu, b, src);
                const uint32_t rk = __shfl_sync(0xffffffffu, rank[side], src);
                slab_write_record_warp(A, bs, buf + SLAB_HDR + (size_t)SLAB_REC * rk, lane);
            }
        }
    }
    if (s >= G.owned_cap)
        return;
    if (f & SLAB_F_VACANT) {
        if (A.bflags[b] != 4) {
            A.bflags[b] = 4;
            slab_write_meta(A, b, 4);
        }
        vacant[rank[2]] = b;
    } else if (f & (SLAB_F_LEFT | SLAB_F_RIGHT)) {
        // A body past the plane now belongs to the neighbour (it is in the buffer just written; being past the plane it
        // is inside the band, hence flagged).  Its slot stays one more step as a ghost -- the neighbour's own ghost
        // layer will only list it from the next exchange on -- and is released by the next pack (bit 3).
        const double p = A.pos[3 * (size_t)b + G.axis];
        if ((G.has_left && p < G.lo) || (G.has_right && p >= G.hi)) {
            A.bflags[b] = 2 | 8;
            slab_write_meta(A, b, 2 | 8);
        }
    }
}

// Rebuild one ghost region from a received buffer; arrivals take vacant owned slots in record order.  One CTA per
// 1024 records: a CTA first counts the arrivals among the records before its chunk (one coordinate each, a few KB),
// which is the rank offset a single sequential CTA would have carried along, then handles its own chunk.
__device__ __forceinline__ bool slab_is_arrival(const SlabGeom& G, const double* __restrict__ recv, uint32_t j) {
    const double p = recv[SLAB_HDR + (size_t)SLAB_REC * j + 1 + G.axis];
    return (!G.has_left || p >= G.lo) && (!G.has_right || p < G.hi);
}

// blockIdx.y = side (0: the left neighbour's buffer, unpacked "first": its arrivals take the first vacant slots).
__global__ void __launch_bounds__(512) k_slab_unpack(SlabGeom G, SlabArrays A, const double* __restrict__ recv_left,
                                                      const double* __restrict__ recv_right, uint32_t ghost_base_left,
                                                      const uint32_t* __restrict__ vacant, uint32_t* __restrict__ scnt,
                                                      DevCounters* __restrict__ cnt) {
    __shared__ uint32_t warp_tot[32];
    if (cnt->overflow)
        return;
    const int first_side = blockIdx.y == 0;
    const double* __restrict__ recv = first_side ? recv_left : recv_right;
    const uint32_t ghost_base = first_side ? ghost_base_left : ghost_base_left + G.ghost_cap;
    const uint32_t n = recv ? (uint32_t)recv[0] : 0u;
    const uint32_t n_vac = scnt[1];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t j0 = blockIdx.x * blockDim.x;
    // arrivals before this chunk (CTA 0 counts all of them instead, for the other side's kernel)
    const uint32_t upto = blockIdx.x == 0 ? n : min(j0, n);
    uint32_t mine = 0;
    for (uint32_t i = threadIdx.x; i < ((upto + 31u) & ~31u); i += blockDim.x) {
        const bool a = i < upto && slab_is_arrival(G, recv, i);
        mine += (uint32_t)__popc(__ballot_sync(0xffffffffu, a));  // warp-uniform
    }
    if (lane == 0)
        warp_tot[wid] = mine;
    __syncthreads();
    uint32_t counted = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++)
        counted += warp_tot[w];
    __syncthreads();
    // arrivals of the side unpacked first: the other side's CTAs count them themselves (one launch for both sides)
    uint32_t side_base = 0;
    if (!first_side && recv_left) {
        const uint32_t nl = (uint32_t)recv_left[0];
        uint32_t m2 = 0;
        for (uint32_t i = threadIdx.x; i < ((nl + 31u) & ~31u); i += blockDim.x) {
            const bool a = i < nl && slab_is_arrival(G, recv_left, i);
            m2 += (uint32_t)__popc(__ballot_sync(0xffffffffu, a));
        }
        if (lane == 0)
            warp_tot[wid] = m2;
        __syncthreads();
        for (int w = 0; w < (int)(blockDim.x >> 5); w++)
            side_base += warp_tot[w];
        __syncthreads();
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
        scnt[first_side ? 2 : 3] = first_side ? counted : side_base + counted;
    const uint32_t base = side_base + (blockIdx.x == 0 ? 0u : counted);

    const uint32_t j = j0 + threadIdx.x;
    bool arrival = false;
    const double* r = nullptr;
    if (j < n) {
        r = recv + SLAB_HDR + (size_t)SLAB_REC * j;
        arrival = slab_is_arrival(G, recv, j);
    }
    const unsigned m = __ballot_sync(0xffffffffu, arrival);
    if (lane == 0)
        warp_tot[wid] = (uint32_t)__popc(m);
    __syncthreads();
    uint32_t before = base;
    for (int w = 0; w < wid; w++)
        before += warp_tot[w];
    const uint32_t rank = before + (uint32_t)__popc(m & ((1u << lane) - 1));
    if (j < G.ghost_cap) {
        const uint32_t gslot = ghost_base + j;
        if (j < n && !arrival) {
            slab_read_record(A, gslot, r, 2);  // ghost: collides, not integrated
        } else {
            if (A.bflags[gslot] != 4) {
                A.bflags[gslot] = 4;
                slab_write_meta(A, gslot, 4);
            }
            if (arrival) {
                if (rank < n_vac)
                    slab_read_record(A, vacant[rank], r, 3);
                else
                    atomicOr(&cnt->overflow, (uint32_t)OVF_SLAB);
            }
        }
    }
}

static SlabArrays slab_arrays(System* S) {
    SlabArrays A;
    A.pos = S->pos.p, A.rot = S->rot.p, A.vel = S->vel.p, A.mass = S->mass.p, A.inertia = S->inertia.p;
    A.inv_inertia = S->inv_inertia.p, A.gid = S->b_gid.p, A.bflags = S->bflags.p;
    A.s_dims = S->s_dims.p, A.s_lpos = S->s_lpos.p, A.s_lrot = S->s_lrot.p;
    A.s_type = S->s_type.p, A.s_mat = S->s_mat.p, A.s_fam = S->s_fam.p, A.meta = S->s_meta.p;
    A.nb_static = S->slab_nb_static, A.ns_static = S->slab_ns_static;
    return A;
}
static SlabGeom slab_geom(const System* S) {
    SlabGeom G;
    G.axis = S->slab_axis, G.has_left = S->slab_has_left, G.has_right = S->slab_has_right;
    G.lo = S->slab_lo, G.hi = S->slab_hi, G.w = S->slab_w;
    G.owned_cap = S->slab_owned_cap, G.ghost_cap = S->slab_ghost_cap;
    return G;
}

}  // namespace cb200

using namespace cb200;
static System* SYSL(cb200_system* s) { return reinterpret_cast<System*>(s); }

extern "C" {

// Global ids of bodies [first, first+n) (what ChBody::GetIdentifier / the body's index in the undecomposed system is).
int cb200_upload_body_ids(cb200_system* sys, int first, int n, const int64_t* ids) {
    System* S = SYSL(sys);
    if (!S || first < 0 || n < 0 || (uint32_t)(first + n) > S->nb || !ids)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_upload_body_ids");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(S->b_gid.reserve(S->pos.cap / 3 + 1, true, S->stream));
    CB_CUDA(cudaMemcpyAsync(S->b_gid.p + first, ids, (size_t)n * 8, cudaMemcpyHostToDevice, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}

// ids (global id, -1 for bodies that never got one) and flags (bit0 active = owned, bit1 collide, bit2 vacant)
int cb200_download_body_ids(cb200_system* sys, int first, int n, int64_t* ids, uint8_t* flags) {
    System* S = SYSL(sys);
    if (!S || first < 0 || n < 0 || (uint32_t)(first + n) > S->nb)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_download_body_ids");
    CB_CUDA(cudaSetDevice(S->device));
    if (ids) {
        if (S->b_gid.cap < S->nb)
            return set_error(CB200_ERR_STATE, "no body ids were uploaded");
        CB_CUDA(cudaMemcpyAsync(ids, S->b_gid.p + first, (size_t)n * 8, cudaMemcpyDeviceToHost, S->stream));
    }
    if (flags)
        CB_CUDA(cudaMemcpyAsync(flags, S->bflags.p + first, (size_t)n, cudaMemcpyDeviceToHost, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}

// Turn the system into one slab of a decomposed system.  Bodies [nb_static, nb) added so far are this rank's
// owned bodies, each with exactly one shape, in order (shape ns_static + k belongs to body nb_static + k).  The call
// appends vacant slots up to owned_cap and the two ghost regions.  lo/hi: this rank's interval along axis (ignored
// on a side without neighbour); width: ghost-layer width, >= 2 r_max + 2 envelope.
int cb200_slab_configure(cb200_system* sys, int nb_static, int ns_static, int axis, double lo, double hi,
                         double width, int has_left, int has_right, int owned_cap, int ghost_cap) {
    System* S = SYSL(sys);
    if (S)
        invalidate_step_graph(S);
    if (!S || nb_static < 0 || (uint32_t)nb_static > S->nb || axis < 0 || axis > 2 || !(width > 0) || ghost_cap <= 0)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_slab_configure");
    if (S->slab_on)
        return set_error(CB200_ERR_STATE, "the system is already configured as a slab");
    if (S->inertia_off.p)
        return set_error(CB200_ERR_UNSUPPORTED, "products of inertia do not travel with the slab records yet");
    if (ghost_cap >= (1 << 21) || owned_cap >= (1 << 22))
        return set_error(CB200_ERR_UNSUPPORTED, "slab capacities above 2^21 ghost / 2^22 owned slots");
    CB_CUDA(cudaSetDevice(S->device));
    const uint32_t n_dyn = S->nb - (uint32_t)nb_static;
    if ((uint32_t)owned_cap < n_dyn)
        return set_error(CB200_ERR_ARG, "owned_cap is smaller than the number of dynamic bodies already added");
    if (S->ns - (uint32_t)ns_static != n_dyn)
        return set_error(CB200_ERR_ARG, "slab mode needs exactly one shape per dynamic body");
    {
        std::vector<uint32_t> sb(n_dyn);
        CB_CUDA(cudaMemcpy(sb.data(), S->s_body.p + ns_static, (size_t)n_dyn * 4, cudaMemcpyDeviceToHost));
        for (uint32_t k = 0; k < n_dyn; k++)
            if (sb[k] != (uint32_t)nb_static + k)
                return set_error(CB200_ERR_ARG, "slab mode needs shape ns_static + k to belong to body nb_static + k");
    }
    const int extra = owned_cap - (int)n_dyn + 2 * ghost_cap;
    {
        std::vector<double> mass(extra, 1.0), inr(3 * (size_t)extra, 1.0), lp(3 * (size_t)extra, 0.0),
            lr(4 * (size_t)extra, 0.0), dims(4 * (size_t)extra, 0.0);
        std::vector<char> zero(extra, 0);
        std::vector<int32_t> type(extra, CB200_SPHERE), mat(extra, 0);
        std::vector<uint32_t> body(extra);
        std::vector<int16_t> fam(2 * (size_t)extra);
        int first = 0;
        CB_TRY(cb200_append_bodies(sys, extra, mass.data(), inr.data(), nullptr, zero.data(), zero.data(), &first));
        for (int k = 0; k < extra; k++) {
            body[k] = (uint32_t)(first + k);
            lr[4 * (size_t)k] = 1.0;
            fam[2 * k] = 1, fam[2 * k + 1] = 0x7fff;
        }
        CB_TRY(cb200_append_shapes(sys, extra, type.data(), body.data(), fam.data(), lp.data(), lr.data(), dims.data(),
                                   mat.data()));
        std::vector<uint8_t> vac(extra, 4);
        CB_CUDA(cudaMemcpy(S->bflags.p + first, vac.data(), (size_t)extra, cudaMemcpyHostToDevice));
    }
    if (S->b_gid.cap < S->nb) {
        const size_t old = S->b_gid.cap;
        CB_TRY(S->b_gid.reserve(S->nb, true, S->stream));
        if (old == 0)
            CB_CUDA(cudaMemset(S->b_gid.p, 0xff, (size_t)S->nb * 8));
    }
    CB_TRY(S->slab_flags.reserve((size_t)owned_cap / 8 + 4));  // one byte per owned slot
    CB_TRY(S->slab_cta.reserve(3 * ((size_t)owned_cap / SLAB_CTA + 2)));
    CB_TRY(S->slab_vacant.reserve((size_t)owned_cap + 1));
    CB_CUDA(cudaMalloc(&S->slab_cnt, 16));
    const uint32_t init[4] = {(uint32_t)owned_cap, 0, 0, 0};
    CB_CUDA(cudaMemcpy(S->slab_cnt, init, 16, cudaMemcpyHostToDevice));
    S->slab_on = true;
    S->slab_axis = axis, S->slab_lo = lo, S->slab_hi = hi, S->slab_w = width;
    S->slab_has_left = has_left ? 1 : 0, S->slab_has_right = has_right ? 1 : 0;
    S->slab_nb_static = (uint32_t)nb_static, S->slab_ns_static = (uint32_t)ns_static;
    S->slab_owned_cap = (uint32_t)owned_cap, S->slab_ghost_cap = (uint32_t)ghost_cap;
    S->meta_dirty = true;
    return CB200_OK;
}

// doubles a send / receive buffer must hold
int64_t cb200_slab_buffer_doubles(cb200_system* sys) {
    System* S = SYSL(sys);
    return S && S->slab_on ? (int64_t)SLAB_HDR + (int64_t)SLAB_REC * S->slab_ghost_cap : 0;
}

// Enqueue the pack pass on the system's stream.  Buffers are DEVICE pointers (NULL on a side without neighbour).
int cb200_slab_pack_dev(cb200_system* sys, void* send_left_dev, void* send_right_dev) {
    System* S = SYSL(sys);
    if (!S || !S->slab_on)
        return set_error(CB200_ERR_STATE, "cb200_slab_configure was not called");
    if ((S->slab_has_left && !send_left_dev) || (S->slab_has_right && !send_right_dev))
        return set_error(CB200_ERR_ARG, "missing send buffer for a side that has a neighbour");
    CB_CUDA(cudaSetDevice(S->device));
    const SlabGeom G = slab_geom(S);
    const SlabArrays A = slab_arrays(S);
    if (S->meta_dirty)  // the kernels below keep the shape records current slot by slot: they must exist first
        CB_TRY(enqueue_build_meta(S));
    const uint32_t ncta = (G.owned_cap + SLAB_CTA - 1) / SLAB_CTA;
    uint8_t* flags = reinterpret_cast<uint8_t*>(S->slab_flags.p);
    k_slab_flags<<<ncta, SLAB_CTA, 0, S->stream>>>(G, A, flags, S->slab_cta.p);
    k_slab_cta_scan<<<1, 1024, 0, S->stream>>>(ncta, S->slab_cta.p);
    k_slab_pack<<<ncta, SLAB_CTA, 0, S->stream>>>(G, A, flags, S->slab_cta.p, ncta, (double*)send_left_dev,
                                                  (double*)send_right_dev, S->slab_vacant.p, S->slab_cnt, S->d_cnt);
    S->launches += 3;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

// Enqueue the unpack pass: ghost regions rebuilt from the two received DEVICE buffers, arrivals adopted.
int cb200_slab_unpack_dev(cb200_system* sys, const void* recv_left_dev, const void* recv_right_dev) {
    System* S = SYSL(sys);
    if (!S || !S->slab_on)
        return set_error(CB200_ERR_STATE, "cb200_slab_configure was not called");
    CB_CUDA(cudaSetDevice(S->device));
    const SlabGeom G = slab_geom(S);
    const SlabArrays A = slab_arrays(S);
    const uint32_t gl = S->slab_nb_static + G.owned_cap, gr = gl + G.ghost_cap;
    const uint32_t ug = (G.ghost_cap + 511u) / 512u;  // (512 threads: room for a whole record in registers)
    (void)gr;
    k_slab_unpack<<<dim3(ug, 2), 512, 0, S->stream>>>(G, A, S->slab_has_left ? (const double*)recv_left_dev : nullptr,
                                                       S->slab_has_right ? (const double*)recv_right_dev : nullptr, gl,
                                                       S->slab_vacant.p, S->slab_cnt, S->d_cnt);
    S->launches++;
    CB_CUDA(cudaGetLastError());
    S->sphere_aabb_valid = false;  // ghost and adopted bodies changed behind the integration's back
    return CB200_OK;
}

}  // extern "C"
