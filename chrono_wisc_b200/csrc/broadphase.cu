// broadphase.cu -- AABB generation and uniform-grid broadphase on the device.
//
// Reference path replaced (CMC = src/chrono/collision/multicore):
//   GenerateAABB                      CMC/ChCollisionSystemMulticore.cpp:458-545   -> k_aabb (+ fused bbox reduce, C1+C2,
//                                                                                   + PreprocessLocalToParent C14)
//   ChBroadphase::DetermineBoundingBox/ComputeTopLevelResolution  CMC/ChBroadphase.cpp:142-208 -> host_grid()
//   OffsetAABB + f_Count_AABB_BIN     CMC/ChBroadphase.cpp:168-176,261-268         -> k_bin_count
//   exclusive_scan / sort_by_key / RLE CMC/ChBroadphase.cpp:271-301                -> counting sort: histogram with
//                                       warp-aggregated atomics, packed u64 scan, scatter, per-bin rank sort
//   f_Count/Store_AABB_AABB_Intersection  CMC/ChCollisionUtilsBroadphase.cpp:86-217 -> k_bin_pairs<false/true>
//
// Design notes: instead of sorting N_i (bin, shape) keys we histogram into a dense bin array (atomics resolve in
// L2), scan it once (entry offsets in the low word, active-bin index in the high word), scatter, and let the warp
// that owns a bin order its few entries by shape id.  That reproduces exactly the (bin, shape) order of the
// reference's sort_by_key, so pair emission order -- and with it the contact order -- equals the reference's.
#include "common.cuh"
#include <float.h>
#include <limits.h>
#include <math.h>

namespace cb200 {

// --------------------------------------------------------------------------------------------------------------
// k_aabb: one thread per shape.
// --------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_aabb(uint32_t ns, const int32_t* __restrict__ s_type,
                                              const uint32_t* __restrict__ s_body, const uint32_t* __restrict__ s_fam,
                                              const int32_t* __restrict__ s_mat, const double* __restrict__ s_lpos,
                                              const double* __restrict__ s_lrot, const double* __restrict__ s_dims,
                                              const double* __restrict__ pos, const double* __restrict__ rot,
                                              const uint8_t* __restrict__ bflags, double envelope,
                                              double* __restrict__ s_wpos, double* __restrict__ s_wrot,
                                              ShapeBP* __restrict__ s_bp, DevCounters* __restrict__ cnt) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    d3 mn = mk3(DBL_MAX), mx = mk3(-DBL_MAX);
    bool counted = false;
    if (i < ns) {
        const uint32_t body = s_body[i];
        ShapeBP bp;
        bp.body = body;
        bp.fam = s_fam[i];
        bp.pad = 0;
        const int type = s_type[i];
        bp.info = ((uint32_t)type & 0xff) << 8 | ((uint32_t)s_mat[i] & 0xffff) << 16;
        d3 amin = mk3(0), amax = mk3(0);
        if (body != UINT_MAX) {
            const uint8_t bf = bflags[body];
            bp.info |= bf & 3u;
            const d3 p = mk3(pos[3 * body], pos[3 * body + 1], pos[3 * body + 2]);
            const double4 qa = reinterpret_cast<const double4*>(rot)[body];
            const q4 q = mkq(qa.x, qa.y, qa.z, qa.w);
            const d3 lp = mk3(s_lpos[3 * i], s_lpos[3 * i + 1], s_lpos[3 * i + 2]);
            const double4 la = reinterpret_cast<const double4*>(s_lrot)[i];
            const q4 lq = mkq(la.x, la.y, la.z, la.w);
            const double4 dm = reinterpret_cast<const double4*>(s_dims)[i];

            const d3 wp = rotate(lp, q) + p;  // == TransformLocalToParent(p, q, lp): IEEE addition commutes
            const q4 wq = qmul(q, lq);
            reinterpret_cast<double4*>(s_wpos)[i] = make_double4(wp.x, wp.y, wp.z, dm.x);
            reinterpret_cast<double4*>(s_wrot)[i] = make_double4(wq.w, wq.x, wq.y, wq.z);

            bool ok = true;
            if (type == CB200_SPHERE) {
                const double r = dm.x + envelope;
                amin = wp - r;
                amax = wp + r;
            } else if (type == CB200_ELLIPSOID || type == CB200_BOX || type == CB200_CYLINDER ||
                       type == CB200_CYLSHELL || type == CB200_CONE) {
                const d3 t = abs_rotate(wq, mk3(dm.x, dm.y, dm.z) + envelope);
                amin = wp - t;
                amax = wp + t;
            } else if (type == CB200_ROUNDEDBOX || type == CB200_ROUNDEDCYL) {
                const d3 t = abs_rotate(wq, mk3(dm.x, dm.y, dm.z) + dm.w + envelope);
                amin = wp - t;
                amax = wp + t;
            } else if (type == CB200_CAPSULE) {
                const d3 t = abs_rotate(wq, mk3(dm.x, dm.x + dm.y, dm.x) + envelope);
                amin = wp - t;
                amax = wp + t;
            } else {
                ok = false;  // type without an AABB rule: the reference `continue`s and keeps stale values
            }
            if (ok && (bf & 2u)) {
                mn = amin;
                mx = amax;
                counted = true;
            }
        }
        bp.mn[0] = amin.x, bp.mn[1] = amin.y, bp.mn[2] = amin.z;
        bp.mx[0] = amax.x, bp.mx[1] = amax.y, bp.mx[2] = amax.z;
        s_bp[i] = bp;
    }
    // union of the AABBs of colliding shapes (BoxInvert/BoxReduce, CMC/ChBroadphase.cpp:38-100): min/max are exact
    // and order-free, so warp shuffles + ordered-integer atomics give the reference's result bit for bit.
    const unsigned any = __ballot_sync(0xffffffffu, counted);
    if (any) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            mn.x = fmin(mn.x, __shfl_xor_sync(0xffffffffu, mn.x, d));
            mn.y = fmin(mn.y, __shfl_xor_sync(0xffffffffu, mn.y, d));
            mn.z = fmin(mn.z, __shfl_xor_sync(0xffffffffu, mn.z, d));
            mx.x = fmax(mx.x, __shfl_xor_sync(0xffffffffu, mx.x, d));
            mx.y = fmax(mx.y, __shfl_xor_sync(0xffffffffu, mx.y, d));
            mx.z = fmax(mx.z, __shfl_xor_sync(0xffffffffu, mx.z, d));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(&cnt->bbox[0], enc_ordered(mn.x));
            atomicMin(&cnt->bbox[1], enc_ordered(mn.y));
            atomicMin(&cnt->bbox[2], enc_ordered(mn.z));
            atomicMax(&cnt->bbox[3], enc_ordered(mx.x));
            atomicMax(&cnt->bbox[4], enc_ordered(mx.y));
            atomicMax(&cnt->bbox[5], enc_ordered(mx.z));
        }
    }
}

__global__ void k_reset_counters(DevCounters* cnt) {
    if (threadIdx.x == 0) {
        cnt->bbox[0] = cnt->bbox[1] = cnt->bbox[2] = ~0ull;
        cnt->bbox[3] = cnt->bbox[4] = cnt->bbox[5] = 0ull;
        cnt->n_large = 0;
        cnt->n_intersections = cnt->n_active_bins = cnt->n_pairs = cnt->n_slots = cnt->n_contacts = 0;
        cnt->multi_slot = 0;
        cnt->n_heavy = 0;
        cnt->ct_body_count = 0;
        cnt->overflow = 0;
        cnt->unsupported = 0;
    }
}

struct GridParams {
    double origin[3];
    double inv[3];
    int bins[3];
    uint32_t num_bins;
};

__device__ __forceinline__ bool bin_range(const ShapeBP& bp, const GridParams& g, int* lo, int* hi) {
    // HashMin / HashMax, CMC/ChCollisionUtils.h:44-60
    lo[0] = (int)floor(bp.mn[0] * g.inv[0]);
    lo[1] = (int)floor(bp.mn[1] * g.inv[1]);
    lo[2] = (int)floor(bp.mn[2] * g.inv[2]);
    hi[0] = (int)ceil(bp.mx[0] * g.inv[0]) - 1;
    hi[1] = (int)ceil(bp.mx[1] * g.inv[1]) - 1;
    hi[2] = (int)ceil(bp.mx[2] * g.inv[2]) - 1;
    // A shape outside the grid (only possible for shapes of non-colliding bodies, which are excluded from the
    // global box) would index outside the dense histogram; it can never form a pair, so it is not binned.
    return lo[0] >= 0 && lo[1] >= 0 && lo[2] >= 0 && hi[0] < g.bins[0] && hi[1] < g.bins[1] && hi[2] < g.bins[2] &&
           hi[0] >= lo[0] && hi[1] >= lo[1] && hi[2] >= lo[2];
}

// Warp-aggregated increment: lanes that target the same counter elect a leader that adds the group size once;
// every lane gets its own slot.  All 32 lanes must call (valid=false lanes pass through).
__device__ __forceinline__ uint32_t warp_agg_inc(uint32_t* counters, uint32_t key, bool valid) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned grp = __match_any_sync(0xffffffffu, valid ? key : 0xffffffffu);
    const int leader = __ffs(grp) - 1;
    const unsigned rank = __popc(grp & ((1u << lane) - 1));
    uint32_t base = 0;
    if (valid && (int)lane == leader)
        base = atomicAdd(&counters[key], (uint32_t)__popc(grp));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + rank;
}

// --------------------------------------------------------------------------------------------------------------
// k_bin_count: OffsetAABB (in place) + bins touched per shape + dense histogram.
// --------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bin_count(uint32_t ns, GridParams g, ShapeBP* __restrict__ s_bp,
                                                   uint32_t* __restrict__ bin_intersections,
                                                   uint32_t* __restrict__ bin_count, uint32_t* __restrict__ large_list,
                                                   uint8_t* __restrict__ heavy, DevCounters* __restrict__ cnt) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    int lo[3] = {0, 0, 0}, hi[3] = {-1, -1, -1};
    uint32_t n = 0;
    if (i < ns) {
        ShapeBP bp = s_bp[i];
        if (bp.body != UINT_MAX) {
            bp.mn[0] -= g.origin[0], bp.mn[1] -= g.origin[1], bp.mn[2] -= g.origin[2];
            bp.mx[0] -= g.origin[0], bp.mx[1] -= g.origin[1], bp.mx[2] -= g.origin[2];
            s_bp[i] = bp;
            if (bin_range(bp, g, lo, hi))
                n = (uint32_t)(hi[0] - lo[0] + 1) * (uint32_t)(hi[1] - lo[1] + 1) * (uint32_t)(hi[2] - lo[2] + 1);
        }
        bin_intersections[i] = n;
        if (n > LARGE_BINS) {
            const uint32_t k = atomicAdd(&cnt->n_large, 1u);
            if (k < MAX_LARGE)
                large_list[k] = i;
            else
                cnt->overflow = 1;
            heavy[bp.body] = 1;
        }
    }
    const uint32_t mine = (n <= LARGE_BINS) ? n : 0;
    uint32_t wmax = mine;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
        wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, d));
    const int ny = hi[1] - lo[1] + 1, nz = hi[2] - lo[2] + 1;
    for (uint32_t it = 0; it < wmax; it++) {
        const bool valid = it < mine;
        uint32_t b = 0;
        if (valid) {
            const int ix = lo[0] + (int)(it / (uint32_t)(ny * nz));
            const int rem = (int)(it % (uint32_t)(ny * nz));
            const int iy = lo[1] + rem / nz, iz = lo[2] + rem % nz;
            b = (uint32_t)((iz * g.bins[1]) * g.bins[0] + iy * g.bins[0] + ix);  // Hash_Index
        }
        warp_agg_inc(bin_count, b, valid);
    }
}

// Shapes spanning many bins (container walls): one CTA per shape, threads stride over its bins.
template <bool FILL>
__global__ void __launch_bounds__(256) k_bin_large(GridParams g, const ShapeBP* __restrict__ s_bp,
                                                   const uint32_t* __restrict__ large_list,
                                                   const DevCounters* __restrict__ cnt,
                                                   uint32_t* __restrict__ bin_count /* or cursor */,
                                                   const uint64_t* __restrict__ bin_start,
                                                   uint32_t* __restrict__ entries) {
    const uint32_t nl = min(cnt->n_large, (uint32_t)MAX_LARGE);
    for (uint32_t s = blockIdx.x; s < nl; s += gridDim.x) {
        const uint32_t shape = large_list[s];
        const ShapeBP bp = s_bp[shape];
        int lo[3], hi[3];
        if (!bin_range(bp, g, lo, hi))
            continue;
        const int ny = hi[1] - lo[1] + 1, nz = hi[2] - lo[2] + 1;
        const uint32_t n = (uint32_t)(hi[0] - lo[0] + 1) * ny * nz;
        for (uint32_t it = threadIdx.x; it < n; it += blockDim.x) {
            const int ix = lo[0] + (int)(it / (uint32_t)(ny * nz));
            const int rem = (int)(it % (uint32_t)(ny * nz));
            const int iy = lo[1] + rem / nz, iz = lo[2] + rem % nz;
            const uint32_t b = (uint32_t)((iz * g.bins[1]) * g.bins[0] + iy * g.bins[0] + ix);
            const uint32_t slot = atomicAdd(&bin_count[b], 1u);
            if (FILL)
                entries[(uint32_t)bin_start[b] + slot] = shape;
        }
    }
}

__global__ void __launch_bounds__(256) k_bin_pack(uint32_t num_bins, const uint32_t* __restrict__ c,
                                                  uint64_t* __restrict__ packed) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < num_bins) {
        const uint32_t v = c[i];
        packed[i] = (uint64_t)v | ((uint64_t)(v != 0) << 32);
    }
}

__global__ void __launch_bounds__(256) k_bin_fill(uint32_t ns, GridParams g, const ShapeBP* __restrict__ s_bp,
                                                  const uint32_t* __restrict__ bin_intersections,
                                                  const uint64_t* __restrict__ bin_start,
                                                  uint32_t* __restrict__ cursor, uint32_t* __restrict__ entries) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    int lo[3] = {0, 0, 0}, hi[3] = {-1, -1, -1};
    uint32_t mine = 0;
    if (i < ns) {
        const uint32_t n = bin_intersections[i];
        if (n != 0 && n <= LARGE_BINS) {
            const ShapeBP bp = s_bp[i];
            bin_range(bp, g, lo, hi);
            mine = n;
        }
    }
    uint32_t wmax = mine;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
        wmax = max(wmax, __shfl_xor_sync(0xffffffffu, wmax, d));
    const int ny = hi[1] - lo[1] + 1, nz = hi[2] - lo[2] + 1;
    for (uint32_t it = 0; it < wmax; it++) {
        const bool valid = it < mine;
        uint32_t b = 0;
        if (valid) {
            const int ix = lo[0] + (int)(it / (uint32_t)(ny * nz));
            const int rem = (int)(it % (uint32_t)(ny * nz));
            const int iy = lo[1] + rem / nz, iz = lo[2] + rem % nz;
            b = (uint32_t)((iz * g.bins[1]) * g.bins[0] + iy * g.bins[0] + ix);
        }
        const uint32_t slot = warp_agg_inc(cursor, b, valid);
        if (valid)
            entries[(uint32_t)bin_start[b] + slot] = i;
    }
}

// --------------------------------------------------------------------------------------------------------------
// k_bin_pairs: one warp per bin.  COUNT pass also orders the bin's entries by shape id (what the reference gets
// from its stable sort_by_key) so both passes walk pairs in the reference's (i, k>i) order.
// --------------------------------------------------------------------------------------------------------------
struct BPRec {
    double mn[3], mx[3];
    uint32_t body, fam, info;
};

__device__ __forceinline__ BPRec load_bp(const ShapeBP* __restrict__ s_bp, uint32_t shape) {
    const double2* p = reinterpret_cast<const double2*>(s_bp + shape);
    const double2 a = p[0], b = p[1], c = p[2];
    const uint4 u = reinterpret_cast<const uint4*>(s_bp + shape)[3];
    BPRec r;
    r.mn[0] = a.x, r.mn[1] = a.y, r.mn[2] = b.x;
    r.mx[0] = b.y, r.mx[1] = c.x, r.mx[2] = c.y;
    r.body = u.x, r.fam = u.y, r.info = u.z;
    return r;
}

__device__ __forceinline__ bool pair_passes(const BPRec& A, uint32_t sA, const BPRec& B, uint32_t sB,
                                            const GridParams& g, uint32_t bin) {
    // filter order of CMC/ChCollisionUtilsBroadphase.cpp:114-140
    if (B.body == UINT_MAX || sA == sB || A.body == B.body)
        return false;
    if (!(B.info & 2u))
        return false;
    if (!(A.info & 1u) && !(B.info & 1u))
        return false;
    const uint32_t gA = A.fam & 0xffffu, mA = A.fam >> 16, gB = B.fam & 0xffffu, mB = B.fam >> 16;
    if (!((mA & gB) && (mB & gA)))
        return false;
    if (!((A.mn[0] <= B.mx[0] && B.mn[0] <= A.mx[0]) && (A.mn[1] <= B.mx[1] && B.mn[1] <= A.mx[1]) &&
          (A.mn[2] <= B.mx[2] && B.mn[2] <= A.mx[2])))
        return false;
    // current_bin (CMC/ChCollisionUtils.h:97-110): the pair belongs to the bin holding max(Amin, Bmin)
    const int cx = (int)floor(fmax(A.mn[0], B.mn[0]) * g.inv[0]);
    const int cy = (int)floor(fmax(A.mn[1], B.mn[1]) * g.inv[1]);
    const int cz = (int)floor(fmax(A.mn[2], B.mn[2]) * g.inv[2]);
    return (uint32_t)((cz * g.bins[1]) * g.bins[0] + cy * g.bins[0] + cx) == bin;
}

template <bool STORE>
__global__ void __launch_bounds__(256) k_bin_pairs(GridParams g, const ShapeBP* __restrict__ s_bp,
                                                   const uint64_t* __restrict__ bin_start,
                                                   const uint32_t* __restrict__ entries_tmp,
                                                   uint32_t* __restrict__ entries, uint32_t* __restrict__ pair_count,
                                                   int64_t* __restrict__ pairs, PairAux* __restrict__ aux,
                                                   DevCounters* __restrict__ cnt) {
    const uint32_t bin = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    if (bin >= g.num_bins)
        return;
    const uint32_t start = (uint32_t)bin_start[bin];
    const uint32_t k = (uint32_t)bin_start[bin + 1] - start;
    if (k == 0) {
        if (!STORE && lane == 0)
            pair_count[bin] = 0;
        return;
    }
    if (!STORE) {
        // rank sort by shape id (ids are unique inside a bin)
        for (uint32_t e = lane; e < k; e += 32) {
            const uint32_t id = entries_tmp[start + e];
            uint32_t rank = 0;
            for (uint32_t j = 0; j < k; j++)
                rank += entries_tmp[start + j] < id;
            entries[start + rank] = id;
        }
        __syncwarp();
        if (k == 1) {
            if (lane == 0)
                pair_count[bin] = 0;
            return;
        }
    } else if (k == 1) {
        return;
    }

    uint32_t out = STORE ? pair_count[bin] : 0;  // exclusive offset in STORE, running count otherwise
    const unsigned lt = (1u << lane) - 1;
    bool multi = false;

    if (k <= 32) {
        // fast path: lane j keeps entry j in registers, entry i arrives by a uniform (broadcast) load
        uint32_t sB = 0;
        BPRec B;
        if (lane < k) {
            sB = entries[start + lane];
            B = load_bp(s_bp, sB);
        }
        for (uint32_t i = 0; i + 1 < k; i++) {
            const uint32_t sA = entries[start + i];
            const BPRec A = load_bp(s_bp, sA);
            if (A.body == UINT_MAX || !(A.info & 2u))
                continue;
            const bool hit = (lane > i && lane < k) && pair_passes(A, sA, B, sB, g, bin);
            const unsigned bal = __ballot_sync(0xffffffffu, hit);
            if (STORE && hit) {
                const uint32_t o = out + __popc(bal & lt);
                pairs[o] = ((int64_t)sA << 32) | (int64_t)sB;
                PairAux a;
                a.bodyA = A.body, a.bodyB = B.body;
                const uint32_t tA = (A.info >> 8) & 0xff, tB = (B.info >> 8) & 0xff;
                a.types = tA | tB << 8;
                a.mats = (A.info >> 16) | (B.info >> 16) << 16;
                aux[o] = a;
                multi |= (tA != CB200_SPHERE && tB != CB200_SPHERE);
            }
            out += __popc(bal);
        }
    } else {
        for (uint32_t i = 0; i + 1 < k; i++) {
            const uint32_t sA = entries[start + i];
            const BPRec A = load_bp(s_bp, sA);
            if (A.body == UINT_MAX || !(A.info & 2u))
                continue;
            for (uint32_t j0 = i + 1; j0 < k; j0 += 32) {
                const uint32_t j = j0 + lane;
                bool hit = false;
                uint32_t sB = 0;
                BPRec B;
                if (j < k) {
                    sB = entries[start + j];
                    B = load_bp(s_bp, sB);
                    hit = pair_passes(A, sA, B, sB, g, bin);
                }
                const unsigned bal = __ballot_sync(0xffffffffu, hit);
                if (STORE && hit) {
                    const uint32_t o = out + __popc(bal & lt);
                    pairs[o] = ((int64_t)sA << 32) | (int64_t)sB;
                    PairAux a;
                    a.bodyA = A.body, a.bodyB = B.body;
                    const uint32_t tA = (A.info >> 8) & 0xff, tB = (B.info >> 8) & 0xff;
                    a.types = tA | tB << 8;
                    a.mats = (A.info >> 16) | (B.info >> 16) << 16;
                    aux[o] = a;
                    multi |= (tA != CB200_SPHERE && tB != CB200_SPHERE);
                }
                out += __popc(bal);
            }
        }
    }
    if (!STORE) {
        if (lane == 0)
            pair_count[bin] = out;
    } else if (__any_sync(0xffffffffu, multi) && lane == 0) {
        cnt->multi_slot = 1;
    }
}

// --------------------------------------------------------------------------------------------------------------
// host side
// --------------------------------------------------------------------------------------------------------------

// DetermineBoundingBox + ComputeTopLevelResolution (CMC/ChBroadphase.cpp:142-208) and Compute_Grid_Resolution
// (CMC/ChCollisionUtilsBroadphase.cpp:36-46), evaluated on the host with the same libm the reference uses.
static void host_grid(System* S, const double* mn, const double* mx) {
    const double fraction = 1e-3;
    double minp[3], maxp[3], diag[3];
    for (int a = 0; a < 3; a++) {
        const double size = mx[a] - mn[a];
        minp[a] = mn[a] - fraction * size;
        maxp[a] = mx[a] + fraction * size;
        S->origin[a] = minp[a];
        S->maxpt[a] = maxp[a];
        diag[a] = fabs(maxp[a] - minp[a]);
    }
    const cb200_settings& st = S->set;
    if (st.broadphase_grid == 0) {
        for (int a = 0; a < 3; a++)
            S->bins[a] = st.bins_per_axis[a];
    } else if (st.broadphase_grid == 1) {
        for (int a = 0; a < 3; a++)
            S->bins[a] = (int)std::ceil(diag[a] / st.bin_size[a]);
    } else {
        const double V = diag[0] * diag[1] * diag[2];
        const double k = st.grid_density;
        const unsigned num_aabb = S->ns;
        for (int a = 0; a < 3; a++) {
            int v = int(diag[a] * std::pow(k * num_aabb / V, 1.0 / 3.0));
            S->bins[a] = v > 1 ? v : 1;
        }
    }
    for (int a = 0; a < 3; a++) {
        if (S->bins[a] < 1)
            S->bins[a] = 1;
        S->bin_size[a] = diag[a] / (double)S->bins[a];
        S->inv_bin_size[a] = 1.0 / S->bin_size[a];
    }
}

static GridParams grid_params(const System* S) {
    GridParams g;
    for (int a = 0; a < 3; a++) {
        g.origin[a] = S->origin[a];
        g.inv[a] = S->inv_bin_size[a];
        g.bins[a] = S->bins[a];
    }
    g.num_bins = S->num_bins;
    return g;
}

int stage_aabb(System* S) {
    if (S->ns == 0)
        return CB200_OK;
    k_reset_counters<<<1, 32, 0, S->stream>>>(S->d_cnt);
    CB_CUDA(cudaMemsetAsync(S->heavy.p, 0, S->nb, S->stream));
    k_aabb<<<blocks_for(S->ns, 256), 256, 0, S->stream>>>(S->ns, S->s_type.p, S->s_body.p, S->s_fam.p, S->s_mat.p,
                                                          S->s_lpos.p, S->s_lrot.p, S->s_dims.p, S->pos.p, S->rot.p,
                                                          S->bflags.p, S->set.collision_envelope, S->s_wpos.p,
                                                          S->s_wrot.p, S->s_bp.p, S->d_cnt);
    S->launches += 2;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

int stage_broadphase(System* S) {
    S->n_pairs = S->n_intersections = S->n_active_bins = 0;
    if (S->ns == 0)
        return CB200_OK;
    // sync point 1: global bounding box -> grid (host libm pow for FIXED_DENSITY)
    CB_TRY(fetch_counters(S));
    double mn[3], mx[3];
    for (int a = 0; a < 3; a++) {
        mn[a] = dec_ordered(S->h_cnt->bbox[a]);
        mx[a] = dec_ordered(S->h_cnt->bbox[3 + a]);
    }
    if (S->h_cnt->bbox[0] == ~0ull) {  // no colliding shape at all
        for (int a = 0; a < 3; a++)
            mn[a] = mx[a] = 0;
    }
    host_grid(S, mn, mx);
    const uint64_t nbins64 = (uint64_t)S->bins[0] * S->bins[1] * S->bins[2];
    if (nbins64 > (1ull << 27))
        return set_error(CB200_ERR_UNSUPPORTED, "broadphase grid has more than 2^27 bins");
    S->num_bins = (uint32_t)nbins64;
    const GridParams g = grid_params(S);

    CB_TRY(S->bin_intersections.reserve(S->ns + 1));
    CB_TRY(S->bin_count.reserve(S->num_bins + 1));
    CB_TRY(S->bin_cursor.reserve(S->num_bins + 1));
    CB_TRY(S->pair_count.reserve(S->num_bins + 1));
    CB_TRY(S->large_list.reserve(MAX_LARGE));

    // histogram (bin_cursor doubles as the 32-bit histogram, then as the scatter cursor)
    CB_CUDA(cudaMemsetAsync(S->bin_cursor.p, 0, (size_t)S->num_bins * 4, S->stream));
    k_bin_count<<<blocks_for(S->ns, 256), 256, 0, S->stream>>>(S->ns, g, S->s_bp.p, S->bin_intersections.p,
                                                               S->bin_cursor.p, S->large_list.p, S->heavy.p, S->d_cnt);
    k_bin_large<false><<<64, 256, 0, S->stream>>>(g, S->s_bp.p, S->large_list.p, S->d_cnt, S->bin_cursor.p, nullptr,
                                                  nullptr);
    k_bin_pack<<<blocks_for(S->num_bins, 256), 256, 0, S->stream>>>(S->num_bins, S->bin_cursor.p, S->bin_count.p);
    S->launches += 3;
    CB_TRY(scan_u64(S, S->bin_count.p, S->bin_count.p, S->num_bins));
    // sync point 2: N_i / N_a size the entry arrays
    uint64_t tot = 0;
    CB_CUDA(cudaMemcpyAsync(&tot, S->bin_count.p + S->num_bins, 8, cudaMemcpyDeviceToHost, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    S->n_intersections = (uint32_t)tot;
    S->n_active_bins = (uint32_t)(tot >> 32);
    if (S->n_intersections == 0)
        return CB200_OK;
    CB_TRY(S->bin_entries_tmp.reserve(S->n_intersections));
    CB_TRY(S->bin_entries.reserve(S->n_intersections));

    CB_CUDA(cudaMemsetAsync(S->bin_cursor.p, 0, (size_t)S->num_bins * 4, S->stream));
    k_bin_fill<<<blocks_for(S->ns, 256), 256, 0, S->stream>>>(S->ns, g, S->s_bp.p, S->bin_intersections.p,
                                                              S->bin_count.p, S->bin_cursor.p, S->bin_entries_tmp.p);
    k_bin_large<true><<<64, 256, 0, S->stream>>>(g, S->s_bp.p, S->large_list.p, S->d_cnt, S->bin_cursor.p,
                                                 S->bin_count.p, S->bin_entries_tmp.p);
    const int pair_blocks = blocks_for((size_t)S->num_bins * 32, 256);
    k_bin_pairs<false><<<pair_blocks, 256, 0, S->stream>>>(g, S->s_bp.p, S->bin_count.p, S->bin_entries_tmp.p,
                                                           S->bin_entries.p, S->pair_count.p, nullptr, nullptr,
                                                           S->d_cnt);
    S->launches += 3;
    CB_TRY(scan_u32(S, S->pair_count.p, S->pair_count.p, S->num_bins));
    // sync point 3: N_p sizes the pair list
    uint32_t np = 0;
    CB_CUDA(cudaMemcpyAsync(&np, S->pair_count.p + S->num_bins, 4, cudaMemcpyDeviceToHost, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    S->n_pairs = np;
    if (np == 0)
        return CB200_OK;
    CB_TRY(S->pairs.reserve(np));
    CB_TRY(S->pair_aux.reserve(np));
    k_bin_pairs<true><<<pair_blocks, 256, 0, S->stream>>>(g, S->s_bp.p, S->bin_count.p, S->bin_entries_tmp.p,
                                                          S->bin_entries.p, S->pair_count.p, S->pairs.p,
                                                          S->pair_aux.p, S->d_cnt);
    S->launches += 1;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

}  // namespace cb200
