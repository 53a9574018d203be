// broadphase.cu -- AABB generation, uniform-grid broadphase and the fused single-contact narrowphase on the device.
//
// Reference path replaced (CMC = src/chrono/collision/multicore):
//   GenerateAABB                      CMC/ChCollisionSystemMulticore.cpp:458-545   -> k_aabb (+ fused bbox reduce C1+C2,
//                                                                                   + PreprocessLocalToParent C14)
//   ChBroadphase::DetermineBoundingBox/ComputeTopLevelResolution  CMC/ChBroadphase.cpp:142-208 -> compute_grid()
//                                       (on the device, so the step needs no host round trip)
//   OffsetAABB + f_Count_AABB_BIN     CMC/ChBroadphase.cpp:168-176,261-268         -> k_bin_count
//   exclusive_scan / sort_by_key / RLE CMC/ChBroadphase.cpp:271-301                -> counting sort: histogram with
//                                       L2 atomics, packed u64 scan, scatter; the in-bin rank sort is
//                                       done by the pair kernel on its staged tile (pairs.cu)
//   f_Count/Store_AABB_AABB_Intersection  CMC/ChCollisionUtilsBroadphase.cpp:86-217
//   + PreprocessCount / DispatchPRIMS / remove_if  CMC/ChNarrowphase.cpp:94-155,264-320,366-385
//                                                                                  -> k_pairs<COUNT>, k_pairs<STORE>
//
// Design notes
//  * Instead of sorting N_i (bin, shape) keys we histogram into a dense bin array (atomics resolve in L2), scan it
//    once (entry offsets in the low word, active-bin index in the high word), scatter, and rank-sort the few entries
//    of each bin by shape id.  That reproduces exactly the (bin, shape) order of the reference's sort_by_key, so
//    pair emission order -- and with it the contact order -- equals the reference's.
//  * The pair kernels give one thread to every (bin, shape) entry; a CTA stages the records of its 256-entry tile
//    plus a forward halo in shared memory (the "neighbour-cell tile"), so the O(k^2) in-bin tests never touch
//    global memory.  A pair that passes the broadphase filter and has a single-contact analytic test (anything
//    against a sphere) is resolved on the spot: the candidate pair list is not written unless somebody asks for
//    it, and contacts come out already compacted, in the reference's order (stable: bin, shape A, shape B).
//  * Spheres carry no AABB record: (centre, radius) is 32 B and the box is recomputed with the reference's own
//    operations, (c - (r + envelope)) - global_origin, bit for bit.
#include "common.cuh"
#include "grid.cuh"
#include <float.h>
#include <limits.h>
#include <math.h>

namespace cb200 {

struct GridSettings {
    int type;  // 0 FIXED_RESOLUTION, 1 FIXED_BIN_SIZE, 2 FIXED_DENSITY
    int bins[3];
    double bin_size[3];
    double density;
    uint32_t ns;
    uint32_t cap_bins;
};

// --------------------------------------------------------------------------------------------------------------
// per-shape integer record; rebuilt only when shapes or body flags change
// --------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_build_meta(uint32_t ns, const int32_t* __restrict__ s_type,
                                                    const uint32_t* __restrict__ s_body,
                                                    const uint32_t* __restrict__ s_fam,
                                                    const int32_t* __restrict__ s_mat,
                                                    const double* __restrict__ s_lpos,
                                                    const uint8_t* __restrict__ bflags,
                                                    const int32_t* __restrict__ b_sphere /* null: slab systems */,
                                                    ShapeMeta* __restrict__ meta) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ns; i += gridDim.x * blockDim.x) {
        ShapeMeta m;
        m.body = s_body[i];
        m.fam = s_fam[i];
        m.pad = 0;
        const uint32_t type = (uint32_t)s_type[i] & 0xff;
        m.info = type << 8 | ((uint32_t)s_mat[i] & 0xffff) << 16;
        bool own = true;  // the body's own centred sphere (the first one it was given)
        if (m.body != UINT_MAX) {
            const uint32_t fl = bflags[m.body];
            m.info |= fl & 3u;
            if (b_sphere)
                own = b_sphere[m.body] == (int32_t)i;
            if (fl & 4u)
                m.body = UINT_MAX;  // vacant slot of a slab system: treated like a removed shape
        } else if (b_sphere) {
            own = false;
        }
        // bit 2: the body's sphere, centred on it (ObA == 0): world centre == body position, no rotation needed
        if (own && type == CB200_SPHERE && s_lpos[3 * i] == 0 && s_lpos[3 * i + 1] == 0 && s_lpos[3 * i + 2] == 0)
            m.info |= 4u;
        meta[i] = m;
    }
}

// keep_bbox: the integration kernel that ended the previous step has already put the bodies' own spheres into the box
__global__ void k_reset_counters(DevCounters* cnt, uint32_t nb, uint32_t ns, bool keep_bbox) {
    if (threadIdx.x == 0) {
        cnt->n_bodies = nb;
        cnt->n_shapes = ns;
        if (!keep_bbox) {
            cnt->bbox[0] = cnt->bbox[1] = cnt->bbox[2] = ~0ull;
            cnt->bbox[3] = cnt->bbox[4] = cnt->bbox[5] = 0ull;
        }
        cnt->n_large = 0;
        cnt->n_heavy = 0;
        cnt->ct_body_count = 0;
        // the counters of the pair pass that follows (pairs.cu pair_cnt_clean)
        cnt->n_pairs = 0;
        cnt->n_slow = cnt->n_long = cnt->n_pairq = 0;
        cnt->n_complex = cnt->n_cslots = 0;
        cnt->n_cls[0] = cnt->n_cls[1] = cnt->n_cls[2] = cnt->n_cls[3] = 0;
        cnt->n_bb[0] = cnt->n_bb[1] = cnt->n_bb[2] = cnt->n_bb[3] = cnt->n_bb[4] = 0;
        // cnt->unsupported is sticky: only the host clears it, after it has read it (api.cu clear_unsupported)
    }
}

// Everything but a body-centred sphere: shape frame, box per type (GenerateAABB, :486-541).
__device__ __forceinline__ bool aabb_general(uint32_t i, int type, uint32_t body, const d3& p, const double4& dm,
                                          const double* __restrict__ s_lpos, const double* __restrict__ s_lrot,
                                          double* s_dims, const double* __restrict__ s_verts,
                                          const double* __restrict__ rot, double envelope,
                                          double* __restrict__ s_wpos, double* __restrict__ s_wrot,
                                          ShapeBox* __restrict__ s_box, d3& amin, d3& amax) {
    bool ok = true;
    const double4 qa = reinterpret_cast<const double4*>(rot)[body];
    const q4 q = mkq(qa.x, qa.y, qa.z, qa.w);
    const d3 lp = mk3(s_lpos[3 * i], s_lpos[3 * i + 1], s_lpos[3 * i + 2]);
    const double4 la = reinterpret_cast<const double4*>(s_lrot)[i];
    const q4 lq = mkq(la.x, la.y, la.z, la.w);
    const d3 wp = rotate(lp, q) + p;  // == TransformLocalToParent(p, q, lp): IEEE addition commutes
    const q4 wq = qmul(q, lq);
    if (type == CB200_TRIANGLE) {
        // GenerateAABB :524-535 and PreprocessLocalToParent (CMC/ChNarrowphase.cpp:186-194): the vertices are
        // given in the BODY frame; no envelope on a triangle's box.  The world vertices (triangle_global) take
        // the shape's wpos / wrot / dims.yzw slots, which a triangle has no other use for (mpr.cuh ShapeView).
        const double* t = s_verts + 3 * (size_t)dm.x;
        const d3 A = rotate(mk3(t[0], t[1], t[2]), q) + p;
        const d3 B = rotate(mk3(t[3], t[4], t[5]), q) + p;
        const d3 C = rotate(mk3(t[6], t[7], t[8]), q) + p;
        reinterpret_cast<double4*>(s_wpos)[i] = make_double4(A.x, A.y, A.z, dm.x);
        reinterpret_cast<double4*>(s_wrot)[i] = make_double4(B.x, B.y, B.z, 0.0);
        s_dims[4 * i + 1] = C.x, s_dims[4 * i + 2] = C.y, s_dims[4 * i + 3] = C.z;
        amin = min3(A, min3(B, C));
        amax = max3(A, max3(B, C));
        ShapeBox b;
        b.mn[0] = amin.x, b.mn[1] = amin.y, b.mn[2] = amin.z;
        b.mx[0] = amax.x, b.mx[1] = amax.y, b.mx[2] = amax.z;
        s_box[i] = b;
        return true;
    }
    reinterpret_cast<double4*>(s_wpos)[i] = make_double4(wp.x, wp.y, wp.z, dm.x);
    reinterpret_cast<double4*>(s_wrot)[i] = make_double4(wq.w, wq.x, wq.y, wq.z);
    if (type == CB200_SPHERE) {
        const double r = dm.x + envelope;
        amin = wp - r;
        amax = wp + r;
    } else {
        d3 t = mk3(0);
        if (type == CB200_ELLIPSOID || type == CB200_BOX || type == CB200_CYLINDER ||
            type == CB200_CYLSHELL || type == CB200_CONE) {
            t = abs_rotate(wq, mk3(dm.x, dm.y, dm.z) + envelope);
        } else if (type == CB200_ROUNDEDBOX || type == CB200_ROUNDEDCYL) {
            t = abs_rotate(wq, mk3(dm.x, dm.y, dm.z) + dm.w + envelope);
        } else if (type == CB200_CAPSULE) {
            t = abs_rotate(wq, mk3(dm.x, dm.x + dm.y, dm.x) + envelope);
        } else if (type == CB200_CONVEXHULL) {
            // ComputeAABBConvex, CMC/ChCollisionSystemMulticore.cpp:424-456: Rotate(point + lpos, body*local)
            const double* h = s_verts + 3 * (size_t)dm.x;
            const int nv = (int)dm.y;
            d3 lo = rotate(mk3(h[0], h[1], h[2]) + lp, wq) + p, hi = lo;
            for (int k = 0; k < nv; k++) {
                const d3 c = rotate(mk3(h[3 * k], h[3 * k + 1], h[3 * k + 2]) + lp, wq) + p;
                lo = min3(lo, c);
                hi = max3(hi, c);
            }
            amin = lo - envelope;
            amax = hi + envelope;
            ShapeBox b;
            b.mn[0] = amin.x, b.mn[1] = amin.y, b.mn[2] = amin.z;
            b.mx[0] = amax.x, b.mx[1] = amax.y, b.mx[2] = amax.z;
            s_box[i] = b;
            return true;
        } else {
            ok = false;  // type without an AABB rule: the reference `continue`s and keeps stale values
        }
        amin = wp - t;
        amax = wp + t;
        if (!ok)
            amin = amax = mk3(0);
        ShapeBox b;
        b.mn[0] = amin.x, b.mn[1] = amin.y, b.mn[2] = amin.z;
        b.mx[0] = amax.x, b.mx[1] = amax.y, b.mx[2] = amax.z;
        s_box[i] = b;
    }
    return ok;
}

// --------------------------------------------------------------------------------------------------------------
// k_aabb: one thread per shape (grid-stride).  Two instantiations, two launches: <false> takes the body-centred
// spheres and nothing else (a lean, latency-bound gather that wants its occupancy), <true> everything else --
// through the list of such shapes kept by cb200_append_shapes, or (ids == nullptr, slab systems, whose shapes are
// written by the device) by scanning all shapes.
// --------------------------------------------------------------------------------------------------------------
template <bool GENERAL>
__global__ void __launch_bounds__(256) k_aabb(uint32_t ns, const uint32_t* __restrict__ ids,
                                              const ShapeMeta* __restrict__ meta,
                                              const double* __restrict__ s_lpos, const double* __restrict__ s_lrot,
                                              double* s_dims, const double* __restrict__ s_verts,
                                              const double* __restrict__ pos,
                                              const double* __restrict__ rot, double envelope,
                                              double* __restrict__ s_wpos, double* __restrict__ s_wrot,
                                              ShapeBox* __restrict__ s_box, DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    d3 mn = mk3(DBL_MAX), mx = mk3(-DBL_MAX);
    bool counted = false;
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < ns; t += gridDim.x * blockDim.x) {
        const uint32_t i = (GENERAL && ids) ? ids[t] : t;
        const ShapeMeta m = meta[i];
        if (m.body == UINT_MAX || ((m.info & 4u) != 0) == GENERAL)
            continue;
        const int type = (m.info >> 8) & 0xff;
        const uint32_t body = m.body;
        const d3 p = mk3(pos[3 * body], pos[3 * body + 1], pos[3 * body + 2]);
        const double4 dm = reinterpret_cast<const double4*>(s_dims)[i];
        d3 amin, amax;
        bool ok = true;
        if (!GENERAL) {
            // centred sphere: Rotate(0, q) + p == p
            reinterpret_cast<double4*>(s_wpos)[i] = make_double4(p.x, p.y, p.z, dm.x);
            const double r = dm.x + envelope;
            amin = p - r;
            amax = p + r;
        } else {
            d3 lo, hi;  // only these have their address taken: the sphere path stays in registers
            ok = aabb_general(i, type, body, p, dm, s_lpos, s_lrot, s_dims, s_verts, rot, envelope, s_wpos, s_wrot, s_box,
                              lo, hi);
            amin = lo, amax = hi;
        }
        if (ok && (m.info & 2u)) {
            mn = min3(mn, amin);
            mx = max3(mx, amax);
            counted = true;
        }
    }
    // union of the AABBs of colliding shapes (BoxInvert/BoxReduce, CMC/ChBroadphase.cpp:38-100)
    bbox_block_merge(cnt, mn, mx, counted);
}

// ---- active bounding box (ChCollisionSystemMulticore::Run, CMC/ChCollisionSystemMulticore.cpp:289-304 with
// GetOverlappingAABB, :547-562): a body that is active and collides stays active only if the box of one of its shapes
// overlaps the active box.  (The reference then puts the others to sleep, :275-283: the change is permanent.)
__global__ void __launch_bounds__(256) k_active_mark(uint32_t ns, const ShapeMeta* __restrict__ meta,
                                                     const double* __restrict__ s_wpos, const ShapeBox* __restrict__ s_box,
                                                     double envelope, d3 amin, d3 amax, uint8_t* __restrict__ inbox,
                                                     const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const double zero[3] = {0, 0, 0};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ns; i += gridDim.x * blockDim.x) {
        const ShapeMeta m = meta[i];
        if (m.body == UINT_MAX)
            continue;
        double mn[3], mx[3];
        shape_box_rel(i, m.info, s_wpos, s_box, zero, envelope, mn, mx);  // absolute: the boxes are not offset yet
        if ((amin.x <= mx[0] && mn[0] <= amax.x) && (amin.y <= mx[1] && mn[1] <= amax.y) && (amin.z <= mx[2] && mn[2] <= amax.z))
            inbox[m.body] = 1;
    }
}
__global__ void __launch_bounds__(256) k_active_apply(uint32_t nb, const uint8_t* __restrict__ inbox,
                                                      uint8_t* __restrict__ bflags, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < nb && (bflags[b] & 3u) == 3u && !inbox[b])
        bflags[b] &= ~1u;
}
__global__ void __launch_bounds__(256) k_active_meta(uint32_t ns, const uint8_t* __restrict__ bflags,
                                                     ShapeMeta* __restrict__ meta, const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ns; i += gridDim.x * blockDim.x) {
        const uint32_t body = meta[i].body;
        if (body != UINT_MAX && !(bflags[body] & 1u))
            meta[i].info &= ~1u;
    }
}

// DetermineBoundingBox + ComputeTopLevelResolution (CMC/ChBroadphase.cpp:142-208) and Compute_Grid_Resolution
// (CMC/ChCollisionUtilsBroadphase.cpp:36-46).  Evaluated by one thread per CTA; all CTAs get the same bits.
// (FIXED_DENSITY uses pow(): CUDA's differs from glibc's by at most 2 ulp, which can change the truncated bin count
// only if diag * cbrt(k N / V) lies within ~1e-15 of an integer.)
struct GridFull {
    GridParams g;
    double maxpt[3], bin_size[3];
};
__device__ GridFull compute_grid(const DevCounters* __restrict__ cnt, const GridSettings& st) {
    GridFull G;
    double mn[3], mx[3], diag[3];
    const bool none = cnt->bbox[0] == ~0ull;  // no colliding shape at all
    for (int a = 0; a < 3; a++) {
        mn[a] = none ? 0.0 : dec_ordered(cnt->bbox[a]);
        mx[a] = none ? 0.0 : dec_ordered(cnt->bbox[3 + a]);
    }
    const double fraction = 1e-3;
    for (int a = 0; a < 3; a++) {
        const double size = mx[a] - mn[a];
        const double minp = mn[a] - fraction * size;
        const double maxp = mx[a] + fraction * size;
        G.g.origin[a] = minp;
        G.maxpt[a] = maxp;
        diag[a] = fabs(maxp - minp);
    }
    if (st.type == 0) {
        for (int a = 0; a < 3; a++)
            G.g.bins[a] = st.bins[a];
    } else if (st.type == 1) {
        for (int a = 0; a < 3; a++)
            G.g.bins[a] = (int)ceil(diag[a] / st.bin_size[a]);
    } else {
        const double V = diag[0] * diag[1] * diag[2];
        const double scale = pow(st.density * st.ns / V, 1.0 / 3.0);
        for (int a = 0; a < 3; a++) {
            const int v = (int)(diag[a] * scale);
            G.g.bins[a] = v > 1 ? v : 1;
        }
    }
    unsigned long long nb = 1;
    for (int a = 0; a < 3; a++) {
        if (G.g.bins[a] < 1)
            G.g.bins[a] = 1;
        G.bin_size[a] = diag[a] / (double)G.g.bins[a];
        G.g.inv[a] = 1.0 / G.bin_size[a];
        nb *= (unsigned long long)G.g.bins[a];
    }
    G.g.num_bins = nb > 0xfffffff0ull ? 0xfffffff0u : (uint32_t)nb;
    return G;
}

__device__ __forceinline__ bool bin_range(const double* mn, const double* mx, const GridParams& g, int* lo, int* hi) {
    // HashMin / HashMax, CMC/ChCollisionUtils.h:44-60
    lo[0] = (int)floor(mn[0] * g.inv[0]);
    lo[1] = (int)floor(mn[1] * g.inv[1]);
    lo[2] = (int)floor(mn[2] * g.inv[2]);
    hi[0] = (int)ceil(mx[0] * g.inv[0]) - 1;
    hi[1] = (int)ceil(mx[1] * g.inv[1]) - 1;
    hi[2] = (int)ceil(mx[2] * g.inv[2]) - 1;
    // A shape outside the grid (only possible for shapes of non-colliding bodies, which are excluded from the
    // global box) would index outside the dense histogram; it can never form a pair, so it is not binned.
    return lo[0] >= 0 && lo[1] >= 0 && lo[2] >= 0 && hi[0] < g.bins[0] && hi[1] < g.bins[1] && hi[2] < g.bins[2] &&
           hi[0] >= lo[0] && hi[1] >= lo[1] && hi[2] >= lo[2];
}

// --------------------------------------------------------------------------------------------------------------
// k_bin_count: grid of this step + OffsetAABB (in place, non-spheres) + bins touched per shape + dense histogram.
// --------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bin_count(GridSettings st, double envelope,
                                                   const ShapeMeta* __restrict__ meta,
                                                   const double* __restrict__ s_wpos, ShapeBox* __restrict__ s_box,
                                                   uint32_t* __restrict__ bin_intersections,
                                                   uint32_t* __restrict__ hist, uint32_t* __restrict__ large_list,
                                                   uint8_t* __restrict__ heavy, DevCounters* __restrict__ cnt) {
    __shared__ GridFull sG;
    if (cnt->overflow)
        return;
    if (threadIdx.x == 0)
        sG = compute_grid(cnt, st);
    __syncthreads();
    const GridParams g = sG.g;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        for (int a = 0; a < 3; a++) {
            cnt->origin[a] = g.origin[a];
            cnt->maxpt[a] = sG.maxpt[a];
            cnt->bin_size[a] = sG.bin_size[a];
            cnt->inv_bin_size[a] = g.inv[a];
            cnt->bins[a] = g.bins[a];
        }
        cnt->num_bins = g.num_bins;
        if (g.num_bins > st.cap_bins) {
            cnt->need_bins = g.num_bins;
            atomicOr(&cnt->overflow, (uint32_t)OVF_BINS);
        }
    }
    if (g.num_bins > st.cap_bins)
        return;
    const uint32_t ns = st.ns;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ns; i += gridDim.x * blockDim.x) {
        int lo[3] = {0, 0, 0}, hi[3] = {-1, -1, -1};
        uint32_t n = 0;
        if (i < ns) {
            const ShapeMeta m = meta[i];
            if (m.body != UINT_MAX) {
                double mn[3], mx[3];
                if (((m.info >> 8) & 0xff) == CB200_SPHERE) {
                    shape_box_rel(i, m.info, s_wpos, s_box, g.origin, envelope, mn, mx);
                } else {
                    ShapeBox b = s_box[i];
                    for (int a = 0; a < 3; a++) {
                        mn[a] = b.mn[a] = b.mn[a] - g.origin[a];
                        mx[a] = b.mx[a] = b.mx[a] - g.origin[a];
                    }
                    s_box[i] = b;
                }
                if (bin_range(mn, mx, g, lo, hi))
                    n = (uint32_t)(hi[0] - lo[0] + 1) * (uint32_t)(hi[1] - lo[1] + 1) * (uint32_t)(hi[2] - lo[2] + 1);
                if (n > LARGE_BINS) {
                    const uint32_t k = atomicAdd(&cnt->n_large, 1u);
                    if (k < MAX_LARGE)
                        large_list[k] = i;
                    else
                        atomicOr(&cnt->overflow, (uint32_t)OVF_LARGE);
                    heavy[m.body] = 1;
                }
            }
            bin_intersections[i] = n;
        }
        const uint32_t mine = (n <= LARGE_BINS) ? n : 0;
        // the shape's bins, x slowest / z fastest, walked with counters (no division per bin).  Plain atomics: they
        // resolve in L2 at a few entries per bin; warp aggregation (match.any) cost more than it saved (measured).
        int ix = lo[0], iy = lo[1], iz = lo[2];
        for (uint32_t it = 0; it < mine; it++) {
            const uint32_t b = (uint32_t)((iz * g.bins[1]) * g.bins[0] + iy * g.bins[0] + ix);  // Hash_Index
            atomicAdd(&hist[b], 1u);
            if (++iz > hi[2]) {
                iz = lo[2];
                if (++iy > hi[1]) {
                    iy = lo[1];
                    ++ix;
                }
            }
        }
    }
}

// Shapes spanning many bins (container walls): one CTA per shape, threads stride over its bins.
template <bool FILL>
__global__ void __launch_bounds__(256) k_bin_large(const ShapeMeta* __restrict__ meta,
                                                   const double* __restrict__ s_wpos,
                                                   const ShapeBox* __restrict__ s_box, double envelope,
                                                   const uint32_t* __restrict__ large_list,
                                                   const DevCounters* __restrict__ cnt, uint32_t cap_entries,
                                                   uint32_t* __restrict__ cursor, const uint64_t* __restrict__ bin_start,
                                                   uint32_t* __restrict__ ent_tmp, uint32_t* __restrict__ ent_bin) {
    if (cnt->overflow)
        return;
    const GridParams g = load_grid(cnt);
    if (FILL && (uint32_t)bin_start[g.num_bins] > cap_entries)
        return;
    const uint32_t nl = min(cnt->n_large, (uint32_t)MAX_LARGE);
    if (nl == 0)
        return;
    // several CTAs share one large shape: work item w = (shape w % nl, part w / nl)
    const uint32_t nparts = max(1u, gridDim.x / nl);
    for (uint32_t w = blockIdx.x; w < nl * nparts; w += gridDim.x) {
        const uint32_t shape = large_list[w % nl];
        const uint32_t part = w / nl;
        double mn[3], mx[3];
        shape_box_rel(shape, meta[shape].info, s_wpos, s_box, g.origin, envelope, mn, mx);
        int lo[3], hi[3];
        if (!bin_range(mn, mx, g, lo, hi))
            continue;
        const int ny = hi[1] - lo[1] + 1, nz = hi[2] - lo[2] + 1;
        const uint32_t n = (uint32_t)(hi[0] - lo[0] + 1) * ny * nz;
        for (uint32_t it = part * blockDim.x + threadIdx.x; it < n; it += nparts * blockDim.x) {
            const int ix = lo[0] + (int)(it / (uint32_t)(ny * nz));
            const int rem = (int)(it % (uint32_t)(ny * nz));
            const int iy = lo[1] + rem / nz, iz = lo[2] + rem % nz;
            const uint32_t b = (uint32_t)((iz * g.bins[1]) * g.bins[0] + iy * g.bins[0] + ix);
            if (FILL) {
                const uint32_t slot = atomicSub(&cursor[b], 1u) - 1u;
                const uint32_t e = (uint32_t)bin_start[b] + slot;
                ent_tmp[e] = shape;
                ent_bin[e] = b;
            } else {
                atomicAdd(&cursor[b], 1u);
            }
        }
    }
}

// scatter: the histogram doubles as a countdown cursor, so no second clear is needed
__global__ void __launch_bounds__(256) k_bin_fill(uint32_t ns, double envelope, const ShapeMeta* __restrict__ meta,
                                                  const double* __restrict__ s_wpos,
                                                  const ShapeBox* __restrict__ s_box,
                                                  const uint32_t* __restrict__ bin_intersections,
                                                  const uint64_t* __restrict__ bin_start, uint32_t cap_entries,
                                                  uint32_t* __restrict__ cursor, uint32_t* __restrict__ ent_tmp,
                                                  uint32_t* __restrict__ ent_bin, DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const GridParams g = load_grid(cnt);
    const uint64_t tot = bin_start[g.num_bins];
    const uint32_t ni = (uint32_t)tot;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        cnt->n_intersections = ni;
        cnt->n_active_bins = (uint32_t)(tot >> 32);
        cnt->n_tiles = (ni + PAIR_TILE - 1) / PAIR_TILE;
        if (ni > cap_entries) {
            cnt->need_entries = ni;
            atomicOr(&cnt->overflow, (uint32_t)OVF_ENTRIES);
        }
    }
    if (ni > cap_entries)
        return;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < ns; i += gridDim.x * blockDim.x) {
        int lo[3] = {0, 0, 0}, hi[3] = {-1, -1, -1};
        uint32_t mine = 0;
        if (i < ns) {
            const uint32_t n = bin_intersections[i];
            if (n != 0 && n <= LARGE_BINS) {
                double mn[3], mx[3];
                shape_box_rel(i, meta[i].info, s_wpos, s_box, g.origin, envelope, mn, mx);
                bin_range(mn, mx, g, lo, hi);
                mine = n;
            }
        }
        int ix = lo[0], iy = lo[1], iz = lo[2];
        // four bins at a time: the returning atomics and the bin_start loads of a group are issued together, the stores
        // follow (one dependent round trip per group instead of one per entry)
        for (uint32_t it = 0; it < mine; it += 4) {
            uint32_t bb[4], slot[4], first[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                bb[k] = (uint32_t)((iz * g.bins[1]) * g.bins[0] + iy * g.bins[0] + ix);
                if (++iz > hi[2]) {
                    iz = lo[2];
                    if (++iy > hi[1]) {
                        iy = lo[1];
                        ++ix;
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (it + k < mine) {
                    slot[k] = atomicSub(&cursor[bb[k]], 1u) - 1u;
                    first[k] = (uint32_t)bin_start[bb[k]];
                }
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (it + k < mine) {
                    const uint32_t e = first[k] + slot[k];
                    ent_tmp[e] = i;
                    ent_bin[e] = bb[k];
                }
        }
    }
}

// --------------------------------------------------------------------------------------------------------------
// host side: enqueue only, no synchronisation
// --------------------------------------------------------------------------------------------------------------
int enqueue_build_meta(System* S) {
    k_build_meta<<<grid_for(S, S->ns, 256), 256, 0, S->stream>>>(S->ns, S->s_type.p, S->s_body.p, S->s_fam.p,
                                                                 S->s_mat.p, S->s_lpos.p, S->bflags.p,
                                                                 S->slab_on ? nullptr : S->b_sphere.p, S->s_meta.p);
    S->meta_dirty = false;
    S->sphere_aabb_valid = false;
    S->launches++;
    return CB200_OK;
}

int enqueue_aabb(System* S) {
    if (S->ns == 0)
        return CB200_OK;
    if (S->meta_dirty)
        CB_TRY(enqueue_build_meta(S));
    // the bodies' own spheres: done by the integration that ended the previous step, unless something touched the
    // state since (uploads, new bodies / shapes, flags, settings)
    const bool spheres_done = S->sphere_aabb_valid && !S->slab_on;
    if (spheres_done) {
        // their records were written to the other buffer (the previous step's stayed readable until now)
        std::swap(S->s_wpos, S->s_wpos_alt);
        S->wpos_parity ^= 1;
    }
    k_reset_counters<<<1, 32, 0, S->stream>>>(S->d_cnt, S->nb, S->ns, spheres_done);
    S->pair_cnt_clean = true;
    if (!S->body_scratch_clean)  // the integration of a whole step left the heavy flags cleared
        CB_CUDA(cudaMemsetAsync(S->heavy.p, 0, S->nb, S->stream));
    if (!spheres_done) {
        k_aabb<false><<<grid_for(S, S->ns, 256), 256, 0, S->stream>>>(
            S->ns, nullptr, S->s_meta.p, S->s_lpos.p, S->s_lrot.p, S->s_dims.p, S->s_verts.p, S->pos.p, S->rot.p,
            S->set.collision_envelope, S->s_wpos.p, S->s_wrot.p, S->s_box.p, S->d_cnt);
        S->launches++;
    }
    S->sphere_aabb_valid = false;  // consumed; the integration of a whole step sets it again
    const bool listed = !S->slab_on && S->general_ids.p;
    const uint32_t ng = listed ? S->n_general : S->ns;
    if (ng)
        k_aabb<true><<<grid_for(S, ng, 256), 256, 0, S->stream>>>(
            ng, listed ? S->general_ids.p : nullptr, S->s_meta.p, S->s_lpos.p, S->s_lrot.p, S->s_dims.p, S->s_verts.p,
            S->pos.p, S->rot.p, S->set.collision_envelope, S->s_wpos.p, S->s_wrot.p, S->s_box.p, S->d_cnt);
    S->launches += 2;
    if (S->use_aabb_active) {
        CB_TRY(S->b_inbox.reserve(S->nb));
        CB_CUDA(cudaMemsetAsync(S->b_inbox.p, 0, S->nb, S->stream));
        k_active_mark<<<grid_for(S, S->ns, 256), 256, 0, S->stream>>>(
            S->ns, S->s_meta.p, S->s_wpos.p, S->s_box.p, S->set.collision_envelope,
            mk3(S->active_min[0], S->active_min[1], S->active_min[2]), mk3(S->active_max[0], S->active_max[1], S->active_max[2]),
            S->b_inbox.p, S->d_cnt);
        k_active_apply<<<blocks_for(S->nb, 256), 256, 0, S->stream>>>(S->nb, S->b_inbox.p, S->bflags.p, S->d_cnt);
        k_active_meta<<<grid_for(S, S->ns, 256), 256, 0, S->stream>>>(S->ns, S->bflags.p, S->s_meta.p, S->d_cnt);
        S->launches += 3;
    }
    S->pairs_valid = false;
    S->box_relative = false;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

int enqueue_bins(System* S) {
    if (S->ns == 0)
        return CB200_OK;
    const cb200_settings& st = S->set;
    GridSettings gs;
    gs.type = st.broadphase_grid;
    for (int a = 0; a < 3; a++) {
        gs.bins[a] = st.bins_per_axis[a];
        gs.bin_size[a] = st.bin_size[a];
    }
    gs.density = st.grid_density;
    gs.ns = S->ns;
    gs.cap_bins = S->cap_bins;
    const double env = st.collision_envelope;
    const int sgrid = grid_for(S, S->ns, 256);
    S->box_relative = true;

    if (!S->bins_clean)
        CB_CUDA(cudaMemsetAsync(S->bin_cursor.p, 0, (size_t)S->cap_bins * 4, S->stream));
    S->bins_clean = false;
    CB_TRY(timer_start(S, "k_bin_count"));
    k_bin_count<<<sgrid, 256, 0, S->stream>>>(gs, env, S->s_meta.p, S->s_wpos.p, S->s_box.p, S->bin_intersections.p,
                                              S->bin_cursor.p, S->large_list.p, S->heavy.p, S->d_cnt);
    CB_TRY(timer_stop(S, "k_bin_count"));
    k_bin_large<false><<<S->num_sms * 4, 256, 0, S->stream>>>(S->s_meta.p, S->s_wpos.p, S->s_box.p, env, S->large_list.p, S->d_cnt,
                                                  S->cap_entries, S->bin_cursor.p, nullptr, nullptr, nullptr);
    S->launches += 2;
    CB_TRY(scan_bins(S, S->bin_cursor.p, S->bin_start.p, &S->d_cnt->num_bins, S->cap_bins));
    cudaStream_t s_large;  // large shapes (a container wall) scatter their entries beside everybody else's
    CB_TRY(aux_fork(S, 0, &s_large));
    CB_TRY(timer_start(S, "k_bin_fill"));
    k_bin_fill<<<sgrid, 256, 0, S->stream>>>(S->ns, env, S->s_meta.p, S->s_wpos.p, S->s_box.p, S->bin_intersections.p,
                                             S->bin_start.p, S->cap_entries, S->bin_cursor.p, S->ent_tmp.p,
                                             S->ent_bin.p, S->d_cnt);
    CB_TRY(timer_stop(S, "k_bin_fill"));
    k_bin_large<true><<<S->num_sms * 4, 256, 0, s_large>>>(S->s_meta.p, S->s_wpos.p, S->s_box.p, env, S->large_list.p, S->d_cnt,
                                                 S->cap_entries, S->bin_cursor.p, S->bin_start.p, S->ent_tmp.p,
                                                 S->ent_bin.p);
    CB_TRY(aux_join(S, 0));
    S->bins_clean = true;  // every count was taken back by the scatter (unless the step overflowed: api.cu resets this)
    // the in-bin sort by shape id happens in the pair kernel, on the staged tile (pairs.cu)
    S->launches += 2;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

int enqueue_broadphase(System* S, bool want_pairs, bool want_contacts) {
    CB_TRY(enqueue_bins(S));
    return enqueue_pairs_only(S, want_pairs, want_contacts);
}

}  // namespace cb200
