// pairs.cu -- in-bin pair tests fused with the narrowphase decision, and contact generation.
//
// Reference path replaced (CMC = src/chrono/collision/multicore):
//   f_Count/Store_AABB_AABB_Intersection  CMC/ChCollisionUtilsBroadphase.cpp:86-217  (per-bin O(k^2) filter)
//   PreprocessCount + Thrust_Expand       CMC/ChNarrowphase.cpp:94-155               (slot allocation)
//   DispatchPRIMS / DispatchHybridMPR     CMC/ChNarrowphase.cpp:264-320 + PRIMSCollision (ChNarrowphasePRIMS.cpp:1406-1607)
//   count + remove_if                     CMC/ChNarrowphase.cpp:366-385              (stable compaction)
//
// Kernels (instead of the reference's two bin passes + five-pass expand + dispatch + seven-array remove_if):
//   k_pair_masks   a CTA stages the records of its 256-entry tile plus a forward halo in shared memory as
//                  struct-of-arrays (conflict-free); every warp spreads the pairs of the bins it owns evenly over its
//                  lanes, applies the broadphase filter and -- for pairs with a single-contact analytic test (anything
//                  against a sphere) -- the cheap part of that test (no sqrt, no division) that decides whether a
//                  contact exists; the result is one 64-bit mask per (bin, shape) entry.
//   (scan of the per-entry counts)
//   k_pair_store   masks -> 64-bit shape-pair keys at their final, compacted slots (contacts; or the candidate pair
//                  list, which is materialised only when somebody asks for it).
//   k_contact_gen  one thread per contact, fully convergent: evaluates the analytic contact (normal, points,
//                  depth, effective radius) and writes the contact SoA.
// Order of the contact list = (bin, shape A, shape B) = the order the reference's stable compaction leaves.
#include "grid.cuh"
#include "prims.cuh"
#include "mpr.cuh"
#include "narrow_single.cuh"
#include "prims_multi.cuh"
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>

namespace cb200 {

constexpr int WIN = PAIR_TILE + PAIR_HALO;
constexpr int PAIR_THREADS = WIN;  // one thread per window entry
constexpr int PAIR_WARPS = PAIR_THREADS / 32;
constexpr int QCAP = 128;    // candidate queue of a warp: 32 rows x 4 candidates per round
constexpr int NNS_MAX = 4;   // non-spheres in a bin up to which the tile kernel keeps the bin's sphere-sphere pairs

// Shared-memory window of the slow (warp-per-bin) kernel, struct-of-arrays: lane-consecutive entries are
// word-consecutive in every array.
template <int N>
struct WindowT {
    double mn[3][N], mx[3][N];  // origin-relative AABB
    double c[4][N];             // world position + dims[0] (sphere: centre, radius)
    uint32_t body[N], fam[N], info[N], id[N];
};

// Window of the tile kernel.  Sphere records only (centre, radius: the AABB is recomputed from them with the very
// operations that produced it), plus a float copy for the conservative pre-test and the float classification.
struct TileSmem {
    double c[4][WIN];
    float4 f[2 * WIN];  // a bin at window offset w0 with k rows holds its records TWICE in a row, [2 w0, 2 w0 + 2k): row
                        // i reaches its partners i + 1 .. i + nd (mod k) as consecutive records, no wrap-around test.
                        // Centre relative to the bin's lower corner, in float; .w = upper bound of radius + envelope
                     // (margin included), its low 4 mantissa bits replaced by flags, stored INVERTED: bits 0..2 clear
                     // = "lower corner hashes to this bin's x / y / z", bit 3 clear = "sphere that can collide"
    float rm[WIN];   // lower bound of radius + envelope (margin subtracted)
    uint32_t body[WIN], fam[WIN], info[WIN], sid[WIN];
    unsigned long long smask[WIN];
    unsigned long long nsm[WIN];  // at the slot where a bin starts: which (sorted) slots of the bin hold non-spheres
    uint32_t queue[PAIR_WARPS][QCAP];
};

struct Rec {
    double mn[3], mx[3];
    double cx, cy, cz, r;
    uint32_t body, fam, info, id;
};

__device__ __forceinline__ void load_rec(Rec& R, uint32_t id, const ShapeMeta* __restrict__ meta,
                                         const double* __restrict__ s_wpos, const ShapeBox* __restrict__ s_box,
                                         const double* origin, double envelope) {
    const ShapeMeta m = meta[id];
    const double4 w = reinterpret_cast<const double4*>(s_wpos)[id];
    R.cx = w.x, R.cy = w.y, R.cz = w.z, R.r = w.w;
    R.body = m.body, R.fam = m.fam, R.info = m.info, R.id = id;
    shape_box_rel(id, m.info, s_wpos, s_box, origin, envelope, R.mn, R.mx);
}

// info bits 3..5 of a record staged for bin (bx, by, bz): the shape's lower corner hashes to the bin's x / y / z
// coordinate.  HashMin is monotone, so current_bin(A, B) (CMC/ChCollisionUtils.h:97-110), i.e.
// Hash(max(Amin, Bmin)) == bin, holds iff on every axis at least one of the two entries carries the flag (both
// lower corners hash to <= the bin coordinate).
__device__ __forceinline__ uint32_t corner_flags(const double* mn, const GridParams& g, int bx, int by, int bz) {
    return ((int)floor(mn[0] * g.inv[0]) == bx ? 8u : 0u) | ((int)floor(mn[1] * g.inv[1]) == by ? 16u : 0u) |
           ((int)floor(mn[2] * g.inv[2]) == bz ? 32u : 0u);
}

__device__ __forceinline__ void bin_coords(uint32_t b, const GridParams& g, int& bx, int& by, int& bz) {
    bx = (int)(b % (uint32_t)g.bins[0]);
    by = (int)((b / (uint32_t)g.bins[0]) % (uint32_t)g.bins[1]);
    bz = (int)(b / ((uint32_t)g.bins[0] * (uint32_t)g.bins[1]));
}

// record of shape id as an entry of bin (bx, by, bz): gathered (the per-shape arrays sit in L2), corner flags set
__device__ __forceinline__ void load_rec_bin(Rec& R, uint32_t id, const ShapeMeta* __restrict__ meta,
                                             const double* __restrict__ s_wpos, const ShapeBox* __restrict__ s_box,
                                             const GridParams& g, int bx, int by, int bz, double envelope) {
    load_rec(R, id, meta, s_wpos, s_box, g.origin, envelope);
    R.info = (R.info & ~56u) | corner_flags(R.mn, g, bx, by, bz);
}

template <int N>
__device__ __forceinline__ void win_put(WindowT<N>& W, int k, const Rec& R) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
        W.mn[a][k] = R.mn[a];
        W.mx[a][k] = R.mx[a];
    }
    W.c[0][k] = R.cx, W.c[1][k] = R.cy, W.c[2][k] = R.cz, W.c[3][k] = R.r;
    W.body[k] = R.body, W.fam[k] = R.fam, W.info[k] = R.info, W.id[k] = R.id;
}

template <int N>
__device__ __forceinline__ void win_get(const WindowT<N>& W, int k, Rec& R) {
#pragma unroll
    for (int a = 0; a < 3; a++) {
        R.mn[a] = W.mn[a][k];
        R.mx[a] = W.mx[a][k];
    }
    R.cx = W.c[0][k], R.cy = W.c[1][k], R.cz = W.c[2][k], R.r = W.c[3][k];
    R.body = W.body[k], R.fam = W.fam[k], R.info = W.info[k], R.id = W.id[k];
}

// CMC/ChCollisionUtilsBroadphase.cpp:114-140 (a pure conjunction, so the evaluation order is free)
// FLAGS: both records come from the bin-sorted copies, whose info bits 3..5 say "lower corner hashes to this bin's
// x / y / z" (k_bin_sort): current_bin then is a bit test instead of three floor(max * inv).
template <bool FLAGS>
__device__ __forceinline__ bool pair_passes(const Rec& A, const Rec& B, const GridParams& g, uint32_t bin) {
    if (!((A.mn[0] <= B.mx[0] && B.mn[0] <= A.mx[0]) && (A.mn[1] <= B.mx[1] && B.mn[1] <= A.mx[1]) &&
          (A.mn[2] <= B.mx[2] && B.mn[2] <= A.mx[2])))
        return false;
    // current_bin (CMC/ChCollisionUtils.h:97-110): the pair belongs to the bin holding max(Amin, Bmin)
    if (FLAGS) {
        if (((A.info | B.info) & 56u) != 56u)
            return false;
    } else {
        const int cx = (int)floor(fmax(A.mn[0], B.mn[0]) * g.inv[0]);
        const int cy = (int)floor(fmax(A.mn[1], B.mn[1]) * g.inv[1]);
        const int cz = (int)floor(fmax(A.mn[2], B.mn[2]) * g.inv[2]);
        if ((uint32_t)((cz * g.bins[1]) * g.bins[0] + cy * g.bins[0] + cx) != bin)
            return false;
    }
    if (B.body == UINT_MAX || A.body == B.body)
        return false;
    if (!(B.info & 2u))
        return false;
    if (!(A.info & 1u) && !(B.info & 1u))
        return false;
    const uint32_t gA = A.fam & 0xffffu, mA = A.fam >> 16, gB = B.fam & 0xffffu, mB = B.fam >> 16;
    return (mA & gB) && (mB & gA);
}

// Full shape view for the complex path: spheres carry their true world rotation too (MPR rotates the search
// direction into the shape frame, ChCollisionUtilsMPR.cpp:71-74); a body-centred sphere has no s_wrot entry.
__device__ __forceinline__ ShapeView make_view_full(uint32_t id, const ShapeMeta& m, const double* __restrict__ s_wpos,
                                                    const double* __restrict__ s_wrot,
                                                    const double* __restrict__ s_dims,
                                                    const double* __restrict__ s_lrot,
                                                    const double* __restrict__ rot,
                                                    const double* __restrict__ s_verts) {
    ShapeView v;
    v.type = (m.info >> 8) & 0xff;
    v.hull = nullptr, v.nv = 0;
    const double4 w = reinterpret_cast<const double4*>(s_wpos)[id];
    v.pos = mk3(w.x, w.y, w.z);
    const double4 d = reinterpret_cast<const double4*>(s_dims)[id];
    v.d0 = d.x, v.d1 = d.y, v.d2 = d.z, v.d3_ = d.w;
    if (v.type == CB200_CONVEXHULL) {  // dims = (first vertex, count) into the vertex pool
        v.hull = s_verts + 3 * (size_t)d.x;
        v.nv = (int)d.y;
    }
    if (m.info & 4u) {
        const double4 qa = reinterpret_cast<const double4*>(rot)[m.body];
        const double4 la = reinterpret_cast<const double4*>(s_lrot)[id];
        v.rot = qmul(mkq(qa.x, qa.y, qa.z, qa.w), mkq(la.x, la.y, la.z, la.w));
    } else {
        const double4 r = reinterpret_cast<const double4*>(s_wrot)[id];
        v.rot = mkq(r.x, r.y, r.z, r.w);
    }
    return v;
}

// slots the reference reserves for a pair (PreprocessCount, CMC/ChNarrowphase.cpp:94-137)
__device__ __forceinline__ uint32_t pair_slots(int tA, int tB, int algorithm) {
    if (algorithm == 0)
        return 1;
    if (tA == CB200_SPHERE || tB == CB200_SPHERE)
        return 1;
    if (tA == CB200_CAPSULE || tB == CB200_CAPSULE)
        return 2;
    if (tA == CB200_CYLSHELL || tB == CB200_CYLSHELL)
        return 8;
    if (tA == CB200_BOX && tB == CB200_BOX)
        return 8;
    if ((tA == CB200_BOX && tB == CB200_TRIANGLE) || (tA == CB200_TRIANGLE && tB == CB200_BOX))
        return 6;  // slots only: PRIMSCollision reports triangle_box as unsupported (:1590-1603), MPR decides in HYBRID
    return 1;
}

__device__ __forceinline__ void flip_contacts(Contact* out, int n) {
    for (int i = 0; i < n && i < MAX_PAIR_CONTACTS; i++) {
        const d3 t = out[i].p1;
        out[i].p1 = out[i].p2;
        out[i].p2 = t;
        out[i].n = -out[i].n;
    }
}

// Narrowphase of one complex pair: DispatchMPR / DispatchPRIMS / DispatchHybridMPR (CMC/ChNarrowphase.cpp:232-320)
// over PRIMSCollision (ChNarrowphasePRIMS.cpp:1406-1607) and MPRCollision (ChNarrowphaseMPR.cpp:627-645).
// Returns the number of contacts written to out[0..].
__device__ __noinline__ int narrow_complex(const ShapeView& A, const ShapeView& B, int algorithm, double envelope,
                                           Contact* out) {
    const double sep = 2 * envelope;
    if (algorithm != 0) {
        int n = prims_single(A, B, sep, out[0]);
        if (n >= 0)
            return n;
        const int tA = A.type, tB = B.type;
        if (tA == CB200_CAPSULE && tB == CB200_CAPSULE)
            return capsule_capsule(A.pos, A.rot, A.d0, A.d1, B.pos, B.rot, B.d0, B.d1, sep, out);
        if (tA == CB200_BOX && tB == CB200_CAPSULE)
            return box_capsule(A.pos, A.rot, mk3(A.d0, A.d1, A.d2), B.pos, B.rot, B.d0, B.d1, sep, out);
        if (tA == CB200_CAPSULE && tB == CB200_BOX) {
            n = box_capsule(B.pos, B.rot, mk3(B.d0, B.d1, B.d2), A.pos, A.rot, A.d0, A.d1, sep, out);
            flip_contacts(out, n);
            return n;
        }
        if (tA == CB200_BOX && tB == CB200_BOX)
            return box_box(A.pos, A.rot, mk3(A.d0, A.d1, A.d2), B.pos, B.rot, mk3(B.d0, B.d1, B.d2), sep, out);
        if (tA == CB200_BOX && tB == CB200_CYLSHELL) {
            n = box_cylshell(A.pos, A.rot, mk3(A.d0, A.d1, A.d2), B.pos, B.rot, B.d0, B.d1, sep, out);
            if (n >= 0)
                return n;  // -1: not an axis-aligned configuration, PRIMSCollision returns false -> MPR in HYBRID
        }
        if (tA == CB200_CYLSHELL && tB == CB200_BOX) {
            n = box_cylshell(B.pos, B.rot, mk3(B.d0, B.d1, B.d2), A.pos, A.rot, A.d0, A.d1, sep, out);
            if (n >= 0) {
                flip_contacts(out, n);
                return n;
            }
        }
        if (algorithm == 1)
            return 0;  // PRIMS only: no analytic test, no contact
    }
    if (mpr_collision(A, B, envelope, out[0].n, out[0].p1, out[0].p2, out[0].depth)) {
        out[0].erad = DEFAULT_EFF_RADIUS;
        return 1;
    }
    return 0;
}

// What the long-bin paths (more than 64 entries in a bin: no masks, no records) need to run a complex pair in place.
struct CplxArgs {
    const double *s_lrot, *rot, *s_verts;
    double* cplx_out;
    uint32_t cap_slots;
};

__device__ __forceinline__ int narrow_complex_ids(uint32_t sA, uint32_t sB, const ShapeMeta* __restrict__ meta,
                                                  const double* __restrict__ s_wpos, const double* __restrict__ s_wrot,
                                                  const double* __restrict__ s_dims, const CplxArgs& X, int algorithm,
                                                  double envelope, Contact* out) {
    const ShapeView A = make_view_full(sA, meta[sA], s_wpos, s_wrot, s_dims, X.s_lrot, X.rot, X.s_verts);
    const ShapeView B = make_view_full(sB, meta[sB], s_wpos, s_wrot, s_dims, X.s_lrot, X.rot, X.s_verts);
    int n = narrow_complex(A, B, algorithm, envelope, out);
    const int slots = (int)pair_slots(A.type, B.type, algorithm);
    return n < 0 ? 0 : (n > slots ? slots : n);
}

// cold path of pair_hit (a non-sphere is involved): kept out of line so that the sphere-sphere loop stays lean
__device__ __noinline__ int pair_hit_general(uint32_t idA, int tA, double4 wA, uint32_t idB, int tB, double4 wB,
                                             double sep, const double* __restrict__ s_wrot,
                                             const double* __restrict__ s_dims) {
    Contact c;
    return prims_single(make_view(idA, tA, wA, s_wrot, s_dims), make_view(idB, tB, wB, s_wrot, s_dims), sep, c);
}

// decision only: does the pair (A, B) produce a contact?  1 / 0 / -1 (not a single-contact analytic pair)
__device__ __forceinline__ int pair_hit(const Rec& A, const Rec& B, double sep, int algorithm,
                                        const double* __restrict__ s_wrot, const double* __restrict__ s_dims) {
    if (algorithm == 0)
        return -1;  // MPR for everything
    const int tA = (A.info >> 8) & 0xff, tB = (B.info >> 8) & 0xff;
    if (tA == CB200_SPHERE && tB == CB200_SPHERE) {
        // the deciding lines of sphere_sphere, ChNarrowphasePRIMS.cpp:52-60
        const d3 delta = mk3(B.cx, B.cy, B.cz) - mk3(A.cx, A.cy, A.cz);
        const double dist2 = dot3(delta, delta);
        const double radSum_s = (A.r + B.r) + sep;
        return !(dist2 >= radSum_s * radSum_s || dist2 < 1e-12);
    }
    return pair_hit_general(A.id, tA, make_double4(A.cx, A.cy, A.cz, A.r), B.id, tB,
                            make_double4(B.cx, B.cy, B.cz, B.r), sep, s_wrot, s_dims);
}

template <typename T>
__device__ __forceinline__ T warp_incl_scan_bp(T v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T o = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d)
            v += o;
    }
    return v;
}


// --------------------------------------------------------------------------------------------------------------
// k_pair_masks: the tile kernel.  A CTA takes a 256-entry tile of the (bin, shape) entry list plus a 64-entry forward
// halo and owns the bins that start in the tile; one thread per window entry.
//
//   stage    every thread fetches one entry (unsorted inside its bin, as the scatter left it), gathers the shape's
//            records (L2-resident), ranks its shape id among the bin's ids in shared memory and drops the record at
//            its sorted slot: the in-bin sort by shape id costs no pass of its own, and ent_shape (bin_aabb_number)
//            is written from here.
//   phase 1  thread t = row t of the sorted window.  Row i meets its partners j = i + d (mod k), d = 1..(k-1)/2
//            (+ half a diagonal when k is even): every unordered pair once, the same work for every row.  The test
//            is a CONSERVATIVE float test (|dc| <= rA + rB + margin on every axis, corner flags, both can collide): it
//            can only say "certainly not", so list membership is still decided by the exact test alone.  ~7 of 8 pairs
//            end here after one 16-byte shared load and a dozen FP32 / integer instructions.
//   phase 2  the survivors of a warp's 32 rows are compacted through a shared queue and spread over all lanes; each
//            gets the reference's test in FP64, operation for operation: AABB overlap
//            (f_Count_AABB_AABB_Intersection, CMC/ChCollisionUtilsBroadphase.cpp:114-140), current_bin, the body /
//            family filters and -- in CONTACTS mode -- the deciding lines of sphere_sphere (ChNarrowphasePRIMS.cpp:52-60).
//            A hit sets bit max(i,j) in the 64-bit mask of entry min(i,j).
// The masks are all the store pass needs: its output slot for (i, j) is offset[i] + popc(mask[i] below bit j), i.e.
// the reference's emission order (bin, shape A, shape B), without re-running a single test.
// Bins with a few non-spheres (container walls): the sphere-sphere pairs stay here, the rows of the non-spheres are
// queued for k_pair_masks_slow.  Bins with many non-spheres, with more than 64 entries, or all bins when MPR is
// requested for everything are queued as a whole.
// --------------------------------------------------------------------------------------------------------------
enum { MODE_PAIRS = 0, MODE_CONTACTS = 1 };

// 0: no output, 1: the pair yields one contact (or, in PAIRS mode, passes the broadphase filter),
// 2: the pair needs the complex path (multi-contact analytic test or MPR)
template <bool FLAGS>
__device__ __forceinline__ int pair_decide(const Rec& A, const Rec& B, const GridParams& g, uint32_t bin, int mode,
                                           double sep, int algorithm, const double* __restrict__ s_wrot,
                                           const double* __restrict__ s_dims, uint32_t& np) {
    if (A.body == UINT_MAX || !(A.info & 2u))
        return 0;
    if (!pair_passes<FLAGS>(A, B, g, bin))
        return 0;
    np++;
    if (mode == MODE_PAIRS)
        return 1;
    const int n = pair_hit(A, B, sep, algorithm, s_wrot, s_dims);
    return n < 0 ? 2 : n;
}

// rank += (v < id): one compare + one predicated add (the compiler's own select sequence is a third longer, and this
// is the inner statement of the in-bin sort)
__device__ __forceinline__ void count_less(uint32_t& rank, uint32_t v, uint32_t id) {
    asm("{ .reg .pred p; setp.lt.u32 p, %1, %2; @p add.u32 %0, %0, 1; }" : "+r"(rank) : "r"(v), "r"(id));
}

template <bool CPLX>
__global__ void __launch_bounds__(PAIR_THREADS, 4) k_pair_masks(const ShapeMeta* __restrict__ meta,
                                                             const double* __restrict__ s_wpos,
                                                             const ShapeBox* __restrict__ s_box,
                                                             const uint32_t* __restrict__ ent_tmp,
                                                             const uint32_t* __restrict__ ent_bin,
                                                             const uint64_t* __restrict__ bin_start, double envelope,
                                                             int algorithm, int mode, uint32_t* __restrict__ ent_shape,
                                                             uint64_t* __restrict__ ent_mask,
                                                             uint64_t* __restrict__ ent_cmask,
                                                             uint32_t* __restrict__ ent_cnt,
                                                             uint32_t* __restrict__ slow_bins,
                                                             uint32_t* __restrict__ long_bins,
                                                             uint2* __restrict__ pair_q, uint32_t cap_pairq,
                                                             DevCounters* __restrict__ cnt) {
    __shared__ TileSmem T;
    if (cnt->overflow)
        return;
    const GridParams g = load_grid(cnt);
    const uint32_t ni = cnt->n_intersections;
    const uint32_t ntiles = cnt->n_tiles;
    const double sep = 2 * envelope;
    // Margins of the float tests.  A centre is kept relative to its bin's lower corner, |coordinate| <= B + R (B the
    // largest bin edge, R = radius + envelope), so with u = 2^-24: a float centre is off by <= u (B + R), a float
    // difference of two by another u |d|, the float norm of the difference by <= 7u (2B + RA + RB) in all.  Each
    // entry carries the margin m = 16u (B + R) + 2^-45 Eabs (Eabs: largest absolute coordinate of the grid; covers the
    // FP64 roundings of the exact tests themselves): a float comparison against RA + RB + mA + mB can only say
    // "certainly not", one against RA + RB - mA - mB only "certainly yes" (error analysis in DESIGN.md section 4.1).
    const double Bmax = fmax(fmax(cnt->bin_size[0], cnt->bin_size[1]), cnt->bin_size[2]);
    const double Eabs = fmax(fmax(fmax(fabs(cnt->origin[0]), fabs(cnt->maxpt[0])), fmax(fabs(cnt->origin[1]), fabs(cnt->maxpt[1]))),
                             fmax(fabs(cnt->origin[2]), fabs(cnt->maxpt[2])));
    const double m_abs = Eabs * (1.0 / 35184372088832.0);  // 2^-45
    const double m_rel = 16.0 / 16777216.0;                // 16 u
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const uint32_t t = threadIdx.x;  // window index (WIN == blockDim.x)
    uint32_t np = 0;

    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const uint32_t e0 = tile * PAIR_TILE;
        const uint32_t wend = min(ni, e0 + WIN);
        const uint32_t e = e0 + t;
        __syncthreads();  // previous tile's readers are done with the window
        T.smask[t] = 0ull;
        T.nsm[t] = 0ull;
        // ---- stage: one entry per thread, unsorted inside its bin
        uint32_t id = 0, bin = 0, s = 0, k = 0;
        ShapeMeta m = {};
        double4 w = make_double4(0, 0, 0, 0);
        const bool have = e < wend;
        if (have) {
            id = ent_tmp[e];
            bin = ent_bin[e];
            s = (uint32_t)bin_start[bin];
            k = (uint32_t)bin_start[bin + 1] - s;
            T.sid[t] = id;
            m = meta[id];
            w = reinterpret_cast<const double4*>(s_wpos)[id];
        }
        // the tile owns the bins that start in its first PAIR_TILE entries, including their rows in the halo
        const bool own = have && s >= e0 && s < e0 + PAIR_TILE;
        const bool sphere = ((m.info >> 8) & 0xff) == CB200_SPHERE;
        const uint32_t w0 = own ? s - e0 : 0u;
        __syncthreads();
        if (own && k <= 64) {
            // rank of this shape id among the bin's ids == its position in the reference's sorted (bin, shape) list
            uint32_t rank = 0, x = w0;
            const uint32_t xe = w0 + k;
            for (; (x & 3u) && x < xe; x++)
                count_less(rank, T.sid[x], id);
            for (; x + 4 <= xe; x += 4) {
                const uint4 v = *reinterpret_cast<const uint4*>(&T.sid[x]);
                count_less(rank, v.x, id);
                count_less(rank, v.y, id);
                count_less(rank, v.z, id);
                count_less(rank, v.w, id);
            }
            for (; x < xe; x++)
                count_less(rank, T.sid[x], id);
            const uint32_t p = w0 + rank;
            ent_shape[s + rank] = id;
            if (!sphere)
                atomicOr(&T.nsm[w0], 1ull << rank);
            double mn[3], mx[3];
            shape_box_rel(id, m.info, s_wpos, s_box, g.origin, envelope, mn, mx);
            int bx, by, bz;
            bin_coords(bin, g, bx, by, bz);
            const uint32_t fl = corner_flags(mn, g, bx, by, bz);
            T.c[0][p] = w.x, T.c[1][p] = w.y, T.c[2][p] = w.z, T.c[3][p] = w.w;
            T.body[p] = m.body, T.fam[p] = m.fam, T.info[p] = (m.info & ~56u) | fl;
            if (!sphere) {
                // nobody reads a non-sphere's row in this kernel (its pairs are decided by k_pair_queue from the global
                // records), so the row keeps the shape's box instead: the sphere rows of the bin test against it before
                // they queue a pair
                T.c[0][p] = mn[0], T.c[1][p] = mn[1], T.c[2][p] = mn[2], T.c[3][p] = mx[0];
                T.body[p] = (uint32_t)__double2hiint(mx[1]), T.fam[p] = (uint32_t)__double2loint(mx[1]);
                T.info[p] = (uint32_t)__double2hiint(mx[2]), T.rm[p] = __int_as_float(__double2loint(mx[2]));
            }
            const bool ok = sphere && m.body != UINT_MAX && (m.info & 2u);
            const double R = w.w + envelope;
            const double me = m_rel * (Bmax + R) + m_abs;
            // upper bound of R + me that survives the replacement of its 4 low mantissa bits by the flags
            const uint32_t rp = ((__float_as_uint(__double2float_ru(R + me)) & ~15u) + 16u) | (~((fl >> 3) | (ok ? 8u : 0u)) & 15u);
            if (sphere)
                T.rm[p] = __double2float_rd(R - me);
            T.f[p + w0] = T.f[p + w0 + k] =
                make_float4(__double2float_rn(w.x - (g.origin[0] + bx * cnt->bin_size[0])),
                            __double2float_rn(w.y - (g.origin[1] + by * cnt->bin_size[1])),
                            __double2float_rn(w.z - (g.origin[2] + bz * cnt->bin_size[2])), __uint_as_float(rp));
        }
        __syncthreads();
        // ---- from here on thread t is row t of the sorted window (same bin, hence same s and k, as its staged entry)
        const unsigned long long nsm = (own && k <= 64) ? T.nsm[w0] : 0ull;
        const uint32_t nn = (uint32_t)__popcll(nsm);
        const bool bin_here = algorithm != 0 && nn <= NNS_MAX && k >= 2 && k <= 64;  // sphere-sphere pairs done here
        const bool whole_slow = own && (k > 64 || (k >= 2 && !bin_here));
        const uint32_t i = e - s;
        uint32_t cand = 0u, nd = 0u;  // bit (nd - d) of cand: partner j = i + d (mod k) may pair with this row
        if (own && bin_here) {
            const float4 A = T.f[w0 + t];
            const uint32_t ab = __float_as_uint(A.w);
            if (!(ab & 8u)) {
                // a partner must be a sphere that can collide (its bit 3 clear) and, on every axis, one of the two
                // lower corners must hash to this bin (not both inverted bits set): one AND against this row's mask
                const uint32_t am = (ab & 7u) | 8u;
                nd = ((k - 1) >> 1) + (((k & 1u) || i >= (k >> 1)) ? 0u : 1u);
                const float4* P = &T.f[w0 + t + 1];
#define CB_PHASE1_TEST(B)                                                                                          \
    {                                                                                                              \
        const float rs = A.w + (B).w;                                                                              \
        const bool maybe = (fabsf((B).x - A.x) <= rs) & (fabsf((B).y - A.y) <= rs) & (fabsf((B).z - A.z) <= rs) &  \
                           ((__float_as_uint((B).w) & am) == 0u);                                                  \
        cand = (cand << 1) | (maybe ? 1u : 0u);                                                                    \
    }
                uint32_t d = 0;
                for (; d + 4 <= nd; d += 4) {
                    const float4 B0 = P[d], B1 = P[d + 1], B2 = P[d + 2], B3 = P[d + 3];
                    CB_PHASE1_TEST(B0) CB_PHASE1_TEST(B1) CB_PHASE1_TEST(B2) CB_PHASE1_TEST(B3)
                }
                for (; d < nd; d++) {
                    const float4 B0 = P[d];
                    CB_PHASE1_TEST(B0)
                }
#undef CB_PHASE1_TEST
            }
        }
        // ---- phase 2: the candidates of the warp's rows, compacted and spread over the lanes
        while (__any_sync(0xffffffffu, cand != 0u)) {
            const uint32_t c = min((uint32_t)__popc(cand), (uint32_t)(QCAP / 32));
            const uint32_t incl = warp_incl_scan_bp(c, lane);
            const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
            uint32_t q = incl - c;
            for (uint32_t x = 0; x < c; x++) {
                const uint32_t d = nd - ((uint32_t)__ffs((int)cand) - 1u);
                cand &= cand - 1u;
                uint32_t j = i + d;
                if (j >= k)
                    j -= k;
                T.queue[wid][q++] = (w0 << 12) | (i << 6) | j;
            }
            __syncwarp();
            for (uint32_t x = lane; x < total; x += 32) {
                const uint32_t item = T.queue[wid][x];
                const uint32_t bw = item >> 12, pi = (item >> 6) & 63u, pj = item & 63u;
                const uint32_t lo = min(pi, pj), hi = max(pi, pj);
                const uint32_t sa = bw + lo, sb = bw + hi;
                // float classification: certainly overlapping boxes?  certainly touching / certainly apart spheres?
                const float4 FA = T.f[sa + bw], FB = T.f[sb + bw];
                const float dx = FB.x - FA.x, dy = FB.y - FA.y, dz = FB.z - FA.z;
                const float rsM = T.rm[sa] + T.rm[sb], rsP = FA.w + FB.w;
                const float d2 = dx * dx + dy * dy + dz * dz;
                const bool sure_box = fmaxf(fmaxf(fabsf(dx), fabsf(dy)), fabsf(dz)) <= rsM;
                const bool sure_hit = d2 <= rsM * rsM * 0.99999904f && d2 >= 1e-10f && rsM > 0.f;
                const bool sure_miss = d2 >= rsP * rsP * 1.00000096f;
                bool ov = true, touch = sure_hit;
                if (!sure_box || (mode == MODE_CONTACTS && !sure_hit && !sure_miss)) {
                    // the reference's tests in FP64, operation for operation.  Both boxes exactly as GenerateAABB +
                    // OffsetAABB leave them: (c -/+ (r + envelope)) - global_origin
                    const double ax = T.c[0][sa], ay = T.c[1][sa], az = T.c[2][sa], ar = T.c[3][sa];
                    const double bx = T.c[0][sb], by = T.c[1][sb], bz = T.c[2][sb], br = T.c[3][sb];
                    const double ra = ar + envelope, rb = br + envelope;
                    const double an0 = (ax - ra) - g.origin[0], an1 = (ay - ra) - g.origin[1], an2 = (az - ra) - g.origin[2];
                    const double ax0 = (ax + ra) - g.origin[0], ax1 = (ay + ra) - g.origin[1], ax2 = (az + ra) - g.origin[2];
                    const double bn0 = (bx - rb) - g.origin[0], bn1 = (by - rb) - g.origin[1], bn2 = (bz - rb) - g.origin[2];
                    const double bx0 = (bx + rb) - g.origin[0], bx1 = (by + rb) - g.origin[1], bx2 = (bz + rb) - g.origin[2];
                    // f_Count_AABB_AABB_Intersection, CMC/ChCollisionUtilsBroadphase.cpp:114-140
                    ov = (an0 <= bx0) & (bn0 <= ax0) & (an1 <= bx1) & (bn1 <= ax1) & (an2 <= bx2) & (bn2 <= ax2);
                    // the deciding lines of sphere_sphere, ChNarrowphasePRIMS.cpp:52-60 (symmetric in the two shapes)
                    const d3 delta = mk3(bx, by, bz) - mk3(ax, ay, az);
                    const double dist2 = dot3(delta, delta);
                    const double radSum_s = (ar + br) + sep;
                    touch = !((dist2 >= radSum_s * radSum_s) | (dist2 < 1e-12));
                }
                const uint32_t ia = T.info[sa], ib = T.info[sb], fa = T.fam[sa], fb = T.fam[sb];
                const uint32_t both = ia | ib;
                // "can collide" and "not removed" were part of the pre-test (exact there: plain bits)
                const bool ok = ((both & 56u) == 56u) & (T.body[sa] != T.body[sb]) & ((both & 1u) != 0u) &
                                (((fa >> 16) & (fb & 0xffffu)) != 0u) & (((fb >> 16) & (fa & 0xffffu)) != 0u);
                bool hit = ov & ok;
                np += hit ? 1u : 0u;
                if (mode == MODE_CONTACTS)
                    hit = hit & touch;
                if (hit)
                    // (a 32-bit shared atomic is one instruction, the 64-bit one a compare-and-swap loop)
                    atomicOr(reinterpret_cast<uint32_t*>(&T.smask[sa]) + (hi >> 5), 1u << (hi & 31u));
            }
            __syncwarp();
        }
        // ---- pairs with one of the few non-spheres of a wall bin: queued one by one for k_pair_queue (a thread per
        // pair).  A sphere row queues itself against every non-sphere, a non-sphere row against the non-spheres after it.
        {
            unsigned long long mine = 0ull;
            if (own && bin_here && nn > 0) {
                const bool row_ns = (nsm >> i) & 1ull;
                mine = row_ns ? (nsm & ~((2ull << i) - 1ull)) : nsm;
                if (!row_ns) {
                    // a pair whose boxes do not overlap is in no list (f_Count_AABB_AABB_Intersection,
                    // CMC/ChCollisionUtilsBroadphase.cpp:114-140): most spheres of a wall bin do not reach the wall.
                    // Both boxes exactly as GenerateAABB + OffsetAABB leave them.
                    const double ar = T.c[3][t] + envelope;
                    const double an0 = (T.c[0][t] - ar) - g.origin[0], an1 = (T.c[1][t] - ar) - g.origin[1],
                                 an2 = (T.c[2][t] - ar) - g.origin[2];
                    const double ax0 = (T.c[0][t] + ar) - g.origin[0], ax1 = (T.c[1][t] + ar) - g.origin[1],
                                 ax2 = (T.c[2][t] + ar) - g.origin[2];
                    for (unsigned long long rest = mine; rest; rest &= rest - 1ull) {
                        const uint32_t a = (uint32_t)__ffsll((long long)rest) - 1u, q = w0 + a;
                        const double bx1 = __hiloint2double((int)T.body[q], (int)T.fam[q]);
                        const double bx2 = __hiloint2double((int)T.info[q], __float_as_int(T.rm[q]));
                        const bool ov = (an0 <= T.c[3][q]) & (T.c[0][q] <= ax0) & (an1 <= bx1) & (T.c[1][q] <= ax1) &
                                        (an2 <= bx2) & (T.c[2][q] <= ax2);
                        if (!ov)
                            mine &= ~(1ull << a);
                    }
                }
            }
            const uint32_t c = (uint32_t)__popcll(mine);
            if (__any_sync(0xffffffffu, c != 0u)) {
                const uint32_t incl = warp_incl_scan_bp(c, lane);
                uint32_t base = 0;
                if (lane == 31)
                    base = atomicAdd(&cnt->n_pairq, incl);
                base = __shfl_sync(0xffffffffu, base, 31) + incl - c;
                for (; mine; mine &= mine - 1ull) {
                    const uint32_t a = (uint32_t)__ffsll((long long)mine) - 1u;
                    if (base < cap_pairq)
                        pair_q[base] = make_uint2(s + min(i, a), s + max(i, a));
                    base++;
                }
            }
        }
        // ---- bins left as a whole to k_pair_masks_slow (a warp per bin)
        if (own && s == e && whole_slow) {
            if (k > 64)
                long_bins[atomicAdd(&cnt->n_long, 1u)] = s;  // k_long_sort / k_long_count / k_long_store
            else
                slow_bins[atomicAdd(&cnt->n_slow, 1u)] = s;
        }
        __syncthreads();
        // masks of the rows this tile owns and has decided
        if (own && !whole_slow) {
            const unsigned long long mk = T.smask[t];
            ent_mask[e] = mk;
            ent_cnt[e] = (uint32_t)__popcll(mk);
            if (CPLX)
                ent_cmask[e] = 0ull;
        }
    }
    // N_p (contacts mode; in pairs mode the store pass sets it from the scanned counts)
#pragma unroll
    for (int d = 16; d > 0; d >>= 1)
        np += __shfl_xor_sync(0xffffffffu, np, d);
    if (lane == 0 && mode == MODE_CONTACTS && np)
        atomicAdd(&cnt->n_pairs, np);
}

// k_pair_queue: the pairs of wall bins that involve a non-sphere, a thread per pair; the result joins the masks the tile
// kernel wrote for those rows (global atomics: two pairs of one row may be decided by different threads).
template <bool CPLX>
__global__ void __launch_bounds__(128) k_pair_queue(const ShapeMeta* __restrict__ meta,
                                                    const double* __restrict__ s_wpos,
                                                    const ShapeBox* __restrict__ s_box,
                                                    const double* __restrict__ s_wrot,
                                                    const double* __restrict__ s_dims,
                                                    const uint32_t* __restrict__ ent_shape,
                                                    const uint32_t* __restrict__ ent_bin,
                                                    const uint64_t* __restrict__ bin_start, double envelope,
                                                    int algorithm, int mode, const uint2* __restrict__ pair_q,
                                                    uint32_t cap_pairq, unsigned long long* __restrict__ ent_mask,
                                                    unsigned long long* __restrict__ ent_cmask,
                                                    uint32_t* __restrict__ ent_cnt, DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nq = cnt->n_pairq;
    if (nq > cap_pairq) {
        if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) {
            cnt->need_pairq = nq;
            __threadfence();
            atomicOr(&cnt->overflow, (uint32_t)OVF_PAIRQ);
        }
        return;
    }
    const GridParams g = load_grid(cnt);
    const double sep = 2 * envelope;
    uint32_t np = 0, ncx = 0;
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += gridDim.x * blockDim.x) {
        const uint2 it = pair_q[q];
        const uint32_t bin = ent_bin[it.x];
        const uint32_t s = (uint32_t)bin_start[bin];
        Rec A, B;
        load_rec(A, ent_shape[it.x], meta, s_wpos, s_box, g.origin, envelope);
        load_rec(B, ent_shape[it.y], meta, s_wpos, s_box, g.origin, envelope);
        const int dec = pair_decide<false>(A, B, g, bin, mode, sep, algorithm, s_wrot, s_dims, np);
        if (dec == 1) {
            atomicOr(&ent_mask[it.x], 1ull << (it.y - s));
            atomicAdd(&ent_cnt[it.x], 1u);
        } else if (dec == 2) {
            if (CPLX)
                atomicOr(&ent_cmask[it.x], 1ull << (it.y - s));
            else
                ncx++;
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        np += __shfl_xor_sync(0xffffffffu, np, d);
        ncx += __shfl_xor_sync(0xffffffffu, ncx, d);
    }
    if ((threadIdx.x & 31) == 0) {
        if (mode == MODE_CONTACTS && np)
            atomicAdd(&cnt->n_pairs, np);
        if (ncx)
            atomicAdd(&cnt->unsupported, ncx);
    }
}

// --------------------------------------------------------------------------------------------------------------
// k_pair_masks_slow: a warp per queued bin; the warp stages up to 64 of the bin's entries in its own shared-memory
// window: all pairs of a bin of <= 64 entries that the tile kernel left as a whole (many non-spheres, or MPR for
// everything).  Bins of more than 64 entries: k_long_sort / k_long_count / k_long_store.
// --------------------------------------------------------------------------------------------------------------
constexpr int SLOW_WIN = 64, SLOW_WARPS = 4;

template <bool CPLX>
__global__ void __launch_bounds__(32 * SLOW_WARPS) k_pair_masks_slow(
    const ShapeMeta* __restrict__ meta, const double* __restrict__ s_wpos, const ShapeBox* __restrict__ s_box,
    const double* __restrict__ s_wrot, const double* __restrict__ s_dims, const uint32_t* __restrict__ ent_tmp,
    uint32_t* ent_shape, const uint32_t* __restrict__ ent_bin, const uint64_t* __restrict__ bin_start,
    double envelope, int algorithm, int mode, uint64_t* ent_mask, uint64_t* ent_cmask, uint32_t* ent_cnt,
    const uint32_t* __restrict__ slow_bins, CplxArgs X, DevCounters* __restrict__ cnt) {
    __shared__ WindowT<SLOW_WIN> W[SLOW_WARPS];
    __shared__ unsigned long long smask[SLOW_WARPS][SLOW_WIN];
    __shared__ unsigned long long scmask[CPLX ? SLOW_WARPS : 1][SLOW_WIN];
    __shared__ uint16_t squeue[SLOW_WARPS][64];  // (lower entry << 8 | higher entry) of pairs that passed the filter
    if (cnt->overflow)
        return;
    const GridParams g = load_grid(cnt);
    const uint32_t nslow = cnt->n_slow;
    const double sep = 2 * envelope;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t np = 0, ncx = 0;
    for (uint32_t q = blockIdx.x * SLOW_WARPS + wid; q < nslow; q += gridDim.x * SLOW_WARPS) {
        const uint32_t bs = slow_bins[q];
        const uint32_t b = ent_bin[bs];
        const uint32_t bk = (uint32_t)bin_start[b + 1] - bs;
        int bx, by, bz;
        bin_coords(b, g, bx, by, bz);
        __syncwarp();
        for (uint32_t x = lane; x < min(bk, (uint32_t)SLOW_WIN); x += 32) {
            Rec R;
            load_rec_bin(R, ent_shape[bs + x], meta, s_wpos, s_box, g, bx, by, bz, envelope);
            win_put(W[wid], (int)x, R);
            smask[wid][x] = 0ull;
            if (CPLX)
                scmask[wid][x] = 0ull;
        }
        __syncwarp();
        if (bk <= 64) {
            // pairs spread evenly over the lanes in the cyclic order (i, i+d mod k); both records from the
            // window, ordered (lower entry, higher entry) for the shape functions
            const uint32_t h = (bk - 1) >> 1;
            const uint32_t full = bk * h;
            const uint32_t total = full + ((bk & 1u) ? 0u : (bk >> 1));
            const uint32_t M = ((1u << 20) + bk - 1) / bk;  // p / k == (p * M) >> 20 for p < 2048, k <= 64
            // Two phases, so that the lanes of the warp do the same thing: (1) the pair filter and the AABB test of 32
            // pairs, survivors compacted into a small queue; (2) whenever 32 of them are waiting (and at the end) the
            // narrowphase decision -- analytic single-contact tests against spheres, "complex" marking for the rest.
            // In a mixed bed two thirds of the entries are not spheres and one pair in seven survives the filter.
            uint32_t qn = 0;  // queued survivors (warp-uniform)
            auto drain = [&](uint32_t upto) {
                if (lane < upto) {
                    const uint32_t ac = squeue[wid][lane];
                    const uint32_t a = ac >> 8, c = ac & 0xffu;
                    Rec RA, RB;
                    win_get(W[wid], (int)a, RA);
                    win_get(W[wid], (int)c, RB);
                    np++;
                    int dec = 1;
                    if (mode != MODE_PAIRS) {
                        const int n = pair_hit(RA, RB, sep, algorithm, s_wrot, s_dims);
                        dec = n < 0 ? 2 : n;
                    }
                    if (dec == 1)
                        atomicOr(reinterpret_cast<uint32_t*>(&smask[wid][a]) + (c >> 5), 1u << (c & 31u));
                    else if (dec == 2) {
                        if (CPLX)
                            atomicOr(&scmask[CPLX ? wid : 0][a], 1ull << c);
                        else
                            ncx++;
                    }
                }
                __syncwarp();
                // what is left moves to the front
                const uint32_t rest = qn - upto;
                const uint32_t mv = lane < rest ? squeue[wid][upto + lane] : 0u;
                __syncwarp();
                if (lane < rest)
                    squeue[wid][lane] = (uint16_t)mv;
                __syncwarp();
                qn = rest;
            };
            for (uint32_t p0 = 0; p0 < total; p0 += 32) {
                const uint32_t p = p0 + lane;
                bool pass = false;
                uint32_t a = 0, c = 0;
                if (p < total) {
                    uint32_t i, j;
                    if (p < full) {
                        const uint32_t d = (p * M) >> 20;
                        i = p - d * bk;
                        j = i + d + 1;
                        if (j >= bk)
                            j -= bk;
                    } else {
                        i = p - full;
                        j = i + (bk >> 1);
                    }
                    a = min(i, j), c = max(i, j);
                    Rec RA, RB;
                    win_get(W[wid], (int)a, RA);
                    win_get(W[wid], (int)c, RB);
                    pass = !(RA.body == UINT_MAX || !(RA.info & 2u)) && pair_passes<true>(RA, RB, g, b);
                }
                const unsigned m = __ballot_sync(0xffffffffu, pass);
                if (pass)
                    squeue[wid][qn + (uint32_t)__popc(m & ((1u << lane) - 1u))] = (uint16_t)(a << 8 | c);
                qn += (uint32_t)__popc(m);
                __syncwarp();
                if (qn >= 32)
                    drain(32);
            }
            if (qn)
                drain(qn);
            __syncwarp();
            for (uint32_t x = lane; x < bk; x += 32) {
                const unsigned long long mk = smask[wid][x];
                ent_mask[bs + x] = mk;
                ent_cnt[bs + x] = (uint32_t)__popcll(mk);
                if (CPLX)
                    ent_cmask[bs + x] = scmask[CPLX ? wid : 0][x];
            }
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        np += __shfl_xor_sync(0xffffffffu, np, d);
        ncx += __shfl_xor_sync(0xffffffffu, ncx, d);
    }
    if (lane == 0) {
        if (mode == MODE_CONTACTS && np)
            atomicAdd(&cnt->n_pairs, np);
        if (ncx)
            atomicAdd(&cnt->unsupported, ncx);
    }
}

// --------------------------------------------------------------------------------------------------------------
// Complex pairs (multi-contact analytic tests and MPR).
//   k_cplx_collect  one thread per entry: reserves records and contact slots for the entry's complex candidates
//                   (atomic cursors; the order of the records is irrelevant, the final position of a contact is
//                   derived from the entry offsets, so the result is deterministic)
//   k_cplx_narrow   one thread per record: narrow_complex -> up to 8 contacts into the record's slots; adds the
//                   number found to the entry's output count
// --------------------------------------------------------------------------------------------------------------
// Narrowphase class of a complex pair: pairs of one class run the same function, so k_cplx_narrow walks the records
// class by class and its warps stay (mostly) convergent.  Has no effect on the results.
__device__ __forceinline__ int pair_class(int tA, int tB, int algorithm) {
    if (algorithm == 0)
        return (tA == CB200_SPHERE || tB == CB200_SPHERE) ? 0 : ((tA == CB200_BOX && tB == CB200_BOX) ? 2 : 3);
    const bool cA = tA == CB200_CAPSULE, cB = tB == CB200_CAPSULE, bA = tA == CB200_BOX, bB = tB == CB200_BOX;
    if (cA && cB)
        return 0;
    if ((bA && cB) || (cA && bB))
        return 1;
    if (bA && bB)
        return 2;
    return 3;
}

__global__ void __launch_bounds__(256) k_cplx_collect(const ShapeMeta* __restrict__ meta,
                                                      const uint32_t* __restrict__ ent_shape,
                                                      const uint32_t* __restrict__ ent_bin,
                                                      const uint64_t* __restrict__ bin_start,
                                                      const uint64_t* __restrict__ ent_cmask, int algorithm,
                                                      uint32_t* __restrict__ ent_cbase, uint32_t* __restrict__ rec,
                                                      uint32_t* __restrict__ cls_list, uint32_t cap_rec,
                                                      DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t ni = cnt->n_intersections;
    const int lane = threadIdx.x & 31;
    // A warp takes 32 entries.  Their complex candidates (the set bits of the 32 masks, 0..64 each) are numbered through
    // by a warp scan and then dealt out evenly to the lanes, 32 at a time -- an entry with 40 candidates does not keep
    // 31 lanes waiting.  Record t of the warp belongs to the entry whose scanned range holds t; inside an entry the
    // records follow the bit order, which is what k_pair_store relies on (ent_cbase + rank of the bit).
    for (uint32_t e0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; e0 < ni; e0 += gridDim.x * blockDim.x) {
        const uint32_t e = e0 + lane;
        const unsigned long long m = e < ni ? ent_cmask[e] : 0ull;
        if (!__any_sync(0xffffffffu, m != 0ull))
            continue;
        uint32_t s = 0, sA = 0;
        int tA = 0;
        if (m) {
            s = (uint32_t)bin_start[ent_bin[e]];
            sA = ent_shape[e];
            tA = (meta[sA].info >> 8) & 0xff;
        }
        const uint32_t count = (uint32_t)__popcll(m);
        uint32_t pc = count;  // inclusive warp scan
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t a = __shfl_up_sync(0xffffffffu, pc, d);
            if (lane >= d)
                pc += a;
        }
        const uint32_t total = __shfl_sync(0xffffffffu, pc, 31);
        uint32_t bc = 0;
        if (lane == 31)
            bc = atomicAdd(&cnt->n_complex, total);
        bc = __shfl_sync(0xffffffffu, bc, 31);
        if (m)
            ent_cbase[e] = bc + pc - count;
        // the partner of record t (the r-th set bit of its owner's mask), its class and its contact slots
        auto record = [&](uint32_t t, int& L, uint32_t& sB, int& c, uint32_t& slots) {
            // owner: the first lane whose inclusive count exceeds t (all lanes take part in the shuffles)
            L = 0;
#pragma unroll
            for (int step = 16; step > 0; step >>= 1) {
                const uint32_t v = __shfl_sync(0xffffffffu, pc, L + step - 1);
                if (v <= t)
                    L += step;
            }
            L = min(L, 31);
            const uint32_t incl = __shfl_sync(0xffffffffu, pc, L), cntL = __shfl_sync(0xffffffffu, count, L);
            const unsigned long long mL = __shfl_sync(0xffffffffu, m, L);
            const uint32_t sL = __shfl_sync(0xffffffffu, s, L);
            const int tAL = __shfl_sync(0xffffffffu, tA, L);
            c = 0, slots = 0, sB = 0;
            if (t < total) {
                const uint32_t r = t - (incl - cntL);
                const uint32_t lo = (uint32_t)mL, hi = (uint32_t)(mL >> 32), nlo = (uint32_t)__popc(lo);
                const int j = r < nlo ? (int)__fns(lo, 0, (int)r + 1) : 32 + (int)__fns(hi, 0, (int)(r - nlo) + 1);
                sB = ent_shape[sL + j];
                const int tB = (meta[sB].info >> 8) & 0xff;
                c = pair_class(tAL, tB, algorithm);
                slots = pair_slots(tAL, tB, algorithm);
            }
        };
        // pass 1: what the warp needs of the contact slots and of each class list -> one atomic each (any disjoint
        // range will do: k_pair_store finds the slots through the record, the order inside a class list is irrelevant)
        uint32_t my_slots = 0;
        unsigned long long my_cls = 0ull;  // 4 x 16 bit
        for (uint32_t t0 = 0; t0 < total; t0 += 32) {
            int L, c;
            uint32_t sB, slots;
            record(t0 + lane, L, sB, c, slots);
            if (t0 + lane < total) {
                my_slots += slots;
                my_cls += 1ull << (16 * c);
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            my_slots += __shfl_xor_sync(0xffffffffu, my_slots, d);
            my_cls += __shfl_xor_sync(0xffffffffu, my_cls, d);  // 32 x 64 records at most: no carry between the fields
        }
        uint32_t base_slot = 0, base_cls[4] = {0, 0, 0, 0};
        if (lane == 0) {
            base_slot = atomicAdd(&cnt->n_cslots, my_slots);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const uint32_t nq = (uint32_t)((my_cls >> (16 * q)) & 0xffff);
                if (nq)
                    base_cls[q] = atomicAdd(&cnt->n_cls[q], nq);
            }
        }
        base_slot = __shfl_sync(0xffffffffu, base_slot, 0);
#pragma unroll
        for (int q = 0; q < 4; q++)
            base_cls[q] = __shfl_sync(0xffffffffu, base_cls[q], 0);
        // pass 2: the records
        for (uint32_t t0 = 0; t0 < total; t0 += 32) {
            const uint32_t t = t0 + lane;
            const bool have = t < total;
            int L, c;
            uint32_t sB, slots;
            record(t, L, sB, c, slots);
            const uint32_t sAL = __shfl_sync(0xffffffffu, sA, L);
            uint32_t ps = slots;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t a = __shfl_up_sync(0xffffffffu, ps, d);
                if (lane >= d)
                    ps += a;
            }
            uint32_t k = 0;
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const unsigned mq = __ballot_sync(0xffffffffu, have && c == q);
                if (have && c == q)
                    k = base_cls[q] + (uint32_t)__popc(mq & ((1u << lane) - 1u));
                base_cls[q] += (uint32_t)__popc(mq);
            }
            const uint32_t base = bc + t;
            if (have && base < cap_rec) {
                reinterpret_cast<uint4*>(rec)[base] = make_uint4(e0 + (uint32_t)L, sAL, sB, base_slot + ps - slots);
                if (k < cap_rec)
                    cls_list[(size_t)c * cap_rec + k] = base;
            }
            base_slot += __shfl_sync(0xffffffffu, ps, 31);
        }
    }
}

// capacity check of the complex path: one thread, before the class kernels (which simply return on overflow)
__global__ void k_cplx_check(uint32_t cap_rec, uint32_t cap_slots, DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nrec = cnt->n_complex, nslots = cnt->n_cslots;
    if (nrec > cap_rec || nslots > cap_slots) {
        cnt->need_complex = nrec;
        cnt->need_cslots = nslots;
        __threadfence();
        atomicOr(&cnt->overflow, (uint32_t)OVF_COMPLEX);
    }
}

// One kernel per narrowphase class (pair_class): a kernel is allocated the registers of the heaviest function it can
// reach, so capsule-capsule pairs should not run at the occupancy box-box pairs force (15-axis SAT + clipping), and
// inside a class the warps stay convergent.  CLS 0..2 call their function directly when the analytic path is on
// (algorithm != MPR); CLS 3 and the all-MPR mode go through the general dispatcher.
template <int CLS>
__global__ void __launch_bounds__(128) k_cplx_narrow(const ShapeMeta* __restrict__ meta,
                                                     const double* __restrict__ s_wpos,
                                                     const double* __restrict__ s_wrot,
                                                     const double* __restrict__ s_dims,
                                                     const double* __restrict__ s_lrot,
                                                     const double* __restrict__ rot,
                                                     const double* __restrict__ s_verts, double envelope, int algorithm,
                                                     const uint32_t* __restrict__ rec,
                                                     const uint32_t* __restrict__ cls_list, uint32_t cap_rec,
                                                     uint32_t* __restrict__ cplx_n, double* __restrict__ cplx_out,
                                                     uint32_t* __restrict__ ent_cnt, DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    uint32_t bad = 0;
    const uint32_t ncls = cnt->n_cls[CLS];
    const double sep = 2 * envelope;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < ncls; k += gridDim.x * blockDim.x) {
        const uint32_t i = cls_list[(size_t)CLS * cap_rec + k];
        const uint4 r = reinterpret_cast<const uint4*>(rec)[i];
        const ShapeMeta mA = meta[r.y], mB = meta[r.z];
        const ShapeView A = make_view_full(r.y, mA, s_wpos, s_wrot, s_dims, s_lrot, rot, s_verts);
        const ShapeView B = make_view_full(r.z, mB, s_wpos, s_wrot, s_dims, s_lrot, rot, s_verts);
        Contact out[MAX_PAIR_CONTACTS];
        int n;
        if (CLS == 0 && algorithm != 0) {
            n = capsule_capsule(A.pos, A.rot, A.d0, A.d1, B.pos, B.rot, B.d0, B.d1, sep, out);
        } else if (CLS == 1 && algorithm != 0) {
            // one call site for both orders (two would be two divergent copies of the function in a warp)
            const bool swap = A.type != CB200_BOX;
            const ShapeView& X = swap ? B : A;
            const ShapeView& Y = swap ? A : B;
            n = box_capsule(X.pos, X.rot, mk3(X.d0, X.d1, X.d2), Y.pos, Y.rot, Y.d0, Y.d1, sep, out);
            if (swap)
                flip_contacts(out, n);
        } else if (CLS == 2 && algorithm != 0) {
            n = box_box(A.pos, A.rot, mk3(A.d0, A.d1, A.d2), B.pos, B.rot, mk3(B.d0, B.d1, B.d2), sep, out);
        } else {
            n = narrow_complex(A, B, algorithm, envelope, out);
        }
        if (n < 0) {
            bad++;
            n = 0;
        }
        const uint32_t slots = pair_slots(A.type, B.type, algorithm);
        if ((uint32_t)n > slots)
            n = (int)slots;
        cplx_n[i] = (uint32_t)n;
        for (int t = 0; t < n; t++) {
            double* o = cplx_out + 11 * (size_t)(r.w + t);
            o[0] = out[t].n.x, o[1] = out[t].n.y, o[2] = out[t].n.z;
            o[3] = out[t].p1.x, o[4] = out[t].p1.y, o[5] = out[t].p1.z;
            o[6] = out[t].p2.x, o[7] = out[t].p2.y, o[8] = out[t].p2.z;
            o[9] = out[t].depth, o[10] = out[t].erad;
        }
        if (n)
            atomicAdd(&ent_cnt[r.x], (uint32_t)n);
    }
    if (bad)
        atomicAdd(&cnt->unsupported, bad);
}

// Box-box pairs (class 2 under PRIMS / HYBRID) in two steps.  box_box has six branches of very different length
// (nothing, one corner contact, edge-edge, face-face clipping with up to 8 contacts, ...); run pair by pair the threads
// of a warp wait for each other (measured: 5.8 of 32 lanes active, 2.2 ms per step in the 518k mixed bed).
//   k_bb_classify  SAT + feature codes only (the same arithmetic as the head of box_box): separated pairs are done,
//                  the others are bucketed by branch (atomic cursors: the order inside a bucket is irrelevant, every
//                  record writes its own slots)
//   k_bb_contacts  walks the buckets branch by branch (padded to whole warps) and runs box_box
__global__ void __launch_bounds__(128) k_bb_classify(const ShapeMeta* __restrict__ meta,
                                                     const double* __restrict__ s_wpos,
                                                     const double* __restrict__ s_wrot,
                                                     const double* __restrict__ s_dims,
                                                     const double* __restrict__ s_lrot, const double* __restrict__ rot,
                                                     const double* __restrict__ s_verts, double envelope,
                                                     const uint32_t* __restrict__ rec,
                                                     const uint32_t* __restrict__ cls_list, uint32_t cap_rec,
                                                     uint32_t* __restrict__ cplx_n, uint32_t* __restrict__ bb_list,
                                                     DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t ncls = cnt->n_cls[2];
    const double sep = 2 * envelope;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < ncls; k += gridDim.x * blockDim.x) {
        const uint32_t i = cls_list[(size_t)2 * cap_rec + k];
        const uint4 r = reinterpret_cast<const uint4*>(rec)[i];
        const ShapeView A = make_view_full(r.y, meta[r.y], s_wpos, s_wrot, s_dims, s_lrot, rot, s_verts);
        const ShapeView B = make_view_full(r.z, meta[r.z], s_wpos, s_wrot, s_dims, s_lrot, rot, s_verts);
        const int key = box_box_case(A.pos, A.rot, mk3(A.d0, A.d1, A.d2), B.pos, B.rot, mk3(B.d0, B.d1, B.d2), sep);
        if (key == 0 || key == 6)
            cplx_n[i] = 0;
        else
            bb_list[(size_t)(key - 1) * cap_rec + atomicAdd(&cnt->n_bb[key - 1], 1u)] = i;
    }
}

__global__ void __launch_bounds__(128) k_bb_contacts(const ShapeMeta* __restrict__ meta,
                                                     const double* __restrict__ s_wpos,
                                                     const double* __restrict__ s_wrot,
                                                     const double* __restrict__ s_dims,
                                                     const double* __restrict__ s_lrot, const double* __restrict__ rot,
                                                     const double* __restrict__ s_verts, double envelope,
                                                     const uint32_t* __restrict__ rec,
                                                     const uint32_t* __restrict__ bb_list, uint32_t cap_rec,
                                                     uint32_t* __restrict__ cplx_n, double* __restrict__ cplx_out,
                                                     uint32_t* __restrict__ ent_cnt, DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const double sep = 2 * envelope;
    uint32_t start[6];
    start[0] = 0;
#pragma unroll
    for (int q = 0; q < 5; q++)
        start[q + 1] = start[q] + ((cnt->n_bb[q] + 31u) & ~31u);
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < start[5]; t += gridDim.x * blockDim.x) {
        int q = 0;
#pragma unroll
        for (int u = 1; u < 5; u++)
            q += t >= start[u] ? 1 : 0;
        const uint32_t k = t - start[q];
        if (k >= cnt->n_bb[q])
            continue;
        const uint32_t i = bb_list[(size_t)q * cap_rec + k];
        const uint4 r = reinterpret_cast<const uint4*>(rec)[i];
        const ShapeView A = make_view_full(r.y, meta[r.y], s_wpos, s_wrot, s_dims, s_lrot, rot, s_verts);
        const ShapeView B = make_view_full(r.z, meta[r.z], s_wpos, s_wrot, s_dims, s_lrot, rot, s_verts);
        Contact out[MAX_PAIR_CONTACTS];
        int n = box_box(A.pos, A.rot, mk3(A.d0, A.d1, A.d2), B.pos, B.rot, mk3(B.d0, B.d1, B.d2), sep, out);
        if (n > 8)
            n = 8;  // pair_slots(box, box)
        cplx_n[i] = (uint32_t)n;
        for (int c = 0; c < n; c++) {
            double* o = cplx_out + 11 * (size_t)(r.w + c);
            o[0] = out[c].n.x, o[1] = out[c].n.y, o[2] = out[c].n.z;
            o[3] = out[c].p1.x, o[4] = out[c].p1.y, o[5] = out[c].p1.z;
            o[6] = out[c].p2.x, o[7] = out[c].p2.y, o[8] = out[c].p2.z;
            o[9] = out[c].depth, o[10] = out[c].erad;
        }
        if (n)
            atomicAdd(&ent_cnt[r.x], (uint32_t)n);
    }
}

// k_pair_store: one thread per entry; turns masks (or, for long bins, a re-walk of the row) into 64-bit shape-pair
// keys at their final, compacted positions.  With CPLX the contacts of complex pairs are interleaved in (j, contact)
// order -- the order the reference's slot layout + stable compaction leaves -- and c_src points at their records.
template <bool CPLX>
__global__ void __launch_bounds__(256) k_pair_store(const ShapeMeta* __restrict__ meta,
                                                    const double* __restrict__ s_wpos,
                                                    const ShapeBox* __restrict__ s_box,
                                                    const double* __restrict__ s_wrot,
                                                    const double* __restrict__ s_dims,
                                                    const uint32_t* __restrict__ ent_shape,
                                                    const uint32_t* __restrict__ ent_bin,
                                                    const uint64_t* __restrict__ bin_start, double envelope,
                                                    int algorithm, int mode, const uint64_t* __restrict__ ent_mask,
                                                    const uint64_t* __restrict__ ent_mask_hi,
                                                    const uint32_t* __restrict__ ent_off, int64_t* __restrict__ out,
                                                    uint32_t cap, const uint64_t* __restrict__ ent_cmask,
                                                    const uint32_t* __restrict__ ent_cbase,
                                                    const uint32_t* __restrict__ rec,
                                                    const uint32_t* __restrict__ cplx_n, uint32_t* __restrict__ c_src,
                                                    int32_t* __restrict__ c_bids, uint32_t* __restrict__ c_mats,
                                                    CplxArgs X, DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t ni = cnt->n_intersections;
    const uint32_t total = ent_off[ni];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (mode == MODE_CONTACTS) {
            cnt->n_contacts = total;
            if (total > cap) {
                cnt->need_contacts = total;
                atomicOr(&cnt->overflow, (uint32_t)OVF_CONTACTS);
            }
        } else {
            cnt->n_pairs = total;
            if (total > cap) {
                cnt->need_pairs = total;
                atomicOr(&cnt->overflow, (uint32_t)OVF_PAIRS);
            }
        }
    }
    if (total > cap)
        return;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < ni; e += gridDim.x * blockDim.x) {
        uint32_t o = ent_off[e];
        if (ent_off[e + 1] == o)
            continue;
        const uint32_t bin = ent_bin[e];
        const uint32_t s = (uint32_t)bin_start[bin], k = (uint32_t)bin_start[bin + 1] - s;
        const uint32_t idA = ent_shape[e];
        const int64_t hiA = (int64_t)idA << 32;
        // contacts mode: body and material ids of every contact are written beside its key (bids, per-contact material
        // pair), so that the force kernel can start its body gathers without waiting for the shape records
        const bool ids = mode == MODE_CONTACTS && c_bids != nullptr;
        ShapeMeta mA = {};
        if (ids)
            mA = meta[idA];
        if (k <= 64) {
            unsigned long long m = ent_mask[e];
            if (CPLX && mode == MODE_CONTACTS) {
                const unsigned long long cm = ent_cmask[e];
                unsigned long long u = m | cm;
                uint32_t ci = cm ? ent_cbase[e] : 0;
                while (u) {
                    const int j = __ffsll((long long)u) - 1;
                    u &= u - 1;
                    const uint32_t idB = ent_shape[s + j];
                    const int64_t key = hiA | (int64_t)idB;
                    const ShapeMeta mB = meta[idB];
                    const int2 bb = make_int2((int)mA.body, (int)mB.body);
                    const uint32_t mm = (mA.info >> 16) | (mB.info >> 16) << 16;
                    if ((cm >> j) & 1ull) {
                        const uint32_t n = cplx_n[ci];
                        const uint32_t slot = rec[4 * (size_t)ci + 3];
                        for (uint32_t t = 0; t < n; t++) {
                            out[o] = key;
                            reinterpret_cast<int2*>(c_bids)[o] = bb;
                            c_mats[o] = mm;
                            c_src[o++] = slot + t + 1;
                        }
                        ci++;
                    } else {
                        out[o] = key;
                        reinterpret_cast<int2*>(c_bids)[o] = bb;
                        c_mats[o] = mm;
                        c_src[o++] = 0;
                    }
                }
            } else {
                while (m) {
                    const int j = __ffsll((long long)m) - 1;
                    m &= m - 1;
                    const uint32_t idB = ent_shape[s + j];
                    if (ids) {
                        const ShapeMeta mB = meta[idB];
                        reinterpret_cast<int2*>(c_bids)[o] = make_int2((int)mA.body, (int)mB.body);
                        c_mats[o] = (mA.info >> 16) | (mB.info >> 16) << 16;
                    }
                    out[o++] = hiA | (int64_t)idB;
                }
            }
        }
        else if (!CPLX && k <= 128) {
            // bins of 65..128 entries: k_long_rows<false> left 128-bit masks (no second walk over the pairs)
            for (int half = 0; half < 2; half++) {
                unsigned long long m = half ? ent_mask_hi[e] : ent_mask[e];
                while (m) {
                    const int j = __ffsll((long long)m) - 1 + 64 * half;
                    m &= m - 1;
                    const uint32_t idB = ent_shape[s + j];
                    if (ids) {
                        const ShapeMeta mB = meta[idB];
                        reinterpret_cast<int2*>(c_bids)[o] = make_int2((int)mA.body, (int)mB.body);
                        c_mats[o] = (mA.info >> 16) | (mB.info >> 16) << 16;
                    }
                    out[o++] = hiA | (int64_t)idB;
                }
            }
        }
        // rows of longer bins (and of all bins of more than 64 entries when complex pairs are in play): k_long_rows<true>
    }
}

// --------------------------------------------------------------------------------------------------------------
// Bins with more than 64 entries (coarse grids; a dense 4 x 4 x 4 bin plus its walls): no masks.  Three small kernels,
// all of which leave at once when there is no such bin:
//   k_long_sort   rank sort of the bin's entries by shape id (a CTA per bin)
//   k_long_count  outputs per row: a warp per row, lanes across the partners (rows of one bin spread over many CTAs)
//   k_long_store  the same walk again after the scan; a lane's slot = row offset + hits of the lanes before it, i.e.
//                 the reference's (shape A, shape B) order
// --------------------------------------------------------------------------------------------------------------
// Bins in flight per launch of the three kernels.  A bed that has compacted for seconds holds thousands of bins of 65 - 80
// entries at a bin edge of 2.5 diameters (not the handful of the first second): with 64 bins in flight these kernels
// grew to a quarter of the step.  CTAs beyond the number of long bins leave at once.
constexpr int LONG_GRID_Y = 592;
constexpr int LONG_STAGE = 128;  // entries of a long bin kept in shared memory by k_long_rows (12 KB)

__global__ void __launch_bounds__(256) k_long_sort(const uint32_t* __restrict__ ent_tmp, uint32_t* __restrict__ ent_shape,
                                                   const uint32_t* __restrict__ ent_bin,
                                                   const uint64_t* __restrict__ bin_start,
                                                   const uint32_t* __restrict__ long_bins,
                                                   const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nl = cnt->n_long;
    for (uint32_t q = blockIdx.x; q < nl; q += gridDim.x) {
        const uint32_t s = long_bins[q];
        const uint32_t k = (uint32_t)bin_start[ent_bin[s] + 1] - s;
        for (uint32_t x = threadIdx.x; x < k; x += blockDim.x) {
            const uint32_t id = ent_tmp[s + x];
            uint32_t rank = 0;
            for (uint32_t y = 0; y < k; y++)
                rank += ent_tmp[s + y] < id ? 1u : 0u;
            ent_shape[s + rank] = id;
        }
    }
}

// decision for the lane's partner j of row A, and the number of outputs it gives (contacts, or 1 in pair-list mode)
template <bool CPLX>
__device__ __forceinline__ int long_pair(const Rec& A, uint32_t idB, const Rec* staged, const GridParams& g,
                                         uint32_t bin, int mode,
                                         double envelope, int algorithm, const ShapeMeta* __restrict__ meta,
                                         const double* __restrict__ s_wpos, const ShapeBox* __restrict__ s_box,
                                         const double* __restrict__ s_wrot, const double* __restrict__ s_dims,
                                         const CplxArgs& X, Contact* tmp, uint32_t& n, uint32_t& np, uint32_t& ncx) {
    Rec B;
    if (staged)
        B = *staged;  // the bin's records sit in shared memory (k_long_rows)
    else
        load_rec(B, idB, meta, s_wpos, s_box, g.origin, envelope);
    const int dec = pair_decide<false>(A, B, g, bin, mode, 2 * envelope, algorithm, s_wrot, s_dims, np);
    n = dec == 1 ? 1u : 0u;
    if (dec == 2) {
        if (!CPLX)
            ncx++;
        else if (mode != MODE_CONTACTS)
            n = 1;  // pair-list mode: a candidate pair like any other
        else  // long bins keep no records: a complex pair is evaluated in place, by the count and by the store pass
            n = (uint32_t)narrow_complex_ids(A.id, idB, meta, s_wpos, s_wrot, s_dims, X, algorithm, envelope, tmp);
    }
    return dec;
}

template <bool CPLX, bool STORE>
__global__ void __launch_bounds__(128) k_long_rows(const ShapeMeta* __restrict__ meta, const double* __restrict__ s_wpos,
                                                   const ShapeBox* __restrict__ s_box,
                                                   const double* __restrict__ s_wrot,
                                                   const double* __restrict__ s_dims,
                                                   const uint32_t* __restrict__ ent_shape,
                                                   const uint32_t* __restrict__ ent_bin,
                                                   const uint64_t* __restrict__ bin_start, double envelope,
                                                   int algorithm, int mode, uint64_t* __restrict__ ent_mask,
                                                   uint64_t* __restrict__ ent_mask_hi,
                                                   uint64_t* __restrict__ ent_cmask, uint32_t* __restrict__ ent_cnt,
                                                   const uint32_t* __restrict__ ent_off, int64_t* __restrict__ out,
                                                   uint32_t cap, uint32_t* __restrict__ c_src,
                                                   int32_t* __restrict__ c_bids, uint32_t* __restrict__ c_mats,
                                                   const uint32_t* __restrict__ long_bins, CplxArgs X,
                                                   DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nl = cnt->n_long;
    if (nl == 0)
        return;
    if (STORE && ent_off[cnt->n_intersections] > cap)
        return;  // k_pair_store raises the flag
    const GridParams g = load_grid(cnt);
    const int lane = threadIdx.x & 31;
    const uint32_t wpb = blockDim.x >> 5;
    uint32_t np = 0, ncx = 0;
    // A bin of up to LONG_STAGE entries is fetched once per CTA into shared memory (a compacted bed has thousands of
    // bins of 65 - 80 entries: a dependent gather per partner was most of these kernels); longer bins go through global.
    __shared__ Rec srec[LONG_STAGE];
    for (uint32_t q = blockIdx.y; q < nl; q += gridDim.y) {
        const uint32_t s = long_bins[q];
        const uint32_t bin = ent_bin[s];
        const uint32_t k = (uint32_t)bin_start[bin + 1] - s;
        const bool staged = k <= LONG_STAGE;
        const bool masked = !CPLX && k <= 128;  // the count pass records 128-bit masks, k_pair_store writes the keys
        if (STORE && masked)
            continue;
        __syncthreads();  // the previous bin's readers are done
        if (staged)
            for (uint32_t x = threadIdx.x; x < k; x += blockDim.x)
                load_rec(srec[x], ent_shape[s + x], meta, s_wpos, s_box, g.origin, envelope);
        __syncthreads();
        for (uint32_t i = blockIdx.x * wpb + (threadIdx.x >> 5); i < k; i += gridDim.x * wpb) {
            const uint32_t e = s + i;
            uint32_t o = 0;
            if (STORE) {
                o = ent_off[e];
                if (ent_off[e + 1] == o)
                    continue;
            }
            Rec A;
            if (staged)
                A = srec[i];
            else
                load_rec(A, ent_shape[e], meta, s_wpos, s_box, g.origin, envelope);
            const int64_t hiA = (int64_t)A.id << 32;
            uint32_t hits = 0;
            unsigned long long mlo = 0ull, mhi = 0ull;
            for (uint32_t j0 = i + 1; j0 < k; j0 += 32) {
                const uint32_t j = j0 + lane;
                Contact tmp[CPLX ? MAX_PAIR_CONTACTS : 1];
                uint32_t n = 0, idB = 0;
                int dec = 0;
                if (j < k) {
                    idB = staged ? srec[j].id : ent_shape[s + j];
                    dec = long_pair<CPLX>(A, idB, staged ? &srec[j] : nullptr, g, bin, mode, envelope, algorithm, meta,
                                          s_wpos, s_box, s_wrot, s_dims, X, tmp, n, np, ncx);
                }
                uint32_t incl = 0;
                if (!STORE && masked) {
                    // (n is 0 or 1 here: one vote instead of a scan)  partners j0 .. j0+31 of this trip as bits
                    // j0 .. j0+31 of the 128-bit mask
                    const unsigned long long w = (unsigned long long)__ballot_sync(0xffffffffu, n != 0);
                    hits += (uint32_t)__popcll(w);
                    const uint32_t sh = j0 & 63u;
                    if (j0 < 64) {
                        mlo |= w << sh;
                        if (sh > 32)
                            mhi |= w >> (64 - sh);
                    } else {
                        mhi |= w << sh;
                    }
                } else {
                    incl = warp_incl_scan_bp(n, lane);
                    hits += __shfl_sync(0xffffffffu, incl, 31);
                }
                if (STORE && n) {
                    uint32_t p = o + incl - n;
                    const int64_t key = hiA | (int64_t)idB;
                    if (mode == MODE_CONTACTS && c_bids) {
                        const ShapeMeta mB = meta[idB];
                        for (uint32_t t = 0; t < n; t++) {
                            reinterpret_cast<int2*>(c_bids)[p + t] = make_int2((int)A.body, (int)mB.body);
                            c_mats[p + t] = (A.info >> 16) | (mB.info >> 16) << 16;
                        }
                    }
                    if (dec == 1 || mode != MODE_CONTACTS) {
                        if (CPLX && mode == MODE_CONTACTS)
                            c_src[p] = 0;
                        out[p] = key;
                    } else if (CPLX) {
                        // contacts of a complex pair go to freshly reserved slots
                        const uint32_t slot = atomicAdd(&cnt->n_cslots, n);
                        if (slot + n > X.cap_slots) {
                            atomicMax(&cnt->need_cslots, slot + n);
                            atomicMax(&cnt->need_complex, cnt->n_complex);
                            atomicOr(&cnt->overflow, (uint32_t)OVF_COMPLEX);
                        } else {
                            for (uint32_t t = 0; t < n; t++) {
                                double* qd = X.cplx_out + 11 * (size_t)(slot + t);
                                qd[0] = tmp[t].n.x, qd[1] = tmp[t].n.y, qd[2] = tmp[t].n.z;
                                qd[3] = tmp[t].p1.x, qd[4] = tmp[t].p1.y, qd[5] = tmp[t].p1.z;
                                qd[6] = tmp[t].p2.x, qd[7] = tmp[t].p2.y, qd[8] = tmp[t].p2.z;
                                qd[9] = tmp[t].depth, qd[10] = tmp[t].erad;
                                out[p] = key;
                                c_src[p++] = slot + t + 1;
                            }
                        }
                    }
                }
                if (STORE)
                    o += __shfl_sync(0xffffffffu, incl, 31);
            }
            if (!STORE && lane == 0) {
                ent_mask[e] = mlo;
                if (masked)
                    ent_mask_hi[e] = mhi;
                ent_cnt[e] = hits;
                if (CPLX)
                    ent_cmask[e] = 0ull;
            }
        }
    }
    if (!STORE) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            np += __shfl_xor_sync(0xffffffffu, np, d);
            ncx += __shfl_xor_sync(0xffffffffu, ncx, d);
        }
        if (lane == 0) {
            if (mode == MODE_CONTACTS && np)
                atomicAdd(&cnt->n_pairs, np);
            if (ncx)
                atomicAdd(&cnt->unsupported, ncx);
        }
    }
}

// --------------------------------------------------------------------------------------------------------------
// k_contact_gen: one thread per contact key -> the contact record (DispatchPRIMS on the compacted list).
// --------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_contact_gen(const ShapeMeta* __restrict__ meta,
                                                     const double* __restrict__ s_wpos,
                                                     const double* __restrict__ s_wrot,
                                                     const double* __restrict__ s_dims,
                                                     const int64_t* __restrict__ c_shapes, double envelope,
                                                     double* __restrict__ c_norm, double* __restrict__ c_pta,
                                                     double* __restrict__ c_ptb, double* __restrict__ c_depth,
                                                     double* __restrict__ c_erad, int32_t* __restrict__ c_bids,
                                                     uint32_t* __restrict__ c_mats,
                                                     const uint32_t* __restrict__ c_src,
                                                     const double* __restrict__ cplx_out,
                                                     const DevCounters* __restrict__ cnt) {
    if (cnt->overflow)
        return;
    const uint32_t nc = cnt->n_contacts;
    const double sep = 2 * envelope;
    for (uint32_t o = blockIdx.x * blockDim.x + threadIdx.x; o < nc; o += gridDim.x * blockDim.x) {
        const int64_t key = c_shapes[o];
        const uint32_t sA = (uint32_t)(key >> 32), sB = (uint32_t)(key & 0xffffffff);
        const ShapeMeta mA = meta[sA], mB = meta[sB];
        const double4 wA = reinterpret_cast<const double4*>(s_wpos)[sA];
        const double4 wB = reinterpret_cast<const double4*>(s_wpos)[sB];
        const int tA = (mA.info >> 8) & 0xff, tB = (mB.info >> 8) & 0xff;
        Contact c;
        const uint32_t src = c_src ? c_src[o] : 0u;
        if (src) {
            const double* q = cplx_out + 11 * (size_t)(src - 1);
            c.n = mk3(q[0], q[1], q[2]), c.p1 = mk3(q[3], q[4], q[5]), c.p2 = mk3(q[6], q[7], q[8]);
            c.depth = q[9], c.erad = q[10];
        } else if (tA == CB200_SPHERE && tB == CB200_SPHERE)
            sphere_sphere(mk3(wA.x, wA.y, wA.z), wA.w, mk3(wB.x, wB.y, wB.z), wB.w, sep, c);
        else
            prims_single(make_view(sA, tA, wA, s_wrot, s_dims), make_view(sB, tB, wB, s_wrot, s_dims), sep, c);
        c_norm[3 * o] = c.n.x, c_norm[3 * o + 1] = c.n.y, c_norm[3 * o + 2] = c.n.z;
        c_pta[3 * o] = c.p1.x, c_pta[3 * o + 1] = c.p1.y, c_pta[3 * o + 2] = c.p1.z;
        c_ptb[3 * o] = c.p2.x, c_ptb[3 * o + 1] = c.p2.y, c_ptb[3 * o + 2] = c.p2.z;
        c_depth[o] = c.depth;
        c_erad[o] = c.erad;
        reinterpret_cast<int2*>(c_bids)[o] = make_int2((int)mA.body, (int)mB.body);
        c_mats[o] = (mA.info >> 16) | (mB.info >> 16) << 16;
    }
}

// --------------------------------------------------------------------------------------------------------------

// --------------------------------------------------------------------------------------------------------------
// host side: enqueue only, no synchronisation
// --------------------------------------------------------------------------------------------------------------
template <bool CPLX>
static int enqueue_pair_pass_t(System* S, int mode, int64_t* out, uint32_t cap) {
    const double env = S->set.collision_envelope;
    const int algorithm = S->set.narrowphase_algorithm;
    // the counters of this pass: cleared by k_reset_counters at the start of a collision pass (broadphase.cu); a second
    // pass over the same bins (pair list download) clears them itself
    const bool clean = S->pair_cnt_clean;
    S->pair_cnt_clean = false;
    if (mode == MODE_CONTACTS && !clean)
        CB_CUDA(cudaMemsetAsync(&S->d_cnt->n_pairs, 0, 4, S->stream));
    const int gm = grid_for(S, S->cap_entries, PAIR_TILE, 4);  // 4 resident CTAs per SM, tiles grid-stride
    if (!clean)
        CB_CUDA(cudaMemsetAsync(&S->d_cnt->n_slow, 0, 12, S->stream));  // n_slow, n_long, n_pairq
    const int gs = S->num_sms * 8;
    CplxArgs X = {};
    X.s_lrot = S->s_lrot.p, X.rot = S->rot.p, X.s_verts = S->s_verts.p, X.cplx_out = S->cplx_out.p;
    X.cap_slots = S->cap_cslots;
    uint64_t* cmask = CPLX ? S->ent_cmask.p : nullptr;
    if (CPLX && !clean) {
        CB_CUDA(cudaMemsetAsync(&S->d_cnt->n_complex, 0, 4, S->stream));
        CB_CUDA(cudaMemsetAsync(&S->d_cnt->n_cslots, 0, 4, S->stream));
        CB_CUDA(cudaMemsetAsync(&S->d_cnt->n_cls[0], 0, 16 + 20, S->stream));  // n_cls, n_bb
    }
    CB_TRY(timer_start(S, "k_pair_masks"));
    k_pair_masks<CPLX><<<gm, PAIR_THREADS, 0, S->stream>>>(
        S->s_meta.p, S->s_wpos.p, S->s_box.p, S->ent_tmp.p, S->ent_bin.p, S->bin_start.p, env, algorithm, mode,
        S->ent_shape.p, S->ent_mask.p, cmask, S->ent_off.p, S->slow_bins.p, S->long_bins.p,
        reinterpret_cast<uint2*>(S->pair_q.p), S->cap_pairq, S->d_cnt);
    CB_TRY(timer_stop(S, "k_pair_masks"));
    // three independent branches consume the queues the tile kernel filled (they touch disjoint bins, or only add
    // bits with atomics): wall-bin pairs here, whole slow bins and long bins on the side streams
    cudaStream_t s_slow, s_long;
    CB_TRY(aux_fork(S, 0, &s_slow));
    CB_TRY(aux_fork(S, 1, &s_long));
    CB_TRY(timer_start(S, "k_pair_queue"));
    k_pair_queue<CPLX><<<grid_for(S, S->cap_pairq, 128, 8), 128, 0, S->stream>>>(
        S->s_meta.p, S->s_wpos.p, S->s_box.p, S->s_wrot.p, S->s_dims.p, S->ent_shape.p, S->ent_bin.p, S->bin_start.p,
        env, algorithm, mode, reinterpret_cast<const uint2*>(S->pair_q.p), S->cap_pairq,
        reinterpret_cast<unsigned long long*>(S->ent_mask.p), reinterpret_cast<unsigned long long*>(cmask),
        S->ent_off.p, S->d_cnt);
    CB_TRY(timer_stop(S, "k_pair_queue"));
    k_pair_masks_slow<CPLX><<<gs, 32 * SLOW_WARPS, 0, s_slow>>>(
        S->s_meta.p, S->s_wpos.p, S->s_box.p, S->s_wrot.p, S->s_dims.p, S->ent_tmp.p, S->ent_shape.p, S->ent_bin.p,
        S->bin_start.p, env, algorithm, mode, S->ent_mask.p, cmask, S->ent_off.p, S->slow_bins.p, X, S->d_cnt);
    k_long_sort<<<LONG_GRID_Y, 256, 0, s_long>>>(S->ent_tmp.p, S->ent_shape.p, S->ent_bin.p, S->bin_start.p, S->long_bins.p,
                                        S->d_cnt);
    k_long_rows<CPLX, false><<<dim3(1, 8 * LONG_GRID_Y), 128, 0, s_long>>>(
        S->s_meta.p, S->s_wpos.p, S->s_box.p, S->s_wrot.p, S->s_dims.p, S->ent_shape.p, S->ent_bin.p, S->bin_start.p,
        env, algorithm, mode, S->ent_mask.p, S->ent_mask_hi.p, cmask, S->ent_off.p, nullptr, nullptr, 0, nullptr,
        nullptr, nullptr,
        S->long_bins.p, X, S->d_cnt);
    CB_TRY(aux_join(S, 0));
    CB_TRY(aux_join(S, 1));
    S->launches += 5;
    if (CPLX) {
        k_cplx_collect<<<grid_for(S, S->cap_entries, 256), 256, 0, S->stream>>>(
            S->s_meta.p, S->ent_shape.p, S->ent_bin.p, S->bin_start.p, S->ent_cmask.p, algorithm, S->ent_cbase.p,
            S->cplx_rec.p, S->cplx_cls.p, S->cap_complex, S->d_cnt);
        k_cplx_check<<<1, 1, 0, S->stream>>>(S->cap_complex, S->cap_cslots, S->d_cnt);
#define CPLX_NARROW(CLS, STREAM)                                                                                      \
    k_cplx_narrow<CLS><<<grid_for(S, S->cap_complex, 128), 128, 0, STREAM>>>(                                         \
        S->s_meta.p, S->s_wpos.p, S->s_wrot.p, S->s_dims.p, S->s_lrot.p, S->rot.p, S->s_verts.p, env, algorithm,      \
        S->cplx_rec.p, S->cplx_cls.p, S->cap_complex, S->cplx_n.p, S->cplx_out.p, S->ent_off.p, S->d_cnt)
        CPLX_NARROW(0, S->stream);
        CPLX_NARROW(1, S->stream);
        if (algorithm != 0) {
            k_bb_classify<<<grid_for(S, S->cap_complex, 128), 128, 0, S->stream>>>(
                S->s_meta.p, S->s_wpos.p, S->s_wrot.p, S->s_dims.p, S->s_lrot.p, S->rot.p, S->s_verts.p, env,
                S->cplx_rec.p, S->cplx_cls.p, S->cap_complex, S->cplx_n.p, S->cplx_bb.p, S->d_cnt);
            k_bb_contacts<<<grid_for(S, S->cap_complex, 128), 128, 0, S->stream>>>(
                S->s_meta.p, S->s_wpos.p, S->s_wrot.p, S->s_dims.p, S->s_lrot.p, S->rot.p, S->s_verts.p, env,
                S->cplx_rec.p, S->cplx_bb.p, S->cap_complex, S->cplx_n.p, S->cplx_out.p, S->ent_off.p, S->d_cnt);
            S->launches++;
        } else {
            CPLX_NARROW(2, S->stream);
        }
        CPLX_NARROW(3, S->stream);
#undef CPLX_NARROW
        S->launches += 6;
    }
    CB_TRY(scan_u32(S, S->ent_off.p, S->ent_off.p, &S->d_cnt->n_intersections, S->cap_entries));
    CB_TRY(aux_fork(S, 1, &s_long));  // the rows of long bins are stored beside the others
    CB_TRY(timer_start(S, "k_pair_store"));
    k_pair_store<CPLX><<<grid_for(S, S->cap_entries, 256), 256, 0, S->stream>>>(
        S->s_meta.p, S->s_wpos.p, S->s_box.p, S->s_wrot.p, S->s_dims.p, S->ent_shape.p, S->ent_bin.p, S->bin_start.p,
        env, algorithm, mode, S->ent_mask.p, S->ent_mask_hi.p, S->ent_off.p, out, cap, cmask,
        CPLX ? S->ent_cbase.p : nullptr,
        CPLX ? S->cplx_rec.p : nullptr, CPLX ? S->cplx_n.p : nullptr, CPLX ? S->c_src.p : nullptr, S->c_bids.p,
        S->c_mats.p, X, S->d_cnt);
    CB_TRY(timer_stop(S, "k_pair_store"));
    k_long_rows<CPLX, true><<<dim3(8, LONG_GRID_Y), 128, 0, s_long>>>(
        S->s_meta.p, S->s_wpos.p, S->s_box.p, S->s_wrot.p, S->s_dims.p, S->ent_shape.p, S->ent_bin.p, S->bin_start.p,
        env, algorithm, mode, nullptr, nullptr, nullptr, nullptr, S->ent_off.p, out, cap, CPLX ? S->c_src.p : nullptr,
        S->c_bids.p, S->c_mats.p, S->long_bins.p, X, S->d_cnt);
    CB_TRY(aux_join(S, 1));
    S->launches += 2;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

static int enqueue_pair_pass(System* S, int mode, int64_t* out, uint32_t cap) {
    if (mode == MODE_CONTACTS && use_complex(S))
        return enqueue_pair_pass_t<true>(S, mode, out, cap);
    return enqueue_pair_pass_t<false>(S, mode, out, cap);
}

// k_contact_gen over the current key list: the contact records (CMC/ChCollisionData.h:159-183).  The shape frames
// it reads (s_wpos, s_wrot) are those of the collision stage that made the keys and stay untouched until the next one.
int enqueue_contact_gen(System* S) {
    CB_TRY(timer_start(S, "k_contact_gen"));
    k_contact_gen<<<grid_for(S, S->cap_contacts, 256), 256, 0, S->stream>>>(
        S->s_meta.p, S->s_wpos.p, S->s_wrot.p, S->s_dims.p, S->c_shapes.p, S->set.collision_envelope, S->c_norm.p,
        S->c_pta.p, S->c_ptb.p, S->c_depth.p, S->c_erad.p, S->c_bids.p, S->c_mats.p,
        use_complex(S) ? S->c_src.p : nullptr, S->cplx_out.p, S->d_cnt);
    CB_TRY(timer_stop(S, "k_contact_gen"));
    S->launches++;
    S->contacts_geom_valid = true;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

// Readers of the contact records call this first: a step of the SMC system leaves them unwritten (the force kernel
// evaluates the geometry in place) until somebody asks.
int ensure_contact_geometry(System* S) {
    if (S->contacts_geom_valid || S->ns == 0)
        return CB200_OK;
    return enqueue_contact_gen(S);
}

int enqueue_pairs_only(System* S, bool want_pairs, bool want_contacts, bool defer_geometry) {
    if (S->ns == 0)
        return CB200_OK;
    if (want_pairs) {
        CB_TRY(enqueue_pair_pass(S, MODE_PAIRS, S->pairs.p, S->cap_pairs));
        S->pairs_valid = true;
    }
    if (want_contacts) {
        CB_TRY(enqueue_pair_pass(S, MODE_CONTACTS, S->c_shapes.p, S->cap_contacts));
        S->contacts_geom_valid = false;
        S->mat_ovr_valid = false;  // per-contact material records belonged to the previous list
        if (!defer_geometry)
            CB_TRY(enqueue_contact_gen(S));
    }
    return CB200_OK;
}

}  // namespace cb200
