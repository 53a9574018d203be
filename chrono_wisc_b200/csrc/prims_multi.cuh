// prims_multi.cuh -- analytic pair tests that can return more than one contact (capsule-capsule, box-capsule,
// box-cylshell, box-box), for the device narrowphase of non-sphere pairs.
//
// Arithmetic follows the reference operation for operation (bit-exact under -fmad=false):
//   capsule_capsule   src/chrono/collision/multicore/ChNarrowphasePRIMS.cpp:446-564
//   box_capsule       ChNarrowphasePRIMS.cpp:575-707
//   box_cylshell      ChNarrowphasePRIMS.cpp:719-915
//   box_box           ChNarrowphasePRIMS.cpp:924-1224
//   helpers           ChCollisionUtils.h:567-730 (snap_to_box_face/edge, get_*_corners, box_farthest_corner,
//                     box_closest_feature), ChCollisionUtilsPRIMS.cpp:142-330 (point_vs_face, segment_vs_edge,
//                     box_intersects_box: 15-axis SAT)
// Outputs go to caller-provided arrays of MAX_PAIR_CONTACTS records.  The reference reserves 8 slots for a box-box
// pair (ChNarrowphase.cpp:128) but its face-face branch does not bound its writes; here a ninth contact is dropped.
#pragma once
#include "prims.cuh"
#include "mpr.cuh"

namespace cb200 {

constexpr int MAX_PAIR_CONTACTS = 8;

struct v3 {  // indexable 3-vector (fully unrolled uses stay in registers)
    double a[3];
    __device__ __forceinline__ double& operator[](int i) { return a[i]; }
    __device__ __forceinline__ double operator[](int i) const { return a[i]; }
};
__device__ __forceinline__ v3 tov(const d3& d) {
    v3 r;
    r.a[0] = d.x, r.a[1] = d.y, r.a[2] = d.z;
    return r;
}
__device__ __forceinline__ d3 tod(const v3& v) { return mk3(v.a[0], v.a[1], v.a[2]); }

// MM/real4.cpp:128-131
__device__ __forceinline__ q4 qinv(const q4& a) {
    const double t1 = a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
    return mkq(a.w / t1, -a.x / t1, -a.y / t1, -a.z / t1);
}

// std::min / std::max as the reference's libstdc++ evaluates them
__device__ __forceinline__ double smin(double a, double b) { return (b < a) ? b : a; }
__device__ __forceinline__ double smax(double a, double b) { return (a < b) ? b : a; }

__device__ __forceinline__ void push_contact(Contact* out, int& j, const d3& n, const d3& p1, const d3& p2,
                                             double depth, double erad) {
    if (j < MAX_PAIR_CONTACTS) {
        out[j].n = n, out[j].p1 = p1, out[j].p2 = p2, out[j].depth = depth, out[j].erad = erad;
    }
    j++;
}

// ChNarrowphasePRIMS.cpp:446-564
__device__ inline int capsule_capsule(const d3& pos1, const q4& rot1, double radius1, double hlen1, const d3& pos2,
                                      const q4& rot2, double radius2, double hlen2, double separation,
                                      Contact* out) {
    const d3 pos = rotate_t(pos2 - pos1, rot1);
    const q4 rot = qmul(qinv(rot1), rot2);
    const d3 W1 = amat_w(rot1), W2 = amat_w(rot2), W = amat_w(rot);
    const double radSum = radius1 + radius2;
    const double radSum_s = radSum + separation;
    const double radSum_s2 = radSum_s * radSum_s;
    int numLocs = 0;
    d3 locs1[2], locs2[2];
    const double denom = 1 - W.z * W.z;
    if (denom < (double)1e-4f) {
        if (pos.x * pos.x + pos.y * pos.y >= radSum_s2)
            return 0;
        const double locs[2] = {smin(hlen1, pos.z + hlen2), smax(-hlen1, pos.z - hlen2)};
        if (locs[0] > locs[1]) {
            numLocs = 2;
            locs1[0] = pos1 + locs[0] * W1;
            locs2[0] = local_to_parent(pos1, rot1, mk3(pos.x, pos.y, locs[0]));
            locs1[1] = pos1 + locs[1] * W1;
            locs2[1] = local_to_parent(pos1, rot1, mk3(pos.x, pos.y, locs[1]));
        } else {
            numLocs = 1;
            locs1[0] = pos1 + locs[pos.z < 0] * W1;
            locs2[0] = local_to_parent(pos1, rot1, mk3(pos.x, pos.y, locs[pos.y > 0]));
        }
    } else {
        double alpha2 = (W.z * pos.z - dot3(W, pos)) / denom;
        double alpha1 = W.z * alpha2 + pos.z;
        if (alpha1 < -hlen1) {
            alpha1 = -hlen1;
            alpha2 = -dot3(pos, W) - hlen1 * W.z;
        } else if (alpha1 > hlen1) {
            alpha1 = hlen1;
            alpha2 = -dot3(pos, W) + hlen1 * W.z;
        }
        if (alpha2 < -hlen2) {
            alpha2 = -hlen2;
            alpha1 = clampd(pos.z - hlen2 * W.z, -hlen1, hlen1);
        } else if (alpha2 > hlen2) {
            alpha2 = hlen2;
            alpha1 = clampd(pos.z + hlen2 * W.z, -hlen1, hlen1);
        }
        numLocs = 1;
        locs1[0] = pos1 + alpha1 * W1;
        locs2[0] = pos2 + alpha2 * W2;
    }
    const double effRad = radius1 * radius2 / radSum;
    int j = 0;
    for (int i = 0; i < numLocs; i++) {
        const d3 delta = locs2[i] - locs1[i];
        const double dist2 = dot3(delta, delta);
        if (dist2 >= radSum_s2 || dist2 < 1e-12)
            continue;
        const double dist = sqrt(dist2);
        const d3 n = delta / dist;
        push_contact(out, j, n, locs1[i] + n * radius1, locs2[i] - n * radius2, dist - radSum, effRad);
    }
    return j;
}

// ChNarrowphasePRIMS.cpp:575-707
__device__ inline int box_capsule(const d3& pos1, const q4& rot1, const d3& hdims1, const d3& pos2, const q4& rot2,
                                  double radius2, double hlen2, double separation, Contact* out) {
    const double radius2_s = radius2 + separation;
    const d3 pos = rotate_t(pos2 - pos1, rot1);
    const q4 rot = qmul(qinv(rot1), rot2);
    const d3 W = amat_w(rot);
    const v3 hexp = tov(hdims1 + radius2_s), p = tov(pos), w = tov(W);
    double tMin = -DBL_MAX, tMax = DBL_MAX;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        if (fabs(w[a]) < 1e-5) {
            if (fabs(p[a]) > hexp[a])
                return 0;
        } else {
            const double t1 = (-hexp[a] - p[a]) / w[a];
            const double t2 = (hexp[a] - p[a]) / w[a];
            tMin = smax(tMin, smin(t1, t2));
            tMax = smin(tMax, smax(t1, t2));
            if (tMin > tMax)
                return 0;
        }
    }
    d3 locs[2] = {pos + tMin * W, pos + tMax * W};
    double t[2];
    for (int i = 0; i < 2; i++) {
        snap_to_box(hdims1, locs[i]);
        t[i] = clampd(dot3(locs[i] - pos, W), -hlen2, hlen2);
    }
    const int numSpheres = is_equal(t[0], t[1]) ? 1 : 2;
    int j = 0;
    for (int i = 0; i < numSpheres; i++) {
        const d3 spherePos = pos + W * t[i];
        d3 boxPos = spherePos;
        const unsigned code = snap_to_box(hdims1, boxPos);
        const d3 delta = spherePos - boxPos;
        const double dist2 = dot3(delta, delta);
        if (dist2 >= radius2_s * radius2_s || dist2 <= 1e-12)
            continue;
        const double dist = sqrt(dist2);
        const d3 n = rotate(delta / dist, rot1);
        const double er = ((code != 1) && (code != 2) && (code != 4)) ? radius2 * EDGE_RADIUS / (radius2 + EDGE_RADIUS)
                                                                    : radius2;
        push_contact(out, j, n, local_to_parent(pos1, rot1, boxPos),
                     local_to_parent(pos1, rot1, spherePos) - n * radius2, dist - radius2, er);
    }
    return j;
}

// ---- box utilities ------------------------------------------------------------------------------------------

// MM/real.h:153-161
__device__ __forceinline__ void clamp_value(double& x, double low, double high) {
    if (low > high) {
        const double t = low;
        low = high;
        high = t;
    }
    if (x < low)
        x = low;
    else if (x > high)
        x = high;
}

// ChCollisionUtils.h:567-590
__device__ __forceinline__ d3 snap_to_box_face(const d3& hdims, const d3& pt_on_box, unsigned code,
                                               const d3& pt_to_snap) {
    const unsigned side = code >> 1;
    d3 r = pt_to_snap;
    if (side == 0) {
        r.x = pt_on_box.x;
        clamp_value(r.y, -hdims.y, hdims.y);
        clamp_value(r.z, -hdims.z, hdims.z);
    } else if (side == 1) {
        r.y = pt_on_box.y;
        clamp_value(r.x, -hdims.x, hdims.x);
        clamp_value(r.z, -hdims.z, hdims.z);
    } else {
        r.z = pt_on_box.z;
        clamp_value(r.x, -hdims.x, hdims.x);
        clamp_value(r.y, -hdims.y, hdims.y);
    }
    return r;
}

// ChCollisionUtils.h:595-616
__device__ __forceinline__ d3 snap_to_box_edge(const d3& hdims, const d3& pt_on_box, unsigned code,
                                               const d3& pt_to_snap) {
    const unsigned axis = (~code & 7) >> 1;
    d3 pt1 = pt_on_box, pt2 = pt_on_box;
    if (axis == 0) {
        pt1.x = -hdims.x;
        pt2.x = hdims.x;
    } else if (axis == 1) {
        pt1.y = -hdims.y;
        pt2.y = hdims.y;
    } else {
        pt1.z = -hdims.z;
        pt2.z = hdims.z;
    }
    const d3 edge = pt2 - pt1;
    const double t = dot3(pt_to_snap - pt1, edge) / (length3(edge) * length3(edge));
    return pt1 + t * edge;
}

// ChCollisionUtils.h:620-633
__device__ __forceinline__ void get_edge_corners(const d3& pt_on_edge, unsigned code, d3* corners) {
    code = ((~code & 7) >> 1);
    corners[0] = pt_on_edge;
    corners[1] = pt_on_edge;
    if (code == 0)
        corners[1].x = -corners[1].x;
    else if (code == 1)
        corners[1].y = -corners[1].y;
    else
        corners[1].z = -corners[1].z;
}

// ChCollisionUtils.h:637-663
__device__ __forceinline__ void get_face_corners(const d3& pt_on_face, unsigned code, d3* corners) {
    code = code >> 1;
    corners[0] = pt_on_face;
    corners[1] = pt_on_face;
    if (code == 0) {
        corners[1].y = -corners[1].y;
        corners[2] = corners[1];
        corners[2].z = -corners[1].z;
        corners[3] = pt_on_face;
        corners[3].z = corners[2].z;
    } else if (code == 1) {
        corners[1].z = -corners[1].z;
        corners[2] = corners[1];
        corners[2].x = -corners[1].x;
        corners[3] = pt_on_face;
        corners[3].x = corners[2].x;
    } else {
        corners[1].x = -corners[1].x;
        corners[2] = corners[1];
        corners[2].y = -corners[1].y;
        corners[3] = pt_on_face;
        corners[3].y = corners[2].y;
    }
}

// ChCollisionUtils.h:713-719, :528-531
__device__ __forceinline__ d3 box_farthest_corner(const d3& hdims, const d3& dir) {
    return mk3((dir.x < 0) ? hdims.x : -hdims.x, (dir.y < 0) ? hdims.y : -hdims.y, (dir.z < 0) ? hdims.z : -hdims.z);
}
__device__ __forceinline__ unsigned box_closest_feature(const d3& dir) {
    const double threshold = 0.01;
    return ((fabs(dir.x) > threshold) << 0) | ((fabs(dir.y) > threshold) << 1) | ((fabs(dir.z) > threshold) << 2);
}

__device__ __forceinline__ double sel3(const d3& v, unsigned i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
__device__ __forceinline__ void set3(d3& v, unsigned i, double s) {
    if (i == 0)
        v.x = s;
    else if (i == 1)
        v.y = s;
    else
        v.z = s;
}

// ChCollisionUtilsPRIMS.cpp:142-181
__device__ inline bool point_vs_face(const d3& hdims, const d3& pt_on_face, unsigned code, const d3& point,
                                     double separation, d3& result, d3& normal, double& dist) {
    const unsigned i = code >> 1, in = (i + 1) % 3, ip = (i + 2) % 3;
    if (fabs(sel3(point, in)) > sel3(hdims, in))
        return false;
    if (fabs(sel3(point, ip)) > sel3(hdims, ip))
        return false;
    const int sign = (sel3(pt_on_face, i) > 0 ? +1 : -1);
    dist = sign * sel3(point, i) - sel3(hdims, i);
    if (dist > separation)
        return false;
    result = point;
    set3(result, i, sign * sel3(hdims, i));
    set3(normal, i, sign);
    set3(normal, in, 0);
    set3(normal, ip, 0);
    return true;
}

// ChCollisionUtilsPRIMS.cpp:190-237
__device__ inline bool segment_vs_edge(const d3& hdims, const d3& pt_on_edge, unsigned code, const d3& pt1,
                                       const d3& pt2, d3& locE, d3& locS) {
    const unsigned i = (~code & 7) >> 1;
    const d3 s = pt2 - pt1;
    const double e = dot3(s, s);
    const double si = sel3(s, i);
    const double d = e - si * si;
    if (d < 3e-4 * e)
        return false;
    const d3 r = pt_on_edge - pt1;
    const double ri = sel3(r, i);
    const double tS = (dot3(s, r) - si * ri) / d;
    const double eps = 1e-8 * e;
    if (tS < eps || tS > 1 - eps)
        return false;
    const double tE = si * tS - ri;
    locE = pt_on_edge;
    set3(locE, i, sel3(locE, i) + tE);
    if (fabs(sel3(locE, i)) > sel3(hdims, i))
        return false;
    locS = pt1 + s * tS;
    return true;
}

// 15-axis SAT, ChCollisionUtilsPRIMS.cpp:262-330.  R(i,j) = Mat33(q) (MM/matrix.h:112-124), column-major.
__device__ inline int box_intersects_box(const d3& hdims1_, const d3& hdims2_, const d3& pos_, const q4& q,
                                         double separation, d3& dir) {
    double R[3][3];  // R[i][j] = row i, column j
    R[0][0] = 1.0 - 2.0 * q.y * q.y - 2.0 * q.z * q.z;
    R[1][0] = 2.0 * q.x * q.y + 2.0 * q.z * q.w;
    R[2][0] = 2.0 * q.x * q.z - 2.0 * q.y * q.w;
    R[0][1] = 2.0 * q.x * q.y - 2.0 * q.z * q.w;
    R[1][1] = 1.0 - 2.0 * q.x * q.x - 2.0 * q.z * q.z;
    R[2][1] = 2.0 * q.y * q.z + 2.0 * q.x * q.w;
    R[0][2] = 2.0 * q.x * q.z + 2.0 * q.y * q.w;
    R[1][2] = 2.0 * q.y * q.z - 2.0 * q.x * q.w;
    R[2][2] = 1.0 - 2.0 * q.x * q.x - 2.0 * q.y * q.y;
    double Ra[3][3];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++)
            Ra[i][j] = fabs(R[i][j]);
    const v3 h1 = tov(hdims1_), h2 = tov(hdims2_), pos = tov(pos_);
    double minOverlap = DBL_MAX;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const double r2 = Ra[i][0] * h2[0] + Ra[i][1] * h2[1] + Ra[i][2] * h2[2];
        const double overlap = h1[i] + r2 - fabs(pos[i]) + separation;
        if (overlap <= 0)
            return 0;
        if (overlap < minOverlap) {
            dir = mk3(i == 0 ? 1.0 : 0.0, i == 1 ? 1.0 : 0.0, i == 2 ? 1.0 : 0.0);
            minOverlap = overlap;
        }
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const double r1 = Ra[0][i] * h1[0] + Ra[1][i] * h1[1] + Ra[2][i] * h1[2];
        const double overlap =
            r1 + h2[i] - fabs(R[0][i] * pos[0] + R[1][i] * pos[1] + R[2][i] * pos[2]) + separation;
        if (overlap <= 0)
            return 0;
        if (overlap < minOverlap) {
            dir = mk3(R[0][i], R[1][i], R[2][i]);
            minOverlap = overlap;
        }
    }
#pragma unroll
    for (int x1 = 0; x1 < 3; x1++) {
        const int y1 = (x1 + 1) % 3, z1 = (x1 + 2) % 3;
#pragma unroll
        for (int x2 = 0; x2 < 3; x2++) {
            const int y2 = (x2 + 1) % 3, z2 = (x2 + 2) % 3;
            v3 cp;
            cp[x1] = 0;
            cp[y1] = -R[z1][x2];
            cp[z1] = R[y1][x2];
            const double lengthSqr = cp[0] * cp[0] + cp[1] * cp[1] + cp[2] * cp[2];
            if (lengthSqr > 1e-6) {
                const double r1 = h1[y1] * Ra[z1][x2] + h1[z1] * Ra[y1][x2];
                const double r2 = h2[y2] * Ra[x1][z2] + h2[z2] * Ra[x1][y2];
                double overlap = r1 + r2 - fabs(pos[z1] * R[y1][x2] - pos[y1] * R[z1][x2]) + separation;
                if (overlap <= 0)
                    return 0;
                const double ooLen = 1 / sqrt(lengthSqr);
                overlap *= ooLen;
                if (overlap < minOverlap) {
                    dir = mk3(cp[0] * ooLen, cp[1] * ooLen, cp[2] * ooLen);
                    minOverlap = overlap;
                }
            }
        }
    }
    return (minOverlap >= separation) ? -1 : +1;
}

// contact from two world points (the edge/corner branches of box_box): delta = penetrated * (pO - pT)
__device__ __forceinline__ void push_pointpair(Contact* out, int& j, const d3& pT, const d3& pO, int penetrated,
                                               double dist, const d3& delta, double erad) {
    push_contact(out, j, delta / dist, pT, pO, penetrated * dist, erad);
}

// Which branch of box_box a pair takes: 0 separated, 1 corner, 2 edge-edge, 3 face-face, 4 face of T / edge of O,
// 5 face of O / edge of T (6: none of them, no contact).  Same arithmetic as the head of box_box; used to order the
// box-box pairs so that the threads of a warp run the same branch (pairs.cu k_bb_classify).
__device__ inline int box_box_case(const d3& posT, const q4& rotT, const d3& hdimsT, const d3& posO, const q4& rotO,
                                   const d3& hdimsO, double separation) {
    const d3 pos = rotate_t(posO - posT, rotT);
    const q4 rot = qmul(qinv(rotT), rotO);
    d3 dirT;
    const int penetrated = box_intersects_box(hdimsT, hdimsO, pos, rot, separation, dirT);
    if (penetrated == 0)
        return 0;
    if (dot3(pos, dirT) > 0)
        dirT = -dirT;
    const d3 dirO = rotate_t(-dirT, rot);
    const unsigned codeT = box_closest_feature(dirT), codeO = box_closest_feature(dirO);
    const bool Tcorner = (codeT == 7), Tedge = (codeT == 3) || (codeT == 5) || (codeT == 6),
               Tface = (codeT == 1) || (codeT == 2) || (codeT == 4);
    const bool Ocorner = (codeO == 7), Oedge = (codeO == 3) || (codeO == 5) || (codeO == 6),
               Oface = (codeO == 1) || (codeO == 2) || (codeO == 4);
    if (Tcorner || Ocorner)
        return 1;
    if (Tedge && Oedge)
        return 2;
    if (Tface && Oface)
        return 3;
    if (Tface)
        return 4;
    if (Oface)
        return 5;
    return 6;
}

// ChNarrowphasePRIMS.cpp:924-1224
__device__ inline int box_box(const d3& posT, const q4& rotT, const d3& hdimsT, const d3& posO, const q4& rotO,
                              const d3& hdimsO, double separation, Contact* out) {
    const d3 pos = rotate_t(posO - posT, rotT);
    const q4 rot = qmul(qinv(rotT), rotO);
    d3 dirT;
    const int penetrated = box_intersects_box(hdimsT, hdimsO, pos, rot, separation, dirT);
    if (penetrated == 0)
        return 0;
    if (dot3(pos, dirT) > 0)
        dirT = -dirT;
    const d3 dirO = rotate_t(-dirT, rot);
    d3 cornerT = box_farthest_corner(hdimsT, dirT);
    d3 cornerO = box_farthest_corner(hdimsO, dirO);
    const unsigned codeT = box_closest_feature(dirT), codeO = box_closest_feature(dirO);
    const bool Tcorner = (codeT == 7), Tedge = (codeT == 3) || (codeT == 5) || (codeT == 6),
               Tface = (codeT == 1) || (codeT == 2) || (codeT == 4);
    const bool Ocorner = (codeO == 7), Oedge = (codeO == 3) || (codeO == 5) || (codeO == 6),
               Oface = (codeO == 1) || (codeO == 2) || (codeO == 4);
    int j = 0;

    // (1) a corner is involved: one contact
    if (Tcorner || Ocorner) {
        if (Tcorner) {
            if (Oedge)
                cornerO = snap_to_box_edge(hdimsO, cornerO, codeO, rotate_t(cornerT - pos, rot));
            else if (Oface)
                cornerO = snap_to_box_face(hdimsO, cornerO, codeO, rotate_t(cornerT - pos, rot));
        } else {
            if (Tedge)
                cornerT = snap_to_box_edge(hdimsT, cornerT, codeT, rotate(cornerO, rot) + pos);
            else if (Tface)
                cornerT = snap_to_box_face(hdimsT, cornerT, codeT, rotate(cornerO, rot) + pos);
        }
        const d3 pT = rotate(cornerT, rotT) + posT;
        const d3 pO = rotate(cornerO, rotO) + posO;
        const d3 delta = (double)penetrated * (pO - pT);
        const double dist = sqrt(dot3(delta, delta));
        push_pointpair(out, j, pT, pO, penetrated, dist, delta, EDGE_RADIUS / 2);
        return 1;
    }

    // (2) edge against edge: 0 or 1 contact
    if (Tedge && Oedge) {
        d3 cornersO[2];
        get_edge_corners(cornerO, codeO, cornersO);
        const d3 corner0 = rotate(cornersO[0], rot) + pos;
        const d3 corner1 = rotate(cornersO[1], rot) + pos;
        d3 locT, locO;
        if (segment_vs_edge(hdimsT, cornerT, codeT, corner0, corner1, locT, locO)) {
            locT = rotate(locT, rotT) + posT;
            locO = rotate(locO, rotT) + posT;
            const d3 delta = (double)penetrated * (locO - locT);
            const double dist = sqrt(dot3(delta, delta));
            if (penetrated == 1 && dist > separation)
                return 0;
            push_pointpair(out, j, locT, locO, penetrated, dist, delta, EDGE_RADIUS / 2);
            return 1;
        }
        return 0;
    }

    // (3) face against face: up to 8 contacts
    if (Tface && Oface) {
        d3 cornersT[4], cornersO[4];
        get_face_corners(cornerT, codeT, cornersT);
        get_face_corners(cornerO, codeO, cornersO);
        d3 loc, nrm;
        double dist;
        for (unsigned i = 0; i < 4; i++) {
            if (point_vs_face(hdimsT, cornerT, codeT, rotate(cornersO[i], rot) + pos, separation, loc, nrm, dist))
                push_contact(out, j, rotate(nrm, rotT), rotate(loc, rotT) + posT, rotate(cornersO[i], rotO) + posO,
                             dist, EDGE_RADIUS);
            if (point_vs_face(hdimsO, cornerO, codeO, rotate_t(cornersT[i] - pos, rot), separation, loc, nrm, dist))
                push_contact(out, j, rotate(-nrm, rotO), rotate(cornersT[i], rotT) + posT, rotate(loc, rotO) + posO,
                             dist, EDGE_RADIUS);
        }
        const unsigned codeTN = (codeT | (codeT >> 1) | (codeT << 2)) & 7;
        const unsigned codeTP = (codeT | (codeT << 1) | (codeT >> 2)) & 7;
        d3 locT, locO;
        for (unsigned i = 0; i < 4; i++) {
            const unsigned codeE = i & 1 ? codeTP : codeTN;
            for (unsigned j1 = 0, j2 = 3; j1 < 4; j2 = j1++) {
                const d3 cornerO1 = rotate(cornersO[j1], rot) + pos;
                const d3 cornerO2 = rotate(cornersO[j2], rot) + pos;
                if (segment_vs_edge(hdimsT, cornersT[i], codeE, cornerO1, cornerO2, locT, locO)) {
                    locT = rotate(locT, rotT) + posT;
                    locO = rotate(locO, rotT) + posT;
                    const d3 delta = (double)penetrated * (locO - locT);
                    dist = sqrt(dot3(delta, delta));
                    if (penetrated == 1 && dist > separation)
                        continue;
                    push_pointpair(out, j, locT, locO, penetrated, dist, delta, EDGE_RADIUS);
                }
            }
        }
        return j < MAX_PAIR_CONTACTS ? j : MAX_PAIR_CONTACTS;
    }

    // (4) face of this box against an edge of the other: up to 2 + 4 contacts
    if (Tface) {
        d3 cornersT[4], cornersO[2];
        get_face_corners(cornerT, codeT, cornersT);
        get_edge_corners(cornerO, codeO, cornersO);
        d3 loc, nrm;
        double dist;
        for (unsigned i = 0; i < 2; i++) {
            if (point_vs_face(hdimsT, cornerT, codeT, rotate(cornersO[i], rot) + pos, separation, loc, nrm, dist))
                push_contact(out, j, rotate(nrm, rotT), rotate(loc, rotT) + posT, rotate(cornersO[i], rotO) + posO,
                             dist, EDGE_RADIUS);
        }
        d3 locT, locO;
        for (unsigned i1 = 0, i2 = 3; i1 < 4; i2 = i1++) {
            const d3 cornerT1 = rotate_t(cornersT[i1] - pos, rot);
            const d3 cornerT2 = rotate_t(cornersT[i2] - pos, rot);
            if (segment_vs_edge(hdimsO, cornerO, codeO, cornerT1, cornerT2, locO, locT)) {
                locT = rotate(locT, rotO) + posO;
                locO = rotate(locO, rotO) + posO;
                const d3 delta = (double)penetrated * (locO - locT);
                dist = sqrt(dot3(delta, delta));
                if (penetrated == 1 && dist > separation)
                    continue;
                push_pointpair(out, j, locT, locO, penetrated, dist, delta, EDGE_RADIUS);
            }
        }
        return j < MAX_PAIR_CONTACTS ? j : MAX_PAIR_CONTACTS;
    }

    // (5) face of the other box against an edge of this one
    if (Oface) {
        d3 cornersT[2], cornersO[4];
        get_edge_corners(cornerT, codeT, cornersT);
        get_face_corners(cornerO, codeO, cornersO);
        d3 loc, nrm;
        double dist;
        for (unsigned i = 0; i < 2; i++) {
            if (point_vs_face(hdimsO, cornerO, codeO, rotate_t(cornersT[i] - pos, rot), separation, loc, nrm, dist))
                push_contact(out, j, rotate(-nrm, rotO), rotate(cornersT[i], rotT) + posT, rotate(loc, rotO) + posO,
                             dist, EDGE_RADIUS);
        }
        d3 locT, locO;
        for (unsigned i1 = 0, i2 = 3; i1 < 4; i2 = i1++) {
            const d3 cornerO1 = rotate(cornersO[i1], rot) + pos;
            const d3 cornerO2 = rotate(cornersO[i2], rot) + pos;
            if (segment_vs_edge(hdimsT, cornerT, codeT, cornerO1, cornerO2, locT, locO)) {
                locT = rotate(locT, rotT) + posT;
                locO = rotate(locO, rotT) + posT;
                const d3 delta = (double)penetrated * (locO - locT);
                dist = sqrt(dot3(delta, delta));
                if (penetrated == 1 && dist > separation)
                    continue;
                push_pointpair(out, j, locT, locO, penetrated, dist, delta, EDGE_RADIUS);
            }
        }
        return j < MAX_PAIR_CONTACTS ? j : MAX_PAIR_CONTACTS;
    }
    return 0;
}

// ChNarrowphasePRIMS.cpp:719-915.  Only axis-parallel / axis-perpendicular configurations are analytic; -1 asks for the
// MPR fall-back.  (The reference clamps the circle/edge intersections with c[i2] in the i2-edge cases and with c[i3] in
// the i3-edge cases; kept as is.)
__device__ inline int box_cylshell(const d3& pos1, const q4& rot1, const d3& hdims_, const d3& pos2, const q4& rot2,
                                   double radius, double hlen, double separation, Contact* out) {
    const v3 c = tov(rotate_t(pos2 - pos1, rot1));
    const q4 rot = qmul(qinv(rot1), rot2);
    const v3 a = tov(amat_w(rot));
    const v3 hdims = tov(hdims_);
    v3 c_abs, a_abs;
#pragma unroll
    for (int x = 0; x < 3; x++)
        c_abs[x] = fabs(c[x]), a_abs[x] = fabs(a[x]);
    const double threshold_par = 1e-4, threshold_perp = 1e-4;
#pragma unroll
    for (int i1 = 0; i1 < 3; i1++) {
        const int i2 = (i1 + 1) % 3, i3 = (i1 + 2) % 3;
        // (1) cylinder axis parallel to the i1 box face
        if (a_abs[i1] < threshold_par) {
            if (c_abs[i1] > hdims[i1] + radius + separation)
                return 0;
            if (c_abs[i1] < hdims[i1])
                continue;
            double tMin = -DBL_MAX, tMax = DBL_MAX;
            if (a_abs[i2] > threshold_par) {
                const double t1 = (-hdims[i2] - c[i2]) / a[i2];
                const double t2 = (+hdims[i2] - c[i2]) / a[i2];
                tMin = smax(tMin, smin(t1, t2));
                tMax = smin(tMax, smax(t1, t2));
                if (tMin > tMax)
                    return 0;
            }
            if (a_abs[i3] > threshold_par) {
                const double t1 = (-hdims[i3] - c[i3]) / a[i3];
                const double t2 = (+hdims[i3] - c[i3]) / a[i3];
                tMin = smax(tMin, smin(t1, t2));
                tMax = smin(tMax, smax(t1, t2));
                if (tMin > tMax)
                    return 0;
            }
            clamp_value(tMin, -hlen, +hlen);
            clamp_value(tMax, -hlen, +hlen);
            const d3 cd = tod(c), ad = tod(a);
            const d3 locs[2] = {cd + tMin * ad, cd + tMax * ad};
            int j = 0;
            for (int i = 0; i < 2; i++) {
                d3 boxPoint = locs[i];
                snap_to_box(hdims_, boxPoint);
                d3 u = locs[i] - boxPoint;
                const double u_nrm = sqrt(dot3(u, u));
                u = u / u_nrm;
                const d3 cylPoint = locs[i] - radius * u;
                push_contact(out, j, rotate(u, rot1), local_to_parent(pos1, rot1, boxPoint),
                             local_to_parent(pos1, rot1, cylPoint), u_nrm - radius, radius);
            }
            return 2;
        }
        // (2) cylinder axis perpendicular to the i1 box face
        if (a_abs[i1] > 1 - threshold_perp) {
            if (c_abs[i1] > hdims[i1] + hlen + separation)
                return 0;
            if (c_abs[i1] < hdims[i1])
                continue;
            if (c_abs[i2] > hdims[i2] || c_abs[i3] > hdims[i3])
                continue;
            const int sign = (c[i1] > 0) ? +1 : -1;
            v3 u;
            u[0] = u[1] = u[2] = 0;
            u[i1] = sign;
            v3 locs[8];
            int nc = 0;
            double discr;
            // negative i2 edge
            discr = radius * radius - (c[i2] + hdims[i2]) * (c[i2] + hdims[i2]);
            if (discr > 0) {
                const double sq = sqrt(discr);
                locs[nc][i2] = -hdims[i2], locs[nc][i3] = clampd(c[i2] + sq, -hdims[i3], +hdims[i3]), nc++;
                locs[nc][i2] = -hdims[i2], locs[nc][i3] = clampd(c[i2] - sq, -hdims[i3], +hdims[i3]), nc++;
            } else {
                locs[nc][i2] = c[i2] - radius, locs[nc][i3] = c[i3], nc++;
            }
            // positive i2 edge
            discr = radius * radius - (c[i2] - hdims[i2]) * (c[i2] - hdims[i2]);
            if (discr > 0) {
                const double sq = sqrt(discr);
                locs[nc][i2] = +hdims[i2], locs[nc][i3] = clampd(c[i2] + sq, -hdims[i3], +hdims[i3]), nc++;
                locs[nc][i2] = +hdims[i2], locs[nc][i3] = clampd(c[i2] - sq, -hdims[i3], +hdims[i3]), nc++;
            } else {
                locs[nc][i2] = c[i2] + radius, locs[nc][i3] = c[i3], nc++;
            }
            // negative i3 edge
            discr = radius * radius - (c[i3] + hdims[i3]) * (c[i3] + hdims[i3]);
            if (discr > 0) {
                const double sq = sqrt(discr);
                locs[nc][i3] = -hdims[i3], locs[nc][i2] = clampd(c[i3] + sq, -hdims[i2], +hdims[i2]), nc++;
                locs[nc][i3] = -hdims[i3], locs[nc][i2] = clampd(c[i3] - sq, -hdims[i2], +hdims[i2]), nc++;
            } else {
                locs[nc][i3] = c[i3] - radius, locs[nc][i2] = c[i2], nc++;
            }
            // positive i3 edge
            discr = radius * radius - (c[i3] - hdims[i3]) * (c[i3] - hdims[i3]);
            if (discr > 0) {
                const double sq = sqrt(discr);
                locs[nc][i3] = +hdims[i3], locs[nc][i2] = clampd(c[i3] + sq, -hdims[i2], +hdims[i2]), nc++;
                locs[nc][i3] = +hdims[i3], locs[nc][i2] = clampd(c[i3] - sq, -hdims[i2], +hdims[i2]), nc++;
            } else {
                locs[nc][i3] = c[i3] + radius, locs[nc][i2] = c[i2], nc++;
            }
            int j = 0;
            const d3 nrm = rotate(tod(u), rot1);
            for (int i = 0; i < nc; i++) {
                locs[i][i1] = sign * hdims[i1];
                const d3 p1 = local_to_parent(pos1, rot1, tod(locs[i]));
                locs[i][i1] = c[i1] - sign * hlen;
                const d3 p2 = local_to_parent(pos1, rot1, tod(locs[i]));
                push_contact(out, j, nrm, p1, p2, c[i1] - sign * hlen - sign * hdims[i1], radius);
            }
            return nc;
        }
    }
    return -1;
}

}  // namespace cb200
