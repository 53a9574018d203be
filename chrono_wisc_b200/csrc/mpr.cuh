// mpr.cuh -- Minkowski Portal Refinement on the device (the "MPR" narrowphase and the HYBRID fallback).
//
// Follows, operation for operation, the arithmetic of
//   src/chrono/collision/multicore/ChNarrowphaseMPR.cpp   (DiscoverPortal :487-555, RefinePortal :556-581,
//       FindPenetration :405-434, FindPos :93-114, ExpandPortal :69-87, MPRGetPoints :613-623, MPRCollision :627-645)
//   src/chrono/collision/multicore/ChCollisionUtilsMPR.cpp:27-97   (support mapping with the swept-sphere envelope)
//   src/chrono/collision/multicore/ChCollisionUtils.h:222-403      (GetSupportPoint_* per shape type)
// so that with -fmad=false the single contact equals the reference's bit for bit (checked in
// tests/test_gpu_narrow.py against the reference's own translation units, oracle/_ref).
// Layout differs from the reference: no virtual ConvexBase accessor, a shape is a plain ShapeView value
// (type, world position, world rotation, four dimension words) held in registers.
#pragma once
#include "vmath.cuh"
#include <float.h>

namespace cb200 {

#define MPR_TOL DBL_EPSILON  /* ChNarrowphaseMPR.cpp:28 */
#define MPR_MAX_IT 100       /* :29 */
#define MPR_WHILE_MAX 1000   /* :30 */
#define DEFAULT_EFF_RADIUS 0.1 /* ChCollisionInfo.cpp:19 */

struct ShapeView {
    int type;
    d3 pos;
    q4 rot;
    double d0, d1, d2, d3_;
    // CONVEXHULL: the hull's points in the shape frame (nv of them, 3 doubles each); unused by every other type.
    // TRIANGLE: the three WORLD vertices (shape_container::triangle_global) ride in pos, (rot.w, rot.x, rot.y) and
    // (d1, d2, d3_) -- a triangle has no frame or dimensions of its own (k_aabb packs them, tri_vertex reads them).
    const double* hull;
    int nv;
};

HD d3 tri_vertex(const ShapeView& S, int i) {
    return i == 0 ? S.pos : (i == 1 ? mk3(S.rot.w, S.rot.x, S.rot.y) : mk3(S.d1, S.d2, S.d3_));
}

// MM/real.h:59-68
HD double sign1(double x) { return x < 0 ? -1.0 : (x > 0 ? 1.0 : 0.0); }
// MM/real.h:90-109
HD bool is_zero(double x) { return fabs(x) < DBL_EPSILON; }
HD bool is_zero3(const d3& v) { return fabs(v.x) < DBL_EPSILON && fabs(v.y) < DBL_EPSILON && fabs(v.z) < DBL_EPSILON; }
HD bool is_equal(double a_, double b_) {
    const double ab = fabs(a_ - b_);
    if (ab < DBL_EPSILON)
        return true;
    const double a = fabs(a_), b = fabs(b_);
    return (b > a) ? (ab < DBL_EPSILON * b) : (ab < DBL_EPSILON * a);
}
// MM/real3.cpp:110-113
HD d3 normalize3(const d3& v) { return v / sqrt(dot3(v)); }

// ChCollisionUtils.h:268-279
HD d3 support_cylinder(double radius, double hheight, const d3& n) {
    const double s = sqrt(n.x * n.x + n.y * n.y);
    const double z = n.z < 0.0 ? -hheight : hheight;
    if (s < 1e-9)
        return mk3(radius, 0, z);
    return mk3(n.x * radius / s, n.y * radius / s, z);
}

// SupportVertNoMargin, ChCollisionUtilsMPR.cpp:27-69 (local frame, direction normalised here)
HD d3 support_local(const ShapeView& S, const d3& nv, double envelope) {
    const d3 n = normalize3(nv);
    d3 p = mk3(0);
    switch (S.type) {
        case CB200_SPHERE:
            p = S.d0 * n;
            break;
        case CB200_ELLIPSOID: {  // ChCollisionUtils.h:259-263
            const d3 B = mk3(S.d0, S.d1, S.d2);
            p = B * B * n / length3(B * n);
            break;
        }
        case CB200_BOX:
            p = mk3(sign1(n.x) * S.d0, sign1(n.y) * S.d1, sign1(n.z) * S.d2);
            break;
        case CB200_CYLINDER:
        case CB200_CYLSHELL:
            p = support_cylinder(S.d0, S.d2, n);
            break;
        case CB200_CONE: {  // ChCollisionUtils.h:297-313
            const double radius = S.d0, hheight = S.d2;
            const double sinAngle = radius / sqrt(radius * radius + 4 * hheight * hheight);
            if (n.z > length3(n) * sinAngle) {
                p = mk3(0, 0, hheight);
            } else {
                const double s = sqrt(n.x * n.x + n.y * n.y);
                if (s < 1e-9)
                    p = mk3(0, 0, -hheight);
                else
                    p = mk3(n.x * radius / s, n.y * radius / s, -hheight);
            }
            break;
        }
        case CB200_CAPSULE:  // Seg(B.y) along x + Sphere(B.x), ChCollisionUtils.h:316-327
            p = mk3(sign1(n.x) * S.d1, 0, 0) + S.d0 * n;
            break;
        case CB200_ROUNDEDBOX:
            p = mk3(sign1(n.x) * S.d0, sign1(n.y) * S.d1, sign1(n.z) * S.d2) + S.d3_ * n;
            break;
        case CB200_ROUNDEDCYL:
            p = support_cylinder(S.d0, S.d2, n) + S.d3_ * n;
            break;
        case CB200_CONVEXHULL: {  // GetSupportPoint_Convex, ChCollisionUtils.h:358-370
            double best = -DBL_MAX;
            p = mk3(S.hull[0], S.hull[1], S.hull[2]);
            for (int i = 0; i < S.nv; i++) {
                const d3 q = mk3(S.hull[3 * i], S.hull[3 * i + 1], S.hull[3 * i + 2]);
                const double dp = dot3(q, n);
                if (dp > best) {
                    best = dp;
                    p = q;
                }
            }
            break;
        }
        default:
            break;
    }
    p += envelope * n;
    return p;
}

// TransformSupportVert, ChCollisionUtilsMPR.cpp:76-97
HD d3 support_world(const ShapeView& S, const d3& n, double envelope) {
    if (S.type == CB200_TRIANGLE) {
        // world vertices, direction used as given (not normalised); GetSupportPoint_Triangle, ChCollisionUtils.h:227-243
        const d3 t0 = tri_vertex(S, 0), t1 = tri_vertex(S, 1), t2 = tri_vertex(S, 2);
        double dist = dot3(t0, n);
        d3 point = t0;
        if (dot3(t1, n) > dist) {
            dist = dot3(t1, n);
            point = t1;
        }
        if (dot3(t2, n) > dist)
            point = t2;
        return point + envelope * n;
    }
    return local_to_parent(S.pos, S.rot, support_local(S, rotate_t(n, S.rot), envelope));
}

struct MprSupport {
    d3 v, v1, v2;
};

HD void mpr_support(const ShapeView& A, const ShapeView& B, const d3& n, double envelope, MprSupport& s) {
    s.v1 = support_world(A, -n, envelope);
    s.v2 = support_world(B, n, envelope);
    s.v = s.v2 - s.v1;
}

struct MprPortal {
    MprSupport s0, s1, s2, s3, s4;
};

// ChNarrowphaseMPR.cpp:89-91
HD d3 portal_dir(const MprPortal& P) { return normalize3(cross3(P.s2.v - P.s1.v, P.s3.v - P.s1.v)); }

// ChNarrowphaseMPR.cpp:69-87
HD void expand_portal(MprPortal& P) {
    if (dot3(cross3(P.s4.v, P.s1.v), P.s0.v) < 0) {
        if (dot3(cross3(P.s4.v, P.s2.v), P.s0.v) < 0)
            P.s1 = P.s4;
        else
            P.s3 = P.s4;
    } else {
        if (dot3(cross3(P.s4.v, P.s3.v), P.s0.v) < 0)
            P.s2 = P.s4;
        else
            P.s1 = P.s4;
    }
}

// ChNarrowphaseMPR.cpp:93-114
HD void find_pos(const MprPortal& P, d3& point) {
    const d3 n = portal_dir(P);
    double b0 = dot3(cross3(P.s1.v, P.s2.v), P.s3.v);
    double b1 = dot3(cross3(P.s3.v, P.s2.v), P.s0.v);
    double b2 = dot3(cross3(P.s0.v, P.s1.v), P.s3.v);
    double b3 = dot3(cross3(P.s2.v, P.s1.v), P.s0.v);
    double sum = b0 + b1 + b2 + b3;
    if (sum <= 0.0) {
        b0 = 0;
        b1 = dot3(cross3(P.s2.v, P.s3.v), n);
        b2 = dot3(cross3(P.s3.v, P.s1.v), n);
        b3 = dot3(cross3(P.s1.v, P.s2.v), n);
        sum = b1 + b2 + b3;
    }
    const double inv = 1.0 / sum;
    const d3 p1 = (b0 * P.s0.v1 + b1 * P.s1.v1 + b2 * P.s2.v1 + b3 * P.s3.v1) * inv;
    const d3 p2 = (b0 * P.s0.v2 + b1 * P.s1.v2 + b2 * P.s2.v2 + b3 * P.s3.v2) * inv;
    point = (p1 + p2) * 0.5;
}

// ChNarrowphaseMPR.cpp:116-135
HD bool portal_encapsules_origin(const MprPortal& P, const d3& n) { return dot3(n, P.s1.v) >= 0.0; }
HD bool portal_can_encapsule_origin(const MprPortal& P, const d3& n) { return dot3(P.s4.v, n) >= 0.0; }
HD bool portal_reach_tolerance(const MprPortal& P, const d3& n) {
    const double dv1 = dot3(P.s1.v, n), dv2 = dot3(P.s2.v, n), dv3 = dot3(P.s3.v, n), dv4 = dot3(P.s4.v, n);
    const double dot1 = dv4 - dv1, dot2 = dv4 - dv2, dot3_ = dv4 - dv3;
    // std::min(a, b) = (b < a) ? b : a
    const double m12 = (dot2 < dot1) ? dot2 : dot1;
    const double m13 = (dot3_ < dot1) ? dot3_ : dot1;
    return is_equal(m12, MPR_TOL) || m13 < MPR_TOL;
}

// GetCenter, ChNarrowphaseMPR.cpp:40-55: a triangle's centroid; a hull's mean point as stored (shape frame, NOT
// rotated) plus the shape's world position [sic]; every other type is centred on its frame.
HD d3 shape_center(const ShapeView& S) {
    if (S.type == CB200_TRIANGLE)
        return (tri_vertex(S, 0) + tri_vertex(S, 1) + tri_vertex(S, 2)) * (1.0 / 3.0);
    if (S.type == CB200_CONVEXHULL) {
        d3 point = mk3(0);
        for (int i = 0; i < S.nv; i++)
            point += mk3(S.hull[3 * i], S.hull[3 * i + 1], S.hull[3 * i + 2]);
        return point / (double)S.nv + S.pos;
    }
    return S.pos;
}

// DiscoverPortal, ChNarrowphaseMPR.cpp:487-555.
__device__ inline int discover_portal(const ShapeView& A, const ShapeView& B, double envelope, MprPortal& P) {
    d3 n;
    P.s0.v1 = shape_center(A);
    P.s0.v2 = shape_center(B);
    P.s0.v = P.s0.v2 - P.s0.v1;
    if (is_zero3(P.s0.v))
        P.s0.v = mk3(1, 0, 0);
    n = normalize3(-P.s0.v);
    mpr_support(A, B, n, envelope, P.s1);
    if (dot3(P.s1.v, n) < 0.0)
        return -1;
    n = cross3(P.s0.v, P.s1.v);
    if (is_zero3(n))
        return is_zero3(P.s1.v) ? 1 : 2;
    n = normalize3(n);
    mpr_support(A, B, n, envelope, P.s2);
    if (dot3(P.s2.v, n) <= 0.0)
        return -1;
    n = normalize3(cross3(P.s1.v - P.s0.v, P.s2.v - P.s0.v));
    if (dot3(n, P.s0.v) > 0.0) {
        const MprSupport t = P.s1;
        P.s1 = P.s2;
        P.s2 = t;
        n = -n;
    }
    for (int wi = 0; wi < MPR_WHILE_MAX; wi++) {
        mpr_support(A, B, n, envelope, P.s3);
        if (dot3(P.s3.v, n) <= 0.0)
            return -1;
        if (dot3(cross3(P.s1.v, P.s3.v), P.s0.v) < 0.0) {
            P.s2 = P.s3;
            n = normalize3(cross3(P.s1.v - P.s0.v, P.s3.v - P.s0.v));
            continue;
        }
        if (dot3(cross3(P.s3.v, P.s2.v), P.s0.v) < 0.0) {
            P.s1 = P.s3;
            n = normalize3(cross3(P.s3.v - P.s0.v, P.s2.v - P.s0.v));
            continue;
        }
        break;
    }
    return 0;
}

// RefinePortal, ChNarrowphaseMPR.cpp:556-581
__device__ inline int refine_portal(const ShapeView& A, const ShapeView& B, double envelope, MprPortal& P) {
    for (int i = 0; i < MPR_MAX_IT; i++) {
        const d3 n = portal_dir(P);
        if (portal_encapsules_origin(P, n))
            return 0;
        mpr_support(A, B, n, envelope, P.s4);
        if (portal_reach_tolerance(P, n) || !portal_can_encapsule_origin(P, n))
            return portal_encapsules_origin(P, n) ? 0 : -1;
        expand_portal(P);
    }
    return -1;
}

// MPRContact + MPRGetPoints + MPRCollision, ChNarrowphaseMPR.cpp:584-645.  Returns the number of contacts (0/1).
__device__ inline int mpr_collision(const ShapeView& A, const ShapeView& B, double envelope, d3& normal, d3& ptA,
                                    d3& ptB, double& depth) {
    MprPortal P;
    d3 point;
    int result = discover_portal(A, B, envelope, P);
    if (result == 0) {
        result = refine_portal(A, B, envelope, P);
        if (result < 0)
            return 0;
        // FindPenetration :405-434
        for (int i = 0; i < MPR_MAX_IT; i++) {
            const d3 dir = portal_dir(P);
            mpr_support(A, B, dir, envelope, P.s4);
            if (portal_reach_tolerance(P, dir) || i == MPR_MAX_IT - 1) {
                normal = normalize3(dir);
                find_pos(P, point);
                break;
            }
            expand_portal(P);
        }
    } else if (result == 1) {
        // FindPenetrationTouch :436-445
        normal = normalize3(P.s1.v - P.s0.v);
        point = (P.s1.v1 + P.s1.v2) * .5;
    } else if (result == 2) {
        // FindPenetrationSegment :447-457
        point = (P.s1.v1 + P.s1.v2) * .5;
        normal = normalize3(P.s1.v);
    } else {
        return 0;
    }
    // MPRGetPoints :613-623
    const d3 p1 = dot3(support_world(A, -normal, envelope) - point, normal) * normal + point;
    const d3 p2 = dot3(support_world(B, normal, envelope) - point, normal) * normal + point;
    normal = -normal;
    ptA = p1 - normal * envelope;
    ptB = p2 + normal * envelope;
    depth = dot3(normal, ptB - ptA);
    return 1;
}

}  // namespace cb200
