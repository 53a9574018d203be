// prims.cuh -- analytic pair tests on the device (the "PRIMS" narrowphase).
//
// Each function follows the arithmetic of its reference counterpart in
// src/chrono/collision/multicore/ChNarrowphasePRIMS.cpp (line ranges cited per function) and the helpers of
// ChCollisionUtils.h:494-560, operation for operation, so that with -fmad=false the contacts equal the
// reference's bit for bit.  Conventions as in the reference: the normal points from shape 1 towards shape 2
// (sphere_sphere builds it from delta = pos2 - pos1), p1/p2 are the points on shape 1/2, depth < 0 when penetrating.
#pragma once
#include "vmath.cuh"

namespace cb200 {

#define EDGE_RADIUS 0.1                   /* ChNarrowphasePRIMS.cpp:31 */
#define TINY_F ((double)1e-12f)           /* the reference compares against the float literal 1e-12f */

struct Contact {
    d3 n, p1, p2;
    double depth, erad;
};

// ChNarrowphasePRIMS.cpp:40-71
HD bool sphere_sphere(const d3& pos1, double r1, const d3& pos2, double r2, double separation, Contact& c) {
    const d3 delta = pos2 - pos1;
    const double dist2 = dot3(delta, delta);
    const double radSum = r1 + r2;
    const double radSum_s = radSum + separation;
    if (dist2 >= radSum_s * radSum_s || dist2 < 1e-12)
        return false;
    const double dist = sqrt(dist2);
    c.n = delta / dist;
    c.p1 = pos1 + c.n * r1;
    c.p2 = pos2 - c.n * r2;
    c.depth = dist - radSum;
    c.erad = r1 * r2 / radSum;
    return true;
}

// ChNarrowphasePRIMS.cpp:81-124
HD bool capsule_sphere(const d3& pos1, const q4& rot1, double r1, double hlen1, const d3& pos2, double r2,
                       double separation, Contact& c) {
    const d3 W = amat_w(rot1);
    double alpha = dot3(pos2 - pos1, W);
    alpha = clampd(alpha, -hlen1, hlen1);
    const d3 loc = pos1 + alpha * W;
    const double radSum = r1 + r2;
    const double radSum_s = radSum + separation;
    const d3 delta = pos2 - loc;
    const double dist2 = dot3(delta, delta);
    if (dist2 >= radSum_s * radSum_s || dist2 <= TINY_F)
        return false;
    const double dist = sqrt(dist2);
    c.n = delta / dist;
    c.p1 = loc + c.n * r1;
    c.p2 = pos2 - c.n * r2;
    c.depth = dist - radSum;
    c.erad = r1 * r2 / radSum;
    return true;
}

// ChCollisionUtils.h:494-516
HD unsigned snap_to_cylinder(double rad, double hlen, d3& loc) {
    unsigned code = 0;
    if (loc.z > hlen) {
        code |= 1;
        loc.z = hlen;
    } else if (loc.z < -hlen) {
        code |= 1;
        loc.z = -hlen;
    }
    const double d2 = loc.x * loc.x + loc.y * loc.y;
    if (d2 > rad * rad) {
        code |= 2;
        const double d = sqrt(d2);
        loc.x *= (rad / d);
        loc.y *= (rad / d);
    }
    return code;
}

// ChCollisionUtils.h:545-562
HD unsigned snap_to_box(const d3& h, d3& loc) {
    unsigned code = 0;
    if (fabs(loc.x) > h.x) {
        code |= 1;
        loc.x = (loc.x > 0) ? h.x : -h.x;
    }
    if (fabs(loc.y) > h.y) {
        code |= 2;
        loc.y = (loc.y > 0) ? h.y : -h.y;
    }
    if (fabs(loc.z) > h.z) {
        code |= 4;
        loc.z = (loc.z > 0) ? h.z : -h.z;
    }
    return code;
}

// ChNarrowphasePRIMS.cpp:134-189
HD bool cylinder_sphere(const d3& pos1, const q4& rot1, double r1, double hlen1, const d3& pos2, double r2,
                        double separation, Contact& c) {
    const d3 spherePos = parent_to_local(pos1, rot1, pos2);
    d3 cylPos = spherePos;
    const unsigned code = snap_to_cylinder(r1, hlen1, cylPos);
    if (code == 0)
        return false;
    const d3 delta = spherePos - cylPos;
    const double dist2 = dot3(delta, delta);
    const double r2_s = r2 + separation;
    if (dist2 >= r2_s * r2_s || dist2 <= TINY_F)
        return false;
    const double dist = sqrt(dist2);
    c.depth = dist - r2;
    c.n = rotate(delta / dist, rot1);
    c.p1 = local_to_parent(pos1, rot1, cylPos);
    c.p2 = pos2 - c.n * r2;
    if (code == 1)
        c.erad = r2;
    else if (code == 2)
        c.erad = r1 * r2 / (r1 + r2);
    else
        c.erad = r2 * EDGE_RADIUS / (r2 + EDGE_RADIUS);
    return true;
}

// ChNarrowphasePRIMS.cpp:200-260
HD bool roundedcyl_sphere(const d3& pos1, const q4& rot1, double r1, double hlen1, double srad1, const d3& pos2,
                          double r2, double separation, Contact& c) {
    const d3 spherePos = parent_to_local(pos1, rot1, pos2);
    d3 cylPos = spherePos;
    const unsigned code = snap_to_cylinder(r1, hlen1, cylPos);
    if (code == 0)
        return false;
    const double radSum = srad1 + r2;
    const double radSum_s = radSum + separation;
    const d3 delta = spherePos - cylPos;
    const double dist2 = dot3(delta, delta);
    if (dist2 >= radSum_s * radSum_s || dist2 <= TINY_F)
        return false;
    const double dist = sqrt(dist2);
    c.depth = dist - radSum;
    c.n = rotate(delta / dist, rot1);
    c.p2 = pos2 - c.n * r2;
    c.p1 = c.p2 - c.depth * c.n;
    if (code == 1)
        c.erad = r2;
    else if (code == 2)
        c.erad = r1 * r2 / (r1 + r2);
    else
        c.erad = srad1 * r2 / (srad1 + r2);
    return true;
}

// ChNarrowphasePRIMS.cpp:269-312
HD bool box_sphere(const d3& pos1, const q4& rot1, const d3& hdims1, const d3& pos2, double r2, double separation,
                   Contact& c) {
    const d3 spherePos = parent_to_local(pos1, rot1, pos2);
    d3 boxPos = spherePos;
    const unsigned code = snap_to_box(hdims1, boxPos);
    const d3 delta = spherePos - boxPos;
    const double dist2 = dot3(delta, delta);
    const double r2_s = r2 + separation;
    if (dist2 >= r2_s * r2_s || dist2 <= TINY_F)
        return false;
    const double dist = sqrt(dist2);
    c.depth = dist - r2;
    c.n = rotate(delta / dist, rot1);
    c.p1 = local_to_parent(pos1, rot1, boxPos);
    c.p2 = pos2 - c.n * r2;
    if ((code != 1) && (code != 2) && (code != 4))
        c.erad = r2 * EDGE_RADIUS / (r2 + EDGE_RADIUS);
    else
        c.erad = r2;
    return true;
}

// ChNarrowphasePRIMS.cpp:323-370
HD bool roundedbox_sphere(const d3& pos1, const q4& rot1, const d3& hdims1, double srad1, const d3& pos2, double r2,
                          double separation, Contact& c) {
    const d3 spherePos = parent_to_local(pos1, rot1, pos2);
    d3 boxPos = spherePos;
    const unsigned code = snap_to_box(hdims1, boxPos);
    const double radSum = srad1 + r2;
    const double radSum_s = radSum + separation;
    const d3 delta = spherePos - boxPos;
    const double dist2 = dot3(delta, delta);
    if (dist2 >= radSum_s * radSum_s || dist2 <= TINY_F)
        return false;
    const double dist = sqrt(dist2);
    c.depth = dist - radSum;
    c.n = rotate(delta / dist, rot1);
    c.p2 = pos2 - c.n * r2;
    c.p1 = c.p2 - c.depth * c.n;
    if ((code != 1) && (code != 2) && (code != 4))
        c.erad = r2 * srad1 / (r2 + srad1);
    else
        c.erad = r2;
    return true;
}

// ChCollisionUtils.h:467-474
HD d3 triangle_normal(const d3& A, const d3& B, const d3& C) {
    const d3 n = cross3(B - A, C - A);
    return n / length3(n);
}

// ChCollisionUtilsPRIMS.cpp:41-106 (Ericson's closest point on a triangle): true when `res` lies on the boundary
HD bool snap_to_triangle(const d3& A, const d3& B, const d3& C, const d3& P, d3& res) {
    const d3 AB = B - A, AC = C - A, AP = P - A;
    const double d1 = dot3(AB, AP), d2 = dot3(AC, AP);
    if (d1 <= 0 && d2 <= 0) {
        res = A;
        return true;
    }
    const d3 BP = P - B;
    const double d3_ = dot3(AB, BP), d4 = dot3(AC, BP);
    if (d3_ >= 0 && d4 <= d3_) {
        res = B;
        return true;
    }
    const double vc = d1 * d4 - d3_ * d2;
    if (vc <= 0 && d1 >= 0 && d3_ <= 0) {
        const double v = d1 / (d1 - d3_);
        res = A + v * AB;
        return true;
    }
    const d3 CP = P - C;
    const double d5 = dot3(AB, CP), d6 = dot3(AC, CP);
    if (d6 >= 0 && d5 <= d6) {
        res = C;
        return true;
    }
    const double vb = d5 * d2 - d1 * d6;
    if (vb <= 0 && d2 >= 0 && d6 <= 0) {
        const double w = d2 / (d2 - d6);
        res = A + w * AC;
        return true;
    }
    const double va = d3_ * d6 - d5 * d4;
    if (va <= 0 && (d4 - d3_) >= 0 && (d5 - d6) >= 0) {
        const double w = (d4 - d3_) / ((d4 - d3_) + (d5 - d6));
        res = B + w * (C - B);
        return true;
    }
    const double denom = 1 / (va + vb + vc);
    const double v = vb * denom;
    const double w = vc * denom;
    res = A + v * AB + w * AC;
    return false;
}

// ChNarrowphasePRIMS.cpp:379-434: one-sided (a sphere centre below the face plane gives no contact)
HD bool triangle_sphere(const d3& A1, const d3& B1, const d3& C1, const d3& pos2, double r2, double separation,
                        Contact& c) {
    const double r2_s = r2 + separation;
    const d3 nrm1 = triangle_normal(A1, B1, C1);
    const double h = dot3(pos2 - A1, nrm1);
    if (h >= r2_s || h <= 0)
        return false;
    d3 faceLoc;
    if (snap_to_triangle(A1, B1, C1, pos2, faceLoc)) {
        const d3 delta = pos2 - faceLoc;
        const double dist2 = dot3(delta, delta);
        if (dist2 >= r2_s * r2_s || dist2 <= TINY_F)
            return false;
        const double dist = sqrt(dist2);
        c.n = delta / dist;
        c.depth = dist - r2;
        c.erad = r2 * EDGE_RADIUS / (r2 + EDGE_RADIUS);
    } else {
        c.n = nrm1;
        c.depth = h - r2;
        c.erad = r2;
    }
    c.p1 = faceLoc;
    c.p2 = pos2 - c.n * r2;
    return true;
}

}  // namespace cb200
