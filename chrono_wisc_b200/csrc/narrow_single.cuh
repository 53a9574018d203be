// narrow_single.cuh -- shape views and the analytic dispatcher for pairs with at most one contact (anything against a
// sphere), shared by the pair kernels (pairs.cu: decision only, contact records) and the SMC force kernel (smc.cu:
// contact geometry evaluated in place).  Reference: PRIMSCollision, CMC/ChNarrowphasePRIMS.cpp:1406-1607.
#pragma once
#include "common.cuh"
#include "mpr.cuh"
#include "prims.cuh"

namespace cb200 {

__device__ __forceinline__ ShapeView make_view(uint32_t id, int type, const double4& w,
                                               const double* __restrict__ s_wrot, const double* __restrict__ s_dims) {
    ShapeView v;
    v.type = type;
    v.pos = mk3(w.x, w.y, w.z);
    v.d0 = w.w;
    v.d1 = v.d2 = v.d3_ = 0;
    v.hull = nullptr, v.nv = 0;  // a hull never has a single-contact analytic test: it is seen by the complex path only
    v.rot = mkq(1, 0, 0, 0);
    if (type != CB200_SPHERE) {
        const double4 r = reinterpret_cast<const double4*>(s_wrot)[id];
        v.rot = mkq(r.x, r.y, r.z, r.w);
        const double4 d = reinterpret_cast<const double4*>(s_dims)[id];
        v.d1 = d.y, v.d2 = d.z, v.d3_ = d.w;
    }
    return v;
}

// Analytic dispatcher for pairs with at most one contact (everything that involves a sphere),
// PRIMSCollision, CMC/ChNarrowphasePRIMS.cpp:1406-1607.
// Returns 1/0 contacts, or -1 when the combination is not of that kind (multi-contact analytic pair or MPR).
// Callers that only use the return value get the decision alone: the compiler drops the sqrt/divisions that
// feed nothing but the unused Contact.
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
            case CB200_TRIANGLE:
                return triangle_sphere(tri_vertex(A, 0), tri_vertex(A, 1), tri_vertex(A, 2), B.pos, B.d0, sep, c) ? 1 : 0;
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
            case CB200_TRIANGLE:
                hit = triangle_sphere(tri_vertex(B, 0), tri_vertex(B, 1), tri_vertex(B, 2), A.pos, A.d0, sep, c);
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

}  // namespace cb200
