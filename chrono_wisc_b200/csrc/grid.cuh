// grid.cuh -- the broadphase grid as the kernels see it, shared by broadphase.cu (binning) and pairs.cu (in-bin tests).
#pragma once
#include "common.cuh"

namespace cb200 {

struct GridParams {
    double origin[3];
    double inv[3];
    int bins[3];
    uint32_t num_bins;
};

__device__ __forceinline__ GridParams load_grid(const DevCounters* __restrict__ cnt) {
    GridParams g;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        g.origin[a] = cnt->origin[a];
        g.inv[a] = cnt->inv_bin_size[a];
        g.bins[a] = cnt->bins[a];
    }
    g.num_bins = cnt->num_bins;
    return g;
}

// origin-relative AABB of any shape: spheres from (centre, radius), others from their box record
__device__ __forceinline__ void shape_box_rel(uint32_t i, uint32_t info, const double* __restrict__ s_wpos,
                                              const ShapeBox* __restrict__ s_box, const double* origin,
                                              double envelope, double* mn, double* mx) {
    if (((info >> 8) & 0xff) == CB200_SPHERE) {
        const double4 w = reinterpret_cast<const double4*>(s_wpos)[i];
        const double r = w.w + envelope;
        mn[0] = (w.x - r) - origin[0], mn[1] = (w.y - r) - origin[1], mn[2] = (w.z - r) - origin[2];
        mx[0] = (w.x + r) - origin[0], mx[1] = (w.y + r) - origin[1], mx[2] = (w.z + r) - origin[2];
    } else {
        const double2* p = reinterpret_cast<const double2*>(s_box + i);
        const double2 a = p[0], b = p[1], c = p[2];
        mn[0] = a.x, mn[1] = a.y, mn[2] = b.x;
        mx[0] = b.y, mx[1] = c.x, mx[2] = c.y;
    }
}

}  // namespace cb200
