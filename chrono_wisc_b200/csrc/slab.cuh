// slab.cuh -- what the pack pass of a slab exchange leaves behind for later readers (slab.cu, comm.cu)
//
// Per owned slot one flag byte (SLAB_F_LEFT: its record went to the left neighbour, SLAB_F_RIGHT: to the right one,
// SLAB_F_VACANT: the slot is free after this exchange) and, per CTA of SLAB_CTA slots, the exclusive offsets of the three
// counts (System::slab_cta: 3 per CTA, the totals behind the last CTA).  The j-th record sent to a side -- j = CTA
// offset + rank inside the CTA, in slot order -- is the j-th ghost slot of that neighbour's region for this rank.
#pragma once
#include <stdint.h>

namespace cb200 {

enum { SLAB_F_LEFT = 1, SLAB_F_RIGHT = 2, SLAB_F_VACANT = 4, SLAB_CTA = 256 };

// Ranks of this thread's slot among the slots of each kind, in slot order.  Must be called by every thread of a CTA of
// SLAB_CTA threads whose thread t handles slot blockIdx.x * SLAB_CTA + t; wtot: 3 x 8 words of shared memory.
__device__ __forceinline__ void slab_ranks(uint32_t f, const uint32_t* __restrict__ cta_off, uint32_t (*wtot)[8],
                                           uint32_t rank[3]) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const unsigned m = __ballot_sync(0xffffffffu, (f >> k) & 1u);
        rank[k] = (uint32_t)__popc(m & ((1u << lane) - 1u));
        if (lane == 0)
            wtot[k][wid] = (uint32_t)__popc(m);
    }
    __syncthreads();
    if (!(f & 7u))
        return;  // only a flagged slot's ranks are ever used (a few per CTA: most warps are done here)
#pragma unroll
    for (int k = 0; k < 3; k++) {
        for (int w = 0; w < wid; w++)
            rank[k] += wtot[k][w];
        rank[k] += cta_off[3 * blockIdx.x + k];
    }
}

}  // namespace cb200
