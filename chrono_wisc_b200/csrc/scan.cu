// scan.cu -- device-wide exclusive prefix sums (replaces the reference's Thrust_Exclusive_Scan / reduce_by_key
// calls, CMC/ChBroadphase.cpp:271,290-301,313 and CMC/ChNarrowphase.cpp:148).  Integer only, so the result is exact
// and independent of the reduction tree.
//
// The element count lives in device memory (*n_dev): the launch geometry is fixed by the capacity and the kernels
// walk the tiles grid-stride, so a scan can be enqueued before its extent is known on the host.
// One launch per scan (decoupled look-back, below); each tile is 512 threads x 16 consecutive items per
// thread, moved as 16-byte vectors.
#include "common.cuh"

namespace cb200 {

constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

struct LoadSame32 {
    typedef uint32_t in_t;
    typedef uint32_t out_t;
    __device__ __forceinline__ static uint32_t load(uint32_t v) { return v; }
};
struct LoadSame64 {
    typedef uint64_t in_t;
    typedef uint64_t out_t;
    __device__ __forceinline__ static uint64_t load(uint64_t v) { return v; }
};
// bin histogram: low word counts entries, high word counts active bins (run-length encoding of the sorted keys)
struct LoadBinPack {
    typedef uint32_t in_t;
    typedef uint64_t out_t;
    __device__ __forceinline__ static uint64_t load(uint32_t v) { return (uint64_t)v | ((uint64_t)(v != 0) << 32); }
};

template <typename T>
__device__ __forceinline__ T warp_incl_scan(T v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T o = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d)
            v += o;
    }
    return v;
}

// Block-wide exclusive scan of one value per thread; returns exclusive prefix, total in *total.
template <typename T>
__device__ __forceinline__ T block_excl_scan(T v, T* total, T* smem /* >= 33 entries */) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    T incl = warp_incl_scan(v, lane);
    if (lane == 31)
        smem[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        T w = (lane < (int)(blockDim.x >> 5)) ? smem[lane] : T(0);
        T wi = warp_incl_scan(w, lane);
        smem[lane] = wi - w;  // exclusive warp offsets
        if (lane == 31)
            smem[32] = wi;
    }
    __syncthreads();
    T off = smem[wid];
    *total = smem[32];
    __syncthreads();
    return off + incl - v;
}

// Single-pass scan with decoupled look-back: every tile publishes its aggregate (flag 1), then walks back over its
// predecessors' status words until it meets an inclusive prefix (flag 2), and publishes its own.  Tiles are handed
// out by an atomic ticket, so a tile only ever waits for tiles that are already running (no deadlock whatever the
// number of resident CTAs).  One launch, each input read once and each output written once.
struct TileState {           // 32 bytes
    uint32_t flag, pad;      // 0 nothing yet, 1 aggregate published, 2 inclusive prefix published
    uint64_t agg, incl, pad2;
};

// 16 consecutive items of a thread, moved as 16-byte vectors (a tile starts on a 16-byte boundary: the arrays come
// from cudaMalloc and a tile holds a multiple of 4 items)
template <typename V>
__device__ __forceinline__ void load_items(const V* __restrict__ p, size_t i0, size_t n, V* v) {
    constexpr int PER = 16 / sizeof(V);
    if (i0 + SCAN_ITEMS <= n) {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS / PER; k++) {
            const uint4 q = reinterpret_cast<const uint4*>(p + i0)[k];
            if (sizeof(V) == 4) {
                v[4 * k] = (V)q.x, v[4 * k + 1] = (V)q.y, v[4 * k + 2] = (V)q.z, v[4 * k + 3] = (V)q.w;
            } else {
                v[2 * k] = (V)((uint64_t)q.x | ((uint64_t)q.y << 32));
                v[2 * k + 1] = (V)((uint64_t)q.z | ((uint64_t)q.w << 32));
            }
        }
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++)
            v[k] = i0 + k < n ? p[i0 + k] : V(0);
    }
}
template <typename V>
__device__ __forceinline__ void store_items(V* __restrict__ p, size_t i0, size_t n, const V* v) {
    constexpr int PER = 16 / sizeof(V);
    if (i0 + SCAN_ITEMS <= n) {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS / PER; k++) {
            uint4 q;
            if (sizeof(V) == 4) {
                q = make_uint4((uint32_t)v[4 * k], (uint32_t)v[4 * k + 1], (uint32_t)v[4 * k + 2], (uint32_t)v[4 * k + 3]);
            } else {
                const uint64_t a = (uint64_t)v[2 * k], b = (uint64_t)v[2 * k + 1];
                q = make_uint4((uint32_t)a, (uint32_t)(a >> 32), (uint32_t)b, (uint32_t)(b >> 32));
            }
            reinterpret_cast<uint4*>(p + i0)[k] = q;
        }
    } else {
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++)
            if (i0 + k < n)
                p[i0 + k] = v[k];
    }
}

template <typename L>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_onepass(const typename L::in_t* in, typename L::out_t* out,
                                                              uint64_t* __restrict__ state_raw,
                                                              const uint32_t* __restrict__ n_dev,
                                                              const DevCounters* __restrict__ cnt) {
    typedef typename L::out_t T;
    typedef typename L::in_t I;
    __shared__ T sm[33];
    __shared__ uint32_t s_ticket;
    // Control block: [0] tile ticket, [1] CTAs done, [2] epoch.  Tile flags carry the epoch of the launch that wrote
    // them (flag = epoch << 2 | status), so records of earlier scans read as "not ready" and nothing has to be cleared
    // between scans: the last CTA to leave resets the ticket and advances the epoch.  Every CTA takes part in that
    // count, also one that skips its work, so the block stays consistent whatever the overflow mask does meanwhile.
    uint32_t* ticket = reinterpret_cast<uint32_t*>(state_raw);
    const uint32_t ep = reinterpret_cast<volatile uint32_t*>(ticket)[2] + 1u;
    const uint32_t F1 = ep << 2 | 1u, F2 = ep << 2 | 2u;
    const size_t n = *n_dev;
    __shared__ uint32_t s_skip;  // one answer per CTA: the mask may be raised by a side branch while this kernel runs
    if (threadIdx.x == 0)
        s_skip = cnt->overflow != 0u || n == 0;
    __syncthreads();
    const bool skip = s_skip != 0u;
    if (n == 0 && blockIdx.x == 0 && threadIdx.x == 0)
        out[0] = T(0);
    const uint32_t ntiles = (uint32_t)((n + SCAN_TILE - 1) / SCAN_TILE);
    volatile TileState* st = reinterpret_cast<volatile TileState*>(state_raw + 4);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    while (!skip) {
        if (threadIdx.x == 0)
            s_ticket = atomicAdd(ticket, 1u);
        __syncthreads();
        const uint32_t tile = s_ticket;
        if (tile >= ntiles)
            break;
        const size_t i0 = (size_t)tile * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
        I raw[SCAN_ITEMS];
        load_items<I>(in, i0, n, raw);
        T loc[SCAN_ITEMS];
        T s = 0;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            loc[k] = i0 + k < n ? L::load(raw[k]) : T(0);
            s += loc[k];
        }
        T total;
        T ex = block_excl_scan(s, &total, sm);
        if (threadIdx.x == 0) {
            st[tile].agg = (uint64_t)total;
            __threadfence();
            st[tile].flag = F1;
        }
        // Look-back by the whole CTA at once: thread i takes the aggregate of tile - 1 - i, thread 0 also the inclusive
        // prefix of the tile just before that window (only systems of more than SCAN_THREADS tiles have one).  The
        // predecessors were handed out earlier, so they are running and publish their aggregate before they wait for
        // anything: one polling round instead of a chain of warp-sized hops.
        T mine = 0;
        {
            const int idx = (int)tile - 1 - (int)threadIdx.x;
            if (idx >= 0) {
                uint32_t f;
                do {
                    f = st[idx].flag;
                } while (f != F1 && f != F2);  // anything else was left by an earlier scan
                __threadfence();
                mine = (T)st[idx].agg;
            }
            const int far = (int)tile - 1 - SCAN_THREADS;
            if (threadIdx.x == 0 && far >= 0) {
                while (st[far].flag != F2) {
                }
                __threadfence();
                mine += (T)st[far].incl;
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
            mine += __shfl_xor_sync(0xffffffffu, mine, d);
        if (lane == 0)
            sm[wid] = mine;
        __syncthreads();
        T carry = 0;
#pragma unroll
        for (int w = 0; w < SCAN_THREADS / 32; w++)
            carry += sm[w];
        if (threadIdx.x == 0) {
            st[tile].incl = (uint64_t)(carry + total);
            __threadfence();
            st[tile].flag = F2;
        }
        ex += carry;
        T o[SCAN_ITEMS];
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            o[k] = ex;
            ex += loc[k];
        }
        store_items<T>(out, i0, n, o);
        if (tile == ntiles - 1 && threadIdx.x == 0)
            out[n] = carry + total;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(ticket + 1, 1u) == gridDim.x - 1) {  // everybody else has left
            ticket[0] = 0u;
            ticket[1] = 0u;
            __threadfence();
            ticket[2] = ep;
        }
    }
}

template <typename L>
static int scan_impl(System* S, const typename L::in_t* in, typename L::out_t* out, const uint32_t* n_dev,
                     size_t n_cap) {
    size_t ntiles = (n_cap + SCAN_TILE - 1) / SCAN_TILE;
    if (ntiles == 0)
        ntiles = 1;
    const size_t words = 4 + 4 * (ntiles + 1);  // ticket block + one 32-byte status record per tile
    if (S->scan_tmp.cap < words) {
        if (S->capturing)
            return set_error(CB200_ERR_STATE, "scan scratch must be allocated before a step is captured");
        invalidate_step_graph(S);
        CB_CUDA(cudaStreamSynchronize(S->stream));
        CB_TRY(S->scan_tmp.reserve(words));
        CB_CUDA(cudaMemsetAsync(S->scan_tmp.p, 0, S->scan_tmp.cap * sizeof(uint64_t), S->stream));  // once: epoch 0
    }
    const int grid = grid_for(S, n_cap, SCAN_TILE, 2);
    k_scan_onepass<L><<<grid, SCAN_THREADS, 0, S->stream>>>(in, out, S->scan_tmp.p, n_dev, S->d_cnt);
    S->launches += 1;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

int scan_u32(System* S, const uint32_t* in, uint32_t* out, const uint32_t* n_dev, size_t n_cap) {
    return scan_impl<LoadSame32>(S, in, out, n_dev, n_cap);
}
int scan_u64(System* S, const uint64_t* in, uint64_t* out, const uint32_t* n_dev, size_t n_cap) {
    return scan_impl<LoadSame64>(S, in, out, n_dev, n_cap);
}
int scan_bins(System* S, const uint32_t* hist, uint64_t* out, const uint32_t* n_dev, size_t n_cap) {
    return scan_impl<LoadBinPack>(S, hist, out, n_dev, n_cap);
}

}  // namespace cb200
