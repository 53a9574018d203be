// scan.cu -- device-wide exclusive prefix sums (replaces the reference's Thrust_Exclusive_Scan / reduce_by_key
// calls, CMC/ChBroadphase.cpp:271,290-301,313 and CMC/ChNarrowphase.cpp:148).  Integer only, so the result is exact
// and independent of the reduction tree.
//
// The element count lives in device memory (*n_dev): the launch geometry is fixed by the capacity and the kernels
// walk the tiles grid-stride, so a scan can be enqueued before its extent is known on the host.
// Three launches (tile reduce -> scan of tile sums in one CTA -> tile scan with carry-in); each tile is
// 256 threads x 16 items, staged through shared memory so that global loads and stores are coalesced.
#include "common.cuh"

namespace cb200 {

constexpr int SCAN_THREADS = 256;
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

template <typename L>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_reduce(const typename L::in_t* __restrict__ in,
                                                             typename L::out_t* __restrict__ tile_sums,
                                                             const uint32_t* __restrict__ n_dev,
                                                             const DevCounters* __restrict__ cnt) {
    typedef typename L::out_t T;
    __shared__ T sm[33];
    if (cnt->overflow)
        return;
    const size_t n = *n_dev;
    const size_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const size_t base = tile * SCAN_TILE;
        T s = 0;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            size_t i = base + (size_t)k * SCAN_THREADS + threadIdx.x;
            if (i < n)
                s += L::load(in[i]);
        }
        T total;
        block_excl_scan(s, &total, sm);
        if (threadIdx.x == 0)
            tile_sums[tile] = total;
    }
}

template <typename T>
__global__ void __launch_bounds__(1024) k_scan_tiles(T* __restrict__ tile_sums, const uint32_t* __restrict__ n_dev,
                                                     T* __restrict__ out, const DevCounters* __restrict__ cnt) {
    __shared__ T sm[33];
    if (cnt->overflow)
        return;
    const size_t n = *n_dev;
    const size_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    T carry = 0;
    for (size_t base = 0; base < ntiles; base += blockDim.x) {
        size_t i = base + threadIdx.x;
        T v = (i < ntiles) ? tile_sums[i] : T(0);
        T total;
        T ex = block_excl_scan(v, &total, sm);
        if (i < ntiles)
            tile_sums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0)
        out[n] = carry;
}

template <typename L>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(const typename L::in_t* __restrict__ in,
                                                            typename L::out_t* __restrict__ out,
                                                            const typename L::out_t* __restrict__ tile_offsets,
                                                            const uint32_t* __restrict__ n_dev,
                                                            const DevCounters* __restrict__ cnt) {
    typedef typename L::out_t T;
    __shared__ T sm[33];
    __shared__ T stage[SCAN_TILE];
    if (cnt->overflow)
        return;
    const size_t n = *n_dev;
    const size_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const size_t base = tile * SCAN_TILE;
        // coalesced load into shared, then each thread owns SCAN_ITEMS consecutive items
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            size_t i = base + (size_t)k * SCAN_THREADS + threadIdx.x;
            stage[k * SCAN_THREADS + threadIdx.x] = (i < n) ? L::load(in[i]) : T(0);
        }
        __syncthreads();
        T loc[SCAN_ITEMS];
        T s = 0;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            loc[k] = stage[threadIdx.x * SCAN_ITEMS + k];
            s += loc[k];
        }
        T total;
        T ex = block_excl_scan(s, &total, sm) + tile_offsets[tile];
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            T v = loc[k];
            stage[threadIdx.x * SCAN_ITEMS + k] = ex;
            ex += v;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            size_t i = base + (size_t)k * SCAN_THREADS + threadIdx.x;
            if (i < n)
                out[i] = stage[k * SCAN_THREADS + threadIdx.x];
        }
        __syncthreads();
    }
}

template <typename L>
static int scan_impl(System* S, const typename L::in_t* in, typename L::out_t* out, const uint32_t* n_dev,
                     size_t n_cap) {
    typedef typename L::out_t T;
    size_t ntiles = (n_cap + SCAN_TILE - 1) / SCAN_TILE;
    if (ntiles == 0)
        ntiles = 1;
    CB_TRY(S->scan_tmp.reserve(ntiles + 2));
    T* tiles = reinterpret_cast<T*>(S->scan_tmp.p);
    const int grid = grid_for(S, n_cap, SCAN_TILE, 4);
    k_scan_reduce<L><<<grid, SCAN_THREADS, 0, S->stream>>>(in, tiles, n_dev, S->d_cnt);
    k_scan_tiles<T><<<1, 1024, 0, S->stream>>>(tiles, n_dev, out, S->d_cnt);
    k_scan_apply<L><<<grid, SCAN_THREADS, 0, S->stream>>>(in, out, tiles, n_dev, S->d_cnt);
    S->launches += 3;
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
