// scan.cu -- device-wide exclusive prefix sums (replaces the reference's Thrust_Exclusive_Scan calls,
// CMC/ChBroadphase.cpp:271,301,313 and CMC/ChNarrowphase.cpp:148).  Integer only, so the result is exact and
// independent of the reduction tree.
//
// Three launches (tile reduce -> scan of tile sums in one CTA -> tile scan with carry-in); each tile is
// 256 threads x 16 items, loads are coalesced 16-item-strided through a warp-transposed register layout.
#include "common.cuh"

namespace cb200 {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

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
__device__ __forceinline__ T block_excl_scan(T v, T* total, T* smem /* >= 32 entries */) {
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

template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_reduce(const T* __restrict__ in, T* __restrict__ tile_sums,
                                                             size_t n) {
    __shared__ T sm[33];
    size_t base = (size_t)blockIdx.x * SCAN_TILE;
    T s = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        size_t i = base + (size_t)k * SCAN_THREADS + threadIdx.x;
        if (i < n)
            s += in[i];
    }
    T total;
    block_excl_scan(s, &total, sm);
    if (threadIdx.x == 0)
        tile_sums[blockIdx.x] = total;
}

template <typename T>
__global__ void __launch_bounds__(1024) k_scan_tiles(T* __restrict__ tile_sums, size_t ntiles, T* __restrict__ total_out) {
    __shared__ T sm[33];
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
        *total_out = carry;
}

template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(const T* __restrict__ in, T* __restrict__ out,
                                                            const T* __restrict__ tile_offsets, size_t n) {
    __shared__ T sm[33];
    __shared__ T stage[SCAN_TILE];
    const size_t base = (size_t)blockIdx.x * SCAN_TILE;
    // coalesced load into shared, then each thread owns SCAN_ITEMS consecutive items
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        size_t i = base + (size_t)k * SCAN_THREADS + threadIdx.x;
        stage[k * SCAN_THREADS + threadIdx.x] = (i < n) ? in[i] : T(0);
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
    T ex = block_excl_scan(s, &total, sm) + tile_offsets[blockIdx.x];
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
}

template <typename T>
static int scan_impl(System* S, const T* in, T* out, size_t n) {
    size_t ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (ntiles == 0)
        ntiles = 1;
    CB_TRY(S->scan_tmp.reserve(ntiles + 2));
    T* tiles = reinterpret_cast<T*>(S->scan_tmp.p);
    k_scan_reduce<T><<<(unsigned)ntiles, SCAN_THREADS, 0, S->stream>>>(in, tiles, n);
    k_scan_tiles<T><<<1, 1024, 0, S->stream>>>(tiles, ntiles, out + n);
    k_scan_apply<T><<<(unsigned)ntiles, SCAN_THREADS, 0, S->stream>>>(in, out, tiles, n);
    S->launches += 3;
    CB_CUDA(cudaGetLastError());
    return CB200_OK;
}

int scan_u32(System* S, const uint32_t* in, uint32_t* out, size_t n) { return scan_impl<uint32_t>(S, in, out, n); }
int scan_u64(System* S, const uint64_t* in, uint64_t* out, size_t n) { return scan_impl<uint64_t>(S, in, out, n); }

}  // namespace cb200
