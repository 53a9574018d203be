// common.cuh -- device buffers, error plumbing and the device-resident system state.
//
// Replaces the reference's struct-of-std::vector data contract (host_container, MC/ChDataManager.h:204-324, and
// ChCollisionData, CMC/ChCollisionData.h:89-222) with persistent device SoA arrays sized by high-water marks.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include <map>

#include "../../include/chrono_b200.h"
#include "vmath.cuh"

namespace cb200 {

extern thread_local std::string g_last_error;
int set_error(int code, const std::string& msg);

#define CB_CUDA(call)                                                                                   \
    do {                                                                                                \
        cudaError_t e__ = (call);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            return cb200::set_error(CB200_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); \
    } while (0)

#define CB_TRY(call)         \
    do {                     \
        int rc__ = (call);   \
        if (rc__ != CB200_OK) \
            return rc__;     \
    } while (0)

// Growable device array (cudaMalloc-backed; grows geometrically, never shrinks: the reference resize()s a
// std::vector every step, we keep the high-water allocation).
template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    int reserve(size_t n, bool keep = false, cudaStream_t s = 0) {
        if (n <= cap)
            return CB200_OK;
        size_t ncap = n + n / 4 + 64;
        T* np = nullptr;
        CB_CUDA(cudaMalloc(&np, ncap * sizeof(T)));
        if (keep && p && cap) {
            CB_CUDA(cudaMemcpyAsync(np, p, cap * sizeof(T), cudaMemcpyDeviceToDevice, s));
            CB_CUDA(cudaStreamSynchronize(s));
        }
        if (p)
            cudaFree(p);
        p = np;
        cap = ncap;
        return CB200_OK;
    }
    void release() {
        if (p)
            cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

// Per-shape broadphase record (64 B, two sectors): AABB (relative to global_origin once OffsetAABB has run,
// CMC/ChBroadphase.cpp:168-176) + everything the per-bin pair filter needs
// (CMC/ChCollisionUtilsBroadphase.cpp:86-146), so a pair test costs one gather per shape.
struct __align__(16) ShapeBP {
    double mn[3];
    double mx[3];
    uint32_t body;   // id_rigid (UINT_MAX = removed)
    uint32_t fam;    // group | mask << 16
    uint32_t info;   // bit0 body active, bit1 body collide, bits 8..15 shape type, bits 16..31 material id
    uint32_t pad;
};
static_assert(sizeof(ShapeBP) == 64, "ShapeBP must be 64 bytes");

// Per-pair side data written with the pair list so that the narrowphase never gathers per-shape integers.
struct __align__(16) PairAux {
    uint32_t bodyA, bodyB;
    uint32_t types;  // typeA | typeB << 8
    uint32_t mats;   // matA | matB << 16
};

// Device-side counters (one pinned-host mirror, read back at the few sync points of a step).
struct DevCounters {
    unsigned long long bbox[6];  // order-preserving encodings of min xyz / max xyz
    uint32_t n_large;            // shapes touching more than LARGE_BINS bins
    uint32_t n_intersections;    // N_i
    uint32_t n_active_bins;      // N_a
    uint32_t n_pairs;            // N_p
    uint32_t n_slots;            // pre-compaction contact slots
    uint32_t n_contacts;         // N_c
    uint32_t multi_slot;         // !=0 if some pair can yield more than one contact
    uint32_t n_heavy;            // bodies reduced by the masked path
    uint32_t ct_body_count;
    uint32_t overflow;           // a capacity was exceeded
    uint32_t unsupported;        // pairs whose shape combination this backend cannot test yet
    uint32_t pad;
};

struct Timer {
    cudaEvent_t a = nullptr, b = nullptr;
    double total_ms = 0;
    bool armed = false;
};

enum { LARGE_BINS = 48, MAX_LARGE = 4096, MAX_HEAVY = 64, SMC_MAT_STRIDE = 13, NSC_MAT_STRIDE = 8 };

struct System {
    int device = 0;
    cudaStream_t stream = nullptr;
    cb200_settings set;
    int64_t launches = 0;

    // ---- bodies (host_container state, MC/ChDataManager.h:228-262) ----
    uint32_t nb = 0;
    DevBuf<double> pos;      // 3 nb
    DevBuf<double> rot;      // 4 nb (w,x,y,z)
    DevBuf<double> vel;      // 6 nb
    DevBuf<double> hf;       // 6 nb
    DevBuf<double> fapp;     // 6 nb applied (non-gravity) force, torque; may stay zero
    bool has_fapp = false;
    DevBuf<double> mass;     // nb
    DevBuf<double> inertia;  // 3 nb (body-frame diagonal)
    DevBuf<double> inv_inertia;  // 3 nb
    DevBuf<uint8_t> bflags;  // bit0 active, bit1 collide
    DevBuf<uint8_t> heavy;   // body owns a large shape this step

    // ---- shapes (shape_container, CMC/ChCollisionData.h:35-59) ----
    uint32_t ns = 0;
    DevBuf<int32_t> s_type;
    DevBuf<uint32_t> s_body;
    DevBuf<uint32_t> s_fam;
    DevBuf<int32_t> s_mat;
    DevBuf<double> s_lpos;   // 3 ns  ObA
    DevBuf<double> s_lrot;   // 4 ns  ObR
    DevBuf<double> s_dims;   // 4 ns
    DevBuf<double> s_wpos;   // 4 ns  world position + dims[0]  (obj_data_A_global)
    DevBuf<double> s_wrot;   // 4 ns  world rotation            (obj_data_R_global)
    DevBuf<ShapeBP> s_bp;    // ns
    bool all_spheres_unit = false;

    // ---- materials: nm x nm composite table computed on the host in float (same arithmetic as the reference) ----
    int nm_smc = 0, nm_nsc = 0;
    DevBuf<double> mat_smc;  // nm*nm*SMC_MAT_STRIDE
    DevBuf<double> mat_nsc;

    // ---- broadphase (CMC/ChCollisionData.h:199-221) ----
    double origin[3], maxpt[3], bin_size[3], inv_bin_size[3];
    int bins[3] = {1, 1, 1};
    uint32_t num_bins = 0;
    DevBuf<uint32_t> bin_intersections;  // ns+1
    DevBuf<uint64_t> bin_count;          // num_bins+1: low 32 = entries in bin, high 32 = 1 if active (packed scan)
    DevBuf<uint32_t> bin_cursor;         // num_bins
    DevBuf<uint32_t> bin_entries_tmp;    // N_i  (unsorted inside a bin)
    DevBuf<uint32_t> bin_entries;        // N_i  (sorted by shape inside a bin) == bin_aabb_number
    DevBuf<uint32_t> large_list;         // MAX_LARGE
    DevBuf<uint32_t> pair_count;         // num_bins+1
    DevBuf<int64_t> pairs;               // N_p
    DevBuf<PairAux> pair_aux;            // N_p

    // ---- narrowphase ----
    DevBuf<uint32_t> slot_index;  // N_p+1 (contact_index)
    DevBuf<uint32_t> pair_nc;     // N_p+1 contacts found per pair -> exclusive scan
    DevBuf<double> t_norm, t_pta, t_ptb, t_depth, t_erad;  // pre-compaction slot arrays
    DevBuf<double> c_norm, c_pta, c_ptb, c_depth, c_erad;  // N_c
    DevBuf<int32_t> c_bids;      // 2 N_c
    DevBuf<int64_t> c_shapes;    // N_c
    DevBuf<uint32_t> c_mats;     // N_c
    uint32_t n_pairs = 0, n_slots = 0, n_contacts = 0, n_intersections = 0, n_active_bins = 0;

    // ---- SMC (MC/solver/ChIterativeSolverMulticoreSMC.cpp) ----
    DevBuf<double> ct_force;    // 3 N_c  force on body 2 (body 1 gets the negative)
    DevBuf<double> ct_torque;   // 6 N_c  torque on body 1, torque on body 2 (body frames)
    DevBuf<uint32_t> b_deg;     // nb+1 incidence counts -> offsets
    DevBuf<uint32_t> b_cursor;  // nb
    DevBuf<uint32_t> b_inc;     // 2 N_c incidence list: contact << 1 | side
    DevBuf<double> b_force;     // 6 nb  per-body contact force + torque
    DevBuf<uint32_t> heavy_list;
    DevBuf<double> heavy_part;

    // ---- NSC ----
    DevBuf<double> n_frame;   // per contact: U,V,W (9) + sbar_a (3) + sbar_b (3)
    DevBuf<double> gamma, n_r, n_E, n_work;  // 3 N_c each (work: several vectors)
    DevBuf<double> n_bodyvec;  // 6 nb scratch
    DevBuf<double> n_scal;     // small scalar block
    uint32_t nsc_iters = 0;
    double nsc_residual = 0, nsc_objective = 0;

    // ---- scan scratch / counters ----
    DevBuf<uint64_t> scan_tmp;
    DevCounters* d_cnt = nullptr;
    DevCounters* h_cnt = nullptr;  // pinned

    std::map<std::string, Timer> timers;
    int timed_steps = 0;
};

// ---- shared host helpers (defined in api.cu) ----
int timer_start(System* S, const char* name);
int timer_stop(System* S, const char* name);
int fetch_counters(System* S);  // D2H of DevCounters + stream sync

// ---- stage entry points ----
int stage_update_forces(System* S);   // integrate.cu
int stage_integrate(System* S);       // integrate.cu
int stage_aabb(System* S);            // broadphase.cu
int stage_broadphase(System* S);      // broadphase.cu
int stage_narrowphase(System* S);     // narrowphase.cu
int stage_smc(System* S);             // smc.cu
int stage_nsc(System* S);             // nsc.cu

// exclusive scans (scan.cu). n+1 outputs: out[n] = total. In-place allowed.
int scan_u32(System* S, const uint32_t* in, uint32_t* out, size_t n);
int scan_u64(System* S, const uint64_t* in, uint64_t* out, size_t n);

inline int blocks_for(size_t n, int threads) { return (int)((n + threads - 1) / threads); }

// Order-preserving map double <-> uint64 so that min/max reductions can use integer atomics (exact, order-free).
HD unsigned long long enc_ordered(double d) {
#ifdef __CUDA_ARCH__
    unsigned long long u = (unsigned long long)__double_as_longlong(d);
#else
    unsigned long long u;
    memcpy(&u, &d, 8);
#endif
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}
HD double dec_ordered(unsigned long long u) {
    u = (u & 0x8000000000000000ull) ? (u & 0x7fffffffffffffffull) : ~u;
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u);
#else
    double d;
    memcpy(&d, &u, 8);
    return d;
#endif
}

}  // namespace cb200
