// common.cuh -- device buffers, error plumbing and the device-resident system state.
//
// Replaces the reference's struct-of-std::vector data contract (host_container, MC/ChDataManager.h:204-324, and
// ChCollisionData, CMC/ChCollisionData.h:89-222) with persistent device SoA arrays.
//
// Execution model ("device-driven step"): every count the reference keeps on the host (N_i, N_a, N_p, N_c, the grid)
// lives in one DevCounters block in device memory.  Kernels read their extents from it and run grid-stride, so a whole
// DoStepDynamics is enqueued without a single host synchronisation.  Output arrays have capacities; a kernel that
// would exceed one raises the sticky `overflow` mask and records the size it needed, every later kernel of the
// stream then returns immediately (state is never advanced with incomplete lists), and the host -- at its one sync
// at the end of cb200_do_steps or of a stage call -- grows the arrays and re-enqueues the steps that did not run.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include <map>
#include <utility>

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

// Per-shape AABB record (48 B): absolute after the AABB stage, relative to global_origin once the binning has run
// (OffsetAABB, CMC/ChBroadphase.cpp:168-176).
struct __align__(16) ShapeBox {
    double mn[3];
    double mx[3];
};

// Per-shape integer record (16 B) holding everything the per-bin pair filter needs
// (CMC/ChCollisionUtilsBroadphase.cpp:86-146), so a pair test costs one gather per shape.
struct __align__(16) ShapeMeta {
    uint32_t body;   // id_rigid (UINT_MAX = removed)
    uint32_t fam;    // group | mask << 16
    uint32_t info;   // bit0 body active, bit1 body collide, bits 8..15 shape type, bits 16..31 material id
    uint32_t pad;
};

// overflow mask bits
enum { OVF_BINS = 1, OVF_ENTRIES = 2, OVF_CONTACTS = 4, OVF_PAIRS = 8, OVF_LARGE = 16, OVF_HEAVY = 32, OVF_COMPLEX = 64,
       OVF_NSC = 128, OVF_SLAB = 256, OVF_PAIRQ = 512 };

// Device-side counters and the broadphase grid (one pinned-host mirror, read back at the single sync of a call).
struct DevCounters {
    unsigned long long bbox[6];  // order-preserving encodings of min xyz / max xyz of the colliding shapes' AABBs
    // grid of this step, ChBroadphase::DetermineBoundingBox/ComputeTopLevelResolution (CMC/ChBroadphase.cpp:142-208)
    double origin[3], maxpt[3], bin_size[3], inv_bin_size[3];
    int32_t bins[3];
    uint32_t num_bins;
    uint32_t n_large;            // shapes touching more than LARGE_BINS bins
    uint32_t n_intersections;    // N_i
    uint32_t n_active_bins;      // N_a
    uint32_t n_tiles;            // ceil(N_i / PAIR_TILE)
    uint32_t n_pairs;            // N_p
    uint32_t n_contacts;         // N_c
    uint32_t n_complex;          // candidate pairs that need the multi-contact / MPR path
    uint32_t n_heavy;            // bodies reduced by the masked path
    uint32_t ct_body_count;
    uint32_t unsupported;        // pairs whose shape combination this backend cannot test yet
    uint32_t overflow;           // sticky OVF_* mask; while set every kernel is a no-op
    uint32_t need_bins, need_entries, need_contacts, need_pairs, need_complex;
    uint32_t steps_done;         // steps fully applied since the host last reset it
    uint32_t nsc_iters;
    uint32_t nsc_slots;          // APGD slots the last solve used: iterations + back-tracking retries
    uint32_t n_bodies, n_shapes; // extents of body / shape arrays (for device-count scans)
    uint32_t need_nsc_slots;     // APGD slot budget that would have sufficed (OVF_NSC)
    double nsc_residual, nsc_objective;
    uint32_t nsc_hist_count;     // iterations recorded in the residual history (includes the one that broke out)
    uint32_t n_cslots;           // contact slots reserved for the complex pairs (PreprocessCount's rule per pair)
    uint32_t need_cslots;
    uint32_t n_slow;             // bins queued for k_pair_masks_slow
    uint32_t n_long;             // ... of which have more than 64 entries (kept right behind n_slow: one memset)
    uint32_t n_pairq;            // (wall-bin pairs with a non-sphere) queued for k_pair_queue (right behind n_long)
    uint32_t need_pairq;
    uint32_t n_cls[4];           // complex pairs per narrowphase class (k_cplx_collect buckets them for k_cplx_narrow)
    uint32_t n_bb[5];            // box-box pairs per branch of box_box (right behind n_cls: cleared together)
};

struct Timer {
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pool;  // event pairs, resolved lazily by flush_timers
    size_t used = 0;
    double total_ms = 0;
    long steps = 0;  // simulation steps the accumulated intervals cover
    bool armed = false;
};

enum { CT_STRIDE = 12, LARGE_BINS = 48, MAX_LARGE = 4096, MAX_HEAVY = 64, SMC_MAT_STRIDE = 14, NSC_MAT_STRIDE = 8, PAIR_TILE = 256,
       PAIR_HALO = 64 };

struct System {
    int device = 0;
    int num_sms = 148;
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
    DevBuf<double> vlim;     // 2 nb (max_speed, max_wvel) of ChBody::SetLimitSpeed; <= 0: no limit
    bool has_vlim = false;
    // ChCollisionSystemMulticore::SetActiveBoundingBox (use_aabb_active, CMC/ChCollisionSystemMulticore.cpp:289-304)
    bool use_aabb_active = false;
    double active_min[3] = {0, 0, 0}, active_max[3] = {0, 0, 0};
    DevBuf<uint8_t> b_inbox;  // body has a shape whose box overlaps the active box
    DevBuf<double> mass;     // nb
    DevBuf<double> inertia;  // 3 nb (body-frame diagonal)
    DevBuf<double> inv_inertia;  // 3 nb
    // products of inertia (use_full_inertia_tensor, MC/ChSettings.h:107): (xy, xz, yz) of the inertia matrix and of its
    // inverse, 3 nb each; null until cb200_set_inertia_full is called (bodies given principal moments only)
    DevBuf<double> inertia_off, inv_inertia_off;
    DevBuf<uint8_t> bflags;  // bit0 active, bit1 collide, bit2 vacant slot (slab decomposition: no body lives here)
    DevBuf<uint8_t> heavy;   // body owns a large shape this step
    // a body's own centred sphere (the first sphere shape with ObA == 0 added for it; -1: none) and its radius: the
    // integration kernel writes that shape's world record and box itself (integrate.cu, k_integrate<.., true>)
    DevBuf<int32_t> b_sphere;
    DevBuf<double> b_rad;
    std::vector<int32_t> h_b_sphere;
    std::vector<double> h_b_rad;
    bool sphere_aabb_valid = false;  // s_wpos of those spheres and their share of the global box match `pos`
    bool fuse_aabb = false;          // whole-step paths: the integration that ends this step prepares the next one

    // ---- shapes (shape_container, CMC/ChCollisionData.h:35-59) ----
    uint32_t ns = 0;
    DevBuf<int32_t> s_type;
    DevBuf<uint32_t> s_body;
    DevBuf<uint32_t> s_fam;
    DevBuf<int32_t> s_mat;
    DevBuf<double> s_lpos;   // 3 ns  ObA
    DevBuf<double> s_lrot;   // 4 ns  ObR
    DevBuf<double> s_dims;   // 4 ns
    DevBuf<uint32_t> general_ids;  // shapes that are not body-centred spheres (k_aabb<true> walks this list)
    uint32_t n_general = 0;
    DevBuf<double> s_verts;  // 3 n_verts: vertex pool of triangles (body frame) and convex hulls (shape frame)
    uint32_t n_verts = 0;
    DevBuf<double> s_wpos;   // 4 ns  world position + dims[0]  (obj_data_A_global)
    DevBuf<double> s_wpos_alt;  // the records of the NEXT step's own spheres, written by the integration of a whole
                                // step; the collision stage's records stay readable until the next step starts
    DevBuf<double> s_wrot;   // 4 ns  world rotation            (obj_data_R_global)
    DevBuf<ShapeBox> s_box;  // ns    AABB
    DevBuf<ShapeMeta> s_meta;  // ns
    bool meta_dirty = true;    // body flags / shapes changed: rebuild s_meta
    uint32_t n_nonsphere = 0;  // shapes that are not spheres (host count)
    uint32_t nonsphere_bodies = 0, nonsphere_body0 = 0;  // 0 / 1 / 2 (= two or more) bodies own non-sphere shapes

    // ---- materials: nm x nm composite table computed on the host in float (same arithmetic as the reference) ----
    int nm_smc = 0, nm_nsc = 0;
    DevBuf<double> mat_smc;  // nm*nm*SMC_MAT_STRIDE
    DevBuf<double> mat_nsc;
    DevBuf<double> mat_ovr;  // per-contact composite records of the current contact list (AddContactCallback), or unused
    bool mat_ovr_valid = false;

    // ---- broadphase (CMC/ChCollisionData.h:199-221); capacities cap_* are what the kernels check against ----
    uint32_t cap_bins = 0, cap_entries = 0, cap_contacts = 0, cap_pairs = 0;
    DevBuf<uint32_t> bin_intersections;  // ns
    DevBuf<uint64_t> bin_start;          // cap_bins+1: low 32 = first entry of bin, high 32 = index among active bins
    DevBuf<uint32_t> bin_cursor;         // cap_bins (histogram, then countdown cursor of the scatter)
    DevBuf<uint32_t> ent_tmp;            // cap_entries  shape ids, unsorted inside a bin
    DevBuf<uint32_t> ent_bin;            // cap_entries  bin of every entry
    DevBuf<uint32_t> ent_shape;          // cap_entries  shape ids sorted inside a bin == bin_aabb_number (written by the pair kernels)
    DevBuf<uint64_t> ent_mask;           // cap_entries  bit j set: the entry pairs with the j-th entry of its bin
    DevBuf<uint64_t> ent_mask_hi;        // cap_entries  bits 64..127 of the same, for bins of 65..128 entries (k_long_rows)
    DevBuf<uint32_t> ent_off;            // cap_entries+2 outputs per entry -> exclusive offsets
    DevBuf<uint32_t> slow_bins;          // cap_entries  first entries of the bins the warp-per-bin pair kernel takes
    DevBuf<uint32_t> long_bins;          // cap_entries / 64 + 1  ... of those with more than 64 entries
    DevBuf<uint64_t> pair_q;             // cap_pairq  (entry, entry) pairs of wall bins that involve a non-sphere
    uint32_t cap_pairq = 0;
    DevBuf<uint32_t> large_list;         // MAX_LARGE
    DevBuf<int64_t> pairs;               // cap_pairs (materialised on demand only)
    bool pairs_valid = false;
    bool contacts_geom_valid = false;  // c_norm/c_pta/c_ptb/c_depth/c_erad hold the records of the current key list
    bool box_relative = false;  // s_box holds origin-relative boxes (the binning ran since the last AABB stage)

    // ---- narrowphase output: the contact list (CMC/ChCollisionData.h:159-183) ----
    DevBuf<double> c_norm, c_pta, c_ptb, c_depth, c_erad;  // cap_contacts
    DevBuf<int32_t> c_bids;      // 2 per contact
    DevBuf<int64_t> c_shapes;    // 1 per contact
    DevBuf<uint32_t> c_mats;     // 1 per contact

    // ---- multi-contact / MPR pairs ("complex" pairs: no sphere involved, or algorithm == MPR) ----
    uint32_t cap_complex = 0, cap_cslots = 0;
    DevBuf<uint64_t> ent_cmask;  // cap_entries  bit j set: the pair (entry, j-th entry of its bin) needs the complex path
    DevBuf<uint32_t> ent_cbase;  // cap_entries  first complex record of the entry
    DevBuf<uint32_t> cplx_rec;   // 4 per record: entry, shape A, shape B, first slot
    DevBuf<uint32_t> cplx_n;     // contacts found per record
    DevBuf<uint32_t> cplx_cls;   // 4 x cap_complex: record indices bucketed by narrowphase class
    DevBuf<uint32_t> cplx_bb;    // 5 x cap_complex: box-box records bucketed by the branch of box_box they take
    DevBuf<double> cplx_out;     // 11 per slot: normal, point A, point B, depth, effective radius
    DevBuf<uint32_t> c_src;      // cap_contacts  slot + 1 of a contact computed by the complex path, 0 otherwise

    // ---- SMC (MC/solver/ChIterativeSolverMulticoreSMC.cpp) ----
    // CT_STRIDE N_c: per contact (force on body 2 [body 1 gets the negative], pad) (torque on body 1, pad) (torque on
    // body 2, pad), torques in the body frames -- three 32-byte sectors
    DevBuf<double> ct_force;
    // ---- bilateral rows (bilateral.cu) ----
    uint32_t n_bil = 0;
    int bil_max_iter = 100;
    double bil_tol = 1e-4;
    DevBuf<int32_t> bil_ids;   // 2 per row
    DevBuf<double> bil_cq;     // 12 per row: Cq_a | Cq_b
    DevBuf<double> bil_b, bil_R, bil_gamma, bil_N, bil_vec, bil_info;
    DevBuf<double> report;      // 21 per contact: the batch of cb200_report_contacts (report.cu)
    DevBuf<double> ct_torque;   // unused (kept for the release list)
    // MultiStep tangential displacement history, 20 slots per body (allocated on first use, smc.cu)
    DevBuf<int32_t> sh_neigh;   // 3 x 20 nb
    DevBuf<uint8_t> sh_touch;   // 20 nb
    DevBuf<double> sh_disp;     // 3 x 20 nb
    DevBuf<double> sh_relvel;   // 20 nb
    DevBuf<double> sh_dur;      // 20 nb
    uint32_t sh_bodies = 0;     // bodies the history arrays are initialised for
    DevBuf<uint32_t> b_deg;     // nb+1 incidence counts -> offsets
    DevBuf<uint32_t> b_cursor;  // nb
    DevBuf<uint32_t> b_inc;     // 2 N_c incidence list: contact << 1 | side
    DevBuf<double> b_force;     // 6 nb  per-body contact force + torque
    DevBuf<uint32_t> heavy_list;
    DevBuf<double> heavy_part;

    // ---- NSC (nsc.cu) ----
    DevBuf<double> n_frame;   // 10 per contact: U (3), world arms s_A, s_B (3 + 3), friction
    DevBuf<double> n_aux;     // 8 per contact: cohesion, E_n, E_t, mu_roll, mu_spin, E_spin, E_roll (only if nsc_has_aux)
    DevBuf<double> gamma, n_gamma_hat, n_y, n_gn, n_g, n_rfull;  // rows per contact each, interleaved (n,u,v[,tn,tu,tv])
    DevBuf<double> n_w, n_w0, n_mk;  // 6 nb: M^-1 D x (world angular part), M_invk world / body frame
    DevBuf<double> n_wy, n_wg;       // 6 nb: M^-1 D y and M^-1 D gamma, carried along by linearity (nsc.cu nsc_phase)
    DevBuf<double> n_part;    // per-CTA partial sums of the fixed-shape reductions
    DevBuf<double> n_hist;    // (residual, objective) per APGD iteration of the last solve
    DevBuf<double> n_rows;    // D_T rows on request (parity path)
    void* n_ctl = nullptr;    // NscCtl
    int nsc_rows_alloc = 0;   // rows per contact the vectors are sized for
    int nsc_has_aux = 0;      // some material pair has cohesion / compliance / rolling / spinning friction
    int nsc_extra_slots = 0;  // APGD slot budget beyond max_iter * 1.5 + 16 (grown on OVF_NSC)
    int nsc_hist_cap = 0;
    bool nsc_fc_current = false;  // b_force also holds the contact forces on inactive bodies (Fc_current)

    // ---- slab decomposition (slab.cu, comm.cu) ----
    struct Comm* comm = nullptr;    // communicator / local group this slab belongs to (comm.cu); null: single system
    DevBuf<uint8_t> c_own;          // NSC: contact enters the global sums on this rank (its first dynamic body is owned)
    bool slab_on = false;
    int slab_axis = 0, slab_has_left = 0, slab_has_right = 0;
    double slab_lo = 0, slab_hi = 0, slab_w = 0;
    uint32_t slab_nb_static = 0, slab_ns_static = 0, slab_owned_cap = 0, slab_ghost_cap = 0;
    DevBuf<int64_t> b_gid;          // global body id per slot
    DevBuf<uint64_t> slab_flags;    // owned_cap BYTES: per owned slot the flags of the last pack (slab.cuh)
    DevBuf<uint32_t> slab_cta;      // 3 per CTA of 256 slots: counts, then exclusive offsets; totals behind the last
    DevBuf<uint32_t> slab_vacant;   // vacant owned slots in slot order
    uint32_t* slab_cnt = nullptr;   // device: [0] owned_cap (scan extent) [1] n_vacant [2] arrivals of last unpack
    cudaStream_t own_stream = nullptr;  // the stream cb200_create made (stream may be the caller's, cb200_set_stream)
    // Side branches of a step (api.cu aux_fork / aux_join): short latency-bound kernels that do not depend on each other
    // (wall-bin pairs | slow bins | long bins; container-body sums | per-body sums) run on two auxiliary streams
    // between an event fork and an event join -- parallel branches of the step graph when the step is captured.
    cudaStream_t aux[2] = {nullptr, nullptr};
    cudaEvent_t aux_ev[32] = {};
    int aux_ev_made = 0, aux_ev_next = 0;
    bool aux_enabled = true;
    // Scratch that a completed step leaves zeroed, so the next one need not clear it (one graph node less each): the
    // bin histogram is counted back down to zero by the scatter; the integration of a whole step clears the body
    // degrees, the incidence cursors and the heavy flags.  Anything that can leave them dirty (an overflow, a
    // reallocation, a stage call) resets the flag and the next enqueue clears them the slow way.
    bool bins_clean = false, body_scratch_clean = false, pair_cnt_clean = false;
    bool whole_step = false;       // inside enqueue_one_step (api.cu): the integration that ends it may clear scratch
    bool bbox_reset_done = false;  // k_reset_bbox of this step already ran on a side branch (smc.cu)

    // ---- scan scratch / counters ----
    DevBuf<uint64_t> scan_tmp;
    DevCounters* d_cnt = nullptr;
    DevCounters* h_cnt = nullptr;  // pinned

    std::map<std::string, Timer> timers;
    int timed_steps = 0;
    bool defer_update_forces = false;  // whole-step paths (SMC): hf is formed in k_body_reduce<true>, k_update_forces is skipped
    bool defer_vel_update = false;   // whole-step paths (SMC): v += M^-1 hf is done by k_integrate<true>, not k_vel_update
    bool capturing = false;          // a step is being captured into a CUDA graph: timers are skipped
    // Two graphs: the world records of the bodies' own spheres are double-buffered (s_wpos / s_wpos_alt swap every
    // whole step, see integrate.cu), so consecutive steps capture different pointers.
    cudaGraphExec_t step_graph[2] = {nullptr, nullptr};
    int step_graph_launches[2] = {0, 0};  // kernel / memset nodes of one captured step
    bool step_graph_meta[2] = {false, false};
    int wpos_parity = 0;             // flips with every s_wpos / s_wpos_alt swap
    int direct_steps = 0;            // steps enqueued directly since the graph was last invalidated
    unsigned async_calls = 0;
    cudaStream_t copy_stream = nullptr;  // second copy engine queue of cb200_step_host
    cudaEvent_t ev_copy = nullptr, ev_main = nullptr;
};

// multi-contact / MPR pairs can occur: two or more non-sphere shapes, or MPR requested for everything
inline bool use_complex(const System* S) { return S->nonsphere_bodies >= 2 || S->set.narrowphase_algorithm == 0; }

// ---- shared host helpers (defined in api.cu) ----
int timer_start(System* S, const char* name);
int timer_stop(System* S, const char* name, int steps = 1);
int flush_timers(System* S);
void invalidate_step_graph(System* S);  // any change to what a step's launches capture (pointers, sizes, settings)
int fetch_counters(System* S);       // D2H of DevCounters + stream sync
int ensure_capacity(System* S);      // first-use allocation of every capacity-bound array
int grow_after_overflow(System* S);  // reads h_cnt->need_*, grows, clears the device overflow mask

// ---- stage enqueue functions: no host synchronisation inside ----
int enqueue_update_forces(System* S);                                    // integrate.cu
int enqueue_integrate(System* S);                                        // integrate.cu (also counts the step as done)
int enqueue_aabb(System* S);                                             // broadphase.cu
int enqueue_bins(System* S);                                             // broadphase.cu: grid + counting sort
int enqueue_broadphase(System* S, bool want_pairs, bool want_contacts);  // broadphase.cu
int enqueue_pairs_only(System* S, bool want_pairs, bool want_contacts, bool defer_geometry = false);  // pairs.cu
int ensure_contact_geometry(System* S);  // pairs.cu: contact records on demand
int alloc_shear(System* S);                                              // smc.cu: MultiStep history arrays
int enqueue_smc(System* S);                                              // smc.cu
int enqueue_nsc(System* S);                                              // nsc.cu
int nsc_contact_forces_inactive(System* S);                              // nsc.cu: on-demand part of CalculateContactForces

// ---- decomposed systems (comm.cu): systems that step in lockstep in this process.  One system per process over NCCL;
// several only when ranks are emulated on one GPU (tests), all on one stream ----
struct Group {
    std::vector<System*> sys;
};
void comm_release(System* S);         // frees the communicator (cb200_destroy)
bool comm_multi(const System* S);     // S is one slab of a decomposed system (has a communicator or a local group)
double* comm_gsum(System* S);         // 24 doubles of device scratch for sums over all ranks
int comm_halo6(Group& g, int which);  // owners' 6 doubles per body (0: n_w, 1: n_w0, 2: n_wy) -> the neighbours' ghost copies
int comm_allreduce(Group& g, int first, int n);  // gsum[first..first+n) summed over all ranks, in place
int enqueue_build_meta(System* S);  // broadphase.cu
int enqueue_report(System* S);  // report.cu
int enqueue_bilateral(System* S);  // bilateral.cu
int aux_fork(System* S, int i, cudaStream_t* out);  // `out`: the stream to launch the branch on
int aux_join(System* S, int i);
bool integrate_fuses_aabb(const System* S);
int enqueue_reset_bbox(System* S, cudaStream_t st);
int enqueue_nsc_group(Group& g);      // nsc.cu: the NSC solve of all systems of the group, slot by slot

// exclusive scans (scan.cu) over a device-side element count (*n_dev <= n_cap). n+1 outputs: out[n] = total.
// In-place allowed.
int scan_u32(System* S, const uint32_t* in, uint32_t* out, const uint32_t* n_dev, size_t n_cap);
int scan_u64(System* S, const uint64_t* in, uint64_t* out, const uint32_t* n_dev, size_t n_cap);
// bin histogram (u32) -> packed u64 (first entry | index among active bins << 32), CMC/ChBroadphase.cpp:290-301
int scan_bins(System* S, const uint32_t* hist, uint64_t* out, const uint32_t* n_dev, size_t n_cap);

inline int blocks_for(size_t n, int threads) { return (int)((n + threads - 1) / threads); }
// grid for a grid-stride kernel over at most n_cap items: enough CTAs to fill the machine, never more than needed
inline int grid_for(const System* S, size_t n_cap, int threads, int per_sm = 8) {
    size_t need = (n_cap + threads - 1) / threads;
    size_t full = (size_t)S->num_sms * per_sm;
    if (need < 1)
        need = 1;
    return (int)(need < full ? need : full);
}

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


#ifdef __CUDACC__
// The boxes of a CTA's threads (counted == false: none) join the global bounding box (BoxInvert / BoxReduce,
// CMC/ChBroadphase.cpp:38-100: min and max are exact and order-free, so any reduction tree gives the reference's bits).
// Keys are the order-preserving 64-bit encodings; a warp reduces them with two 32-bit hardware reductions per value
// (high word, then the low words of the lanes that hold that high word), the CTA combines its warps in shared memory and
// issues at most six atomics, and only those that still improve on a fresh read of the global value.
// Must be called by all threads of the CTA.
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
    const uint32_t hi = (uint32_t)(v >> 32), lo = (uint32_t)v;
    const uint32_t mh = __reduce_min_sync(0xffffffffu, hi);
    const uint32_t ml = __reduce_min_sync(0xffffffffu, hi == mh ? lo : 0xffffffffu);
    return ((unsigned long long)mh << 32) | ml;
}
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
    const uint32_t hi = (uint32_t)(v >> 32), lo = (uint32_t)v;
    const uint32_t mh = __reduce_max_sync(0xffffffffu, hi);
    const uint32_t ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
    return ((unsigned long long)mh << 32) | ml;
}
__device__ __forceinline__ void bbox_block_merge(DevCounters* cnt, const d3& mn, const d3& mx, bool counted) {
    __shared__ unsigned long long part[32][6];
    unsigned long long k[6];
    k[0] = counted ? enc_ordered(mn.x) : ~0ull, k[1] = counted ? enc_ordered(mn.y) : ~0ull;
    k[2] = counted ? enc_ordered(mn.z) : ~0ull, k[3] = counted ? enc_ordered(mx.x) : 0ull;
    k[4] = counted ? enc_ordered(mx.y) : 0ull, k[5] = counted ? enc_ordered(mx.z) : 0ull;
#pragma unroll
    for (int a = 0; a < 3; a++) {
        k[a] = warp_min_u64(k[a]);
        k[3 + a] = warp_max_u64(k[3 + a]);
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    if (lane == 0)
#pragma unroll
        for (int a = 0; a < 6; a++)
            part[wid][a] = k[a];
    __syncthreads();
    if (threadIdx.x < 6) {
        const int a = threadIdx.x;
        unsigned long long v = part[0][a];
        for (int w = 1; w < nw; w++)
            v = a < 3 ? min(v, part[w][a]) : max(v, part[w][a]);
        const unsigned long long cur = reinterpret_cast<volatile unsigned long long*>(cnt->bbox)[a];
        if (a < 3 ? v < cur : v > cur) {
            if (a < 3)
                atomicMin(&cnt->bbox[a], v);
            else
                atomicMax(&cnt->bbox[a], v);
        }
    }
}
#endif

}  // namespace cb200
