// api.cu -- the C ABI declared in include/chrono_b200.h: handle management, uploads/downloads, step driver.
#include "common.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits.h>
#include <limits>

namespace cb200 {

thread_local std::string g_last_error;
int set_error(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

// Timers never block the stream: every start/stop records into a pool of event pairs, and the pairs are
// resolved in one go (flush_timers) at the end of cb200_do_steps or when the pool holds 64 intervals.
int timer_start(System* S, const char* name) {
    if (S->capturing)
        return CB200_OK;
    Timer& t = S->timers[name];
    if (t.used == t.pool.size()) {
        cudaEvent_t a, b;
        CB_CUDA(cudaEventCreate(&a));
        CB_CUDA(cudaEventCreate(&b));
        t.pool.emplace_back(a, b);
    }
    CB_CUDA(cudaEventRecord(t.pool[t.used].first, S->stream));
    t.armed = true;
    return CB200_OK;
}

int timer_stop(System* S, const char* name, int steps) {
    if (S->capturing)
        return CB200_OK;
    Timer& t = S->timers[name];
    if (!t.armed)
        return CB200_OK;
    CB_CUDA(cudaEventRecord(t.pool[t.used].second, S->stream));
    t.used++;
    t.steps += steps;
    t.armed = false;
    return CB200_OK;
}

int flush_timers(System* S) {
    CB_CUDA(cudaStreamSynchronize(S->stream));
    for (auto& kv : S->timers) {
        Timer& t = kv.second;
        for (size_t i = 0; i < t.used; i++) {
            float ms = 0;
            CB_CUDA(cudaEventElapsedTime(&ms, t.pool[i].first, t.pool[i].second));
            t.total_ms += ms;
        }
        t.used = 0;
    }
    return CB200_OK;
}

void invalidate_step_graph(System* S) {
    for (int p = 0; p < 2; p++)
        if (S->step_graph[p]) {
            cudaGraphExecDestroy(S->step_graph[p]);
            S->step_graph[p] = nullptr;
        }
    S->direct_steps = 0;
}

// `unsupported` is sticky on the device (a step in the middle of a chunk must not hide it): the host clears it at the
// start of a synchronising call, i.e. right after the previous call has read and reported it.
static int clear_unsupported(System* S) {
    CB_CUDA(cudaMemsetAsync(&S->d_cnt->unsupported, 0, 4, S->stream));
    return CB200_OK;
}

int fetch_counters(System* S) {
    CB_CUDA(cudaMemcpyAsync(S->h_cnt, S->d_cnt, sizeof(DevCounters), cudaMemcpyDeviceToHost, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}

template <typename T>
static int upload(System* S, DevBuf<T>& buf, size_t offset, const T* src, size_t n) {
    CB_CUDA(cudaMemcpyAsync(buf.p + offset, src, n * sizeof(T), cudaMemcpyHostToDevice, S->stream));
    return CB200_OK;
}
template <typename T>
static int download(System* S, const T* dev, T* dst, size_t n) {
    if (n == 0 || !dst)
        return CB200_OK;
    CB_CUDA(cudaMemcpyAsync(dst, dev, n * sizeof(T), cudaMemcpyDeviceToHost, S->stream));
    return CB200_OK;
}

// ---- capacities -------------------------------------------------------------------------------------------
static int alloc_bins(System* S, uint32_t n) {
    invalidate_step_graph(S);
    S->bins_clean = false;
    CB_TRY(S->bin_start.reserve((size_t)n + 2));
    CB_TRY(S->bin_cursor.reserve((size_t)n + 1));
    S->cap_bins = (uint32_t)std::min<size_t>(S->bin_cursor.cap - 1, S->bin_start.cap - 2);
    return CB200_OK;
}
static int alloc_pairq(System* S, uint32_t n) {
    invalidate_step_graph(S);
    CB_TRY(S->pair_q.reserve(n));
    S->cap_pairq = (uint32_t)S->pair_q.cap;
    return CB200_OK;
}
static int alloc_entries(System* S, uint32_t n) {
    invalidate_step_graph(S);
    CB_TRY(S->ent_tmp.reserve(n));
    CB_TRY(S->ent_bin.reserve(n));
    CB_TRY(S->ent_shape.reserve(n));
    CB_TRY(S->ent_mask.reserve(n));
    CB_TRY(S->ent_mask_hi.reserve(n));
    S->cap_entries = (uint32_t)S->ent_tmp.cap;
    CB_TRY(S->ent_off.reserve((size_t)S->cap_entries + 4));
    CB_TRY(S->slow_bins.reserve((size_t)S->cap_entries + 4));
    CB_TRY(S->long_bins.reserve((size_t)S->cap_entries / 64 + 4));
    if (S->cap_pairq == 0)
        CB_TRY(alloc_pairq(S, S->cap_entries / 4 + 1024));
    if (use_complex(S)) {
        CB_TRY(S->ent_cmask.reserve(S->cap_entries));
        CB_TRY(S->ent_cbase.reserve(S->cap_entries));
    }
    return CB200_OK;
}
static int alloc_contacts(System* S, uint32_t n) {
    invalidate_step_graph(S);
    CB_TRY(S->c_depth.reserve(n));
    const size_t c = S->c_depth.cap;
    CB_TRY(S->c_norm.reserve(3 * c));
    CB_TRY(S->c_pta.reserve(3 * c));
    CB_TRY(S->c_ptb.reserve(3 * c));
    CB_TRY(S->c_erad.reserve(c));
    CB_TRY(S->c_bids.reserve(2 * c));
    CB_TRY(S->c_shapes.reserve(c));
    CB_TRY(S->c_mats.reserve(c));
    CB_TRY(S->ct_force.reserve(CT_STRIDE * (size_t)c));
    CB_TRY(S->b_inc.reserve(2 * c));
    CB_TRY(S->heavy_part.reserve((size_t)MAX_HEAVY * (c / 4096 + 2) * 7));
    S->cap_contacts = (uint32_t)c;
    if (use_complex(S))
        CB_TRY(S->c_src.reserve(c));
    return CB200_OK;
}
static int alloc_complex(System* S, uint32_t nrec, uint32_t nslots) {
    invalidate_step_graph(S);
    if (nslots < 2 * nrec)
        nslots = 2 * nrec;
    CB_TRY(S->cplx_rec.reserve(4 * (size_t)nrec));
    S->cap_complex = (uint32_t)(S->cplx_rec.cap / 4);
    CB_TRY(S->cplx_n.reserve(S->cap_complex));
    CB_TRY(S->cplx_cls.reserve(4 * (size_t)S->cap_complex));
    CB_TRY(S->cplx_bb.reserve(5 * (size_t)S->cap_complex));
    CB_TRY(S->cplx_out.reserve(11 * (size_t)nslots));
    S->cap_cslots = (uint32_t)(S->cplx_out.cap / 11);
    return CB200_OK;
}
static int alloc_pairs(System* S, uint32_t n) {
    invalidate_step_graph(S);
    CB_TRY(S->pairs.reserve(n));
    S->cap_pairs = (uint32_t)S->pairs.cap;
    return CB200_OK;
}

// First-use allocation of everything a step touches.  Sizes are guesses; a guess that turns out too small costs one
// re-run of the step that found out (grow_after_overflow).
int ensure_capacity(System* S) {
    const size_t nb = S->nb, ns = S->ns;
    if (S->b_deg.cap < nb + 2 || S->b_cursor.cap < nb + 1)
        S->body_scratch_clean = false;
    CB_TRY(S->b_deg.reserve(nb + 2));
    CB_TRY(S->b_cursor.reserve(nb + 1));
    CB_TRY(S->b_force.reserve(6 * nb + 6));
    CB_TRY(S->heavy_list.reserve(MAX_HEAVY));
    CB_TRY(S->large_list.reserve(MAX_LARGE));
    CB_TRY(S->bin_intersections.reserve(ns + 1));
    if (S->cap_bins == 0)
        CB_TRY(alloc_bins(S, (uint32_t)std::max<size_t>(4 * ns, 1 << 14)));
    if (S->cap_entries < 4 * ns + 1024)
        CB_TRY(alloc_entries(S, (uint32_t)(8 * ns + 4096)));
    if (S->cap_contacts < 2 * ns + 1024)
        CB_TRY(alloc_contacts(S, (uint32_t)(6 * ns + 4096)));
    if (use_complex(S)) {
        CB_TRY(S->ent_cmask.reserve(S->cap_entries));
        CB_TRY(S->ent_cbase.reserve(S->cap_entries));
        CB_TRY(S->c_src.reserve(S->cap_contacts));
        if (S->cap_complex == 0)
            CB_TRY(alloc_complex(S, (uint32_t)(4 * (size_t)S->n_nonsphere + 4096), 0));
    }
    if (S->set.system_type == 1 && S->set.tangential_displ_mode == 2)
        CB_TRY(alloc_shear(S));
    return CB200_OK;
}

// Called after fetch_counters() found the overflow mask set: grow what was too small and clear the mask.
int grow_after_overflow(System* S) {
    const DevCounters& c = *S->h_cnt;
    // the steps enqueued behind the overflow were no-ops on the device (their integration prepared nothing) while the
    // host-side bookkeeping went on: the next step computes the spheres' records itself
    S->sphere_aabb_valid = false;
    S->bins_clean = S->body_scratch_clean = false;  // the step stopped half way
    invalidate_step_graph(S);  // the clears are part of the next (directly enqueued) step, not of the graphs
    if (c.overflow & OVF_LARGE)
        return set_error(CB200_ERR_UNSUPPORTED, "more than 4096 shapes span more than 48 bins each");
    if (c.overflow & OVF_HEAVY)
        return set_error(CB200_ERR_UNSUPPORTED, "more than 64 bodies own a shape spanning more than 48 bins");
    if (c.overflow & OVF_SLAB)
        return set_error(CB200_ERR_STATE, "slab ghost / arrival capacity exceeded: bodies that crossed into this slab "
                                          "were dropped (recreate the slab system with larger owned_cap / ghost_cap)");
    if (c.overflow & OVF_BINS) {
        if (c.need_bins > (1u << 28))
            return set_error(CB200_ERR_UNSUPPORTED, "broadphase grid has more than 2^28 bins");
        CB_TRY(alloc_bins(S, c.need_bins + c.need_bins / 8));
    }
    if (c.overflow & OVF_ENTRIES)
        CB_TRY(alloc_entries(S, c.need_entries + c.need_entries / 4));
    if (c.overflow & OVF_CONTACTS)
        CB_TRY(alloc_contacts(S, c.need_contacts + c.need_contacts / 4));
    if (c.overflow & OVF_PAIRS)
        CB_TRY(alloc_pairs(S, c.need_pairs + c.need_pairs / 8));
    if (c.overflow & OVF_COMPLEX)
        CB_TRY(alloc_complex(S, c.need_complex + c.need_complex / 4, c.need_cslots + c.need_cslots / 4));
    if (c.overflow & OVF_PAIRQ)
        CB_TRY(alloc_pairq(S, c.need_pairq + c.need_pairq / 4));
    if (c.overflow & OVF_NSC)
        S->nsc_extra_slots = 2 * S->nsc_extra_slots + 32 + (int)c.need_nsc_slots / 4;
    CB_CUDA(cudaMemsetAsync(&S->d_cnt->overflow, 0, 4, S->stream));
    S->h_cnt->overflow = 0;
    return CB200_OK;
}

// Enqueue `what`, wait, and repeat while a capacity was exceeded (stage-level calls; one sync per call).
template <typename F>
static int run_stage(System* S, F what) {
    CB_TRY(ensure_capacity(S));
    for (int attempt = 0; attempt < 8; attempt++) {
        CB_TRY(what());
        CB_TRY(fetch_counters(S));
        if (!S->h_cnt->overflow)
            return CB200_OK;
        CB_TRY(grow_after_overflow(S));
    }
    return set_error(CB200_ERR_STATE, "capacity growth did not converge");
}

}  // namespace cb200

using namespace cb200;

static System* SYS(cb200_system* s) { return reinterpret_cast<System*>(s); }

extern "C" {

const char* cb200_last_error(void) { return g_last_error.c_str(); }

int cb200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
        return 0;
    return n;
}

int cb200_host_alloc(uint64_t bytes, void** out) {
    if (!out)
        return set_error(CB200_ERR_ARG, "null output pointer");
    CB_CUDA(cudaMallocHost(out, bytes ? bytes : 1));
    return CB200_OK;
}

int cb200_host_free(void* p) {
    if (p)
        CB_CUDA(cudaFreeHost(p));
    return CB200_OK;
}

void cb200_default_settings(cb200_settings* s) {
    memset(s, 0, sizeof(*s));
    s->collision_envelope = 0;
    s->broadphase_grid = 0;
    s->bins_per_axis[0] = s->bins_per_axis[1] = s->bins_per_axis[2] = 10;
    s->bin_size[0] = s->bin_size[1] = s->bin_size[2] = 1;
    s->grid_density = 5;
    s->narrowphase_algorithm = 2;
    s->system_type = 0;
    s->step_size = 0.01;
    s->contact_force_model = 1;
    s->adhesion_force_model = 0;
    s->tangential_displ_mode = 1;
    s->use_material_properties = 1;
    s->characteristic_vel = 1;
    s->min_slip_vel = 1e-4;
    s->min_roll_vel = 1e-4;
    s->min_spin_vel = 1e-4;
    s->solver_type = 0;
    s->solver_mode = 1;
    s->max_iteration_normal = 0;
    s->max_iteration_sliding = 100;
    s->max_iteration_spinning = 0;
    s->cache_step_length = 0;
    s->alpha = .2;
    s->contact_recovery_speed = .6;
    s->tolerance = 1e-4;
    s->deterministic = 1;
    s->test_objective = 0;
    s->tolerance_objective = 1e-6;
}

int cb200_create(int device, cb200_system** out) {
    if (!out)
        return set_error(CB200_ERR_ARG, "null output handle");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return set_error(CB200_ERR_CUDA, std::string("no CUDA device available (this backend has no CPU fallback): ") +
                                             cudaGetErrorString(e));
    if (device < 0 || device >= n)
        return set_error(CB200_ERR_ARG, "device index out of range");
    CB_CUDA(cudaSetDevice(device));
    System* S = new System();
    S->device = device;
    CB_CUDA(cudaDeviceGetAttribute(&S->num_sms, cudaDevAttrMultiProcessorCount, device));
    cb200_default_settings(&S->set);
    CB_CUDA(cudaStreamCreateWithFlags(&S->stream, cudaStreamNonBlocking));
    S->own_stream = S->stream;
    CB_CUDA(cudaMalloc(&S->d_cnt, sizeof(DevCounters)));
    CB_CUDA(cudaMemset(S->d_cnt, 0, sizeof(DevCounters)));
    CB_CUDA(cudaMallocHost(&S->h_cnt, sizeof(DevCounters)));
    memset(S->h_cnt, 0, sizeof(DevCounters));
    *out = reinterpret_cast<cb200_system*>(S);
    return CB200_OK;
}

void cb200_destroy(cb200_system* sys) {
    System* S = SYS(sys);
    if (!S)
        return;
    cudaSetDevice(S->device);
    cudaStreamSynchronize(S->stream);
    comm_release(S);
#define REL(x) S->x.release()
    REL(pos); REL(rot); REL(vel); REL(hf); REL(fapp); REL(mass); REL(inertia); REL(inv_inertia); REL(inertia_off); REL(inv_inertia_off); REL(bflags); REL(heavy); REL(b_sphere); REL(b_rad); REL(c_own); REL(vlim); REL(b_inbox); REL(report); REL(mat_ovr); REL(bil_ids); REL(bil_cq); REL(bil_b); REL(bil_R); REL(bil_gamma); REL(bil_N); REL(bil_vec); REL(bil_info);
    REL(s_type); REL(s_body); REL(s_fam); REL(s_mat); REL(s_lpos); REL(s_lrot); REL(s_dims); REL(general_ids); REL(s_verts); REL(sh_neigh); REL(sh_touch); REL(sh_disp); REL(sh_relvel); REL(sh_dur); REL(s_wpos); REL(s_wpos_alt); REL(s_wrot);
    REL(s_box); REL(s_meta); REL(mat_smc); REL(mat_nsc); REL(bin_intersections); REL(bin_start); REL(bin_cursor);
    REL(ent_tmp); REL(ent_bin); REL(ent_shape); REL(ent_mask); REL(ent_mask_hi); REL(ent_off); REL(slow_bins); REL(long_bins); REL(pair_q); REL(large_list); REL(pairs);
    REL(c_norm); REL(c_pta); REL(c_ptb); REL(c_depth); REL(c_erad); REL(c_bids); REL(c_shapes); REL(c_mats);
    REL(ent_cmask); REL(ent_cbase); REL(cplx_rec); REL(cplx_n); REL(cplx_cls); REL(cplx_bb); REL(cplx_out); REL(c_src);
    REL(ct_force); REL(ct_torque); REL(b_deg); REL(b_cursor); REL(b_inc); REL(b_force); REL(heavy_list);
    REL(heavy_part); REL(n_frame); REL(n_aux); REL(gamma); REL(n_gamma_hat); REL(n_y); REL(n_gn); REL(n_g); REL(n_rfull);
    REL(n_w); REL(n_w0); REL(n_wy); REL(n_wg); REL(n_mk); REL(n_part); REL(n_hist); REL(n_rows);
    invalidate_step_graph(S);
    if (S->copy_stream) {
        cudaStreamDestroy(S->copy_stream);
        cudaEventDestroy(S->ev_copy);
        cudaEventDestroy(S->ev_main);
    }
    if (S->n_ctl)
        cudaFree(S->n_ctl);
    REL(scan_tmp);
#undef REL
    for (auto& kv : S->timers)
        for (auto& ev : kv.second.pool) {
            cudaEventDestroy(ev.first);
            cudaEventDestroy(ev.second);
        }
    cudaFree(S->d_cnt);
    cudaFreeHost(S->h_cnt);
    S->b_gid.release(); S->slab_flags.release(); S->slab_vacant.release(); S->slab_cta.release();
    if (S->slab_cnt)
        cudaFree(S->slab_cnt);
    for (int i = 0; i < 2; i++)
        if (S->aux[i])
            cudaStreamDestroy(S->aux[i]);
    for (int i = 0; i < S->aux_ev_made; i++)
        cudaEventDestroy(S->aux_ev[i]);
    cudaStreamDestroy(S->own_stream);
    delete S;
}

}  // extern "C"

namespace cb200 {
static int aux_event(System* S, cudaEvent_t* ev) {
    if (S->aux_ev_next == 32)  // a step uses about a dozen: an event is re-recorded long after its waits were enqueued
        S->aux_ev_next = 0;
    if (S->aux_ev_next == S->aux_ev_made) {
        CB_CUDA(cudaEventCreateWithFlags(&S->aux_ev[S->aux_ev_made], cudaEventDisableTiming));
        S->aux_ev_made++;
    }
    *ev = S->aux_ev[S->aux_ev_next++];
    return CB200_OK;
}
// Branch i of the step starts here: everything enqueued on the main stream so far happens before it.
int aux_fork(System* S, int i, cudaStream_t* out) {
    *out = S->stream;
    if (!S->aux_enabled)
        return CB200_OK;
    if (!S->aux[i])
        CB_CUDA(cudaStreamCreateWithFlags(&S->aux[i], cudaStreamNonBlocking));
    cudaEvent_t ev;
    CB_TRY(aux_event(S, &ev));
    CB_CUDA(cudaEventRecord(ev, S->stream));
    CB_CUDA(cudaStreamWaitEvent(S->aux[i], ev, 0));
    *out = S->aux[i];
    return CB200_OK;
}
// ... and ends here: what follows on the main stream happens after it.
int aux_join(System* S, int i) {
    if (!S->aux_enabled || !S->aux[i])
        return CB200_OK;
    cudaEvent_t ev;
    CB_TRY(aux_event(S, &ev));
    CB_CUDA(cudaEventRecord(ev, S->aux[i]));
    CB_CUDA(cudaStreamWaitEvent(S->stream, ev, 0));
    return CB200_OK;
}
}  // namespace cb200

namespace cb200 {
// gather of fixed-size records: dst[j] = src[idx[j]], W 4-byte words per record
template <int W>
__global__ void __launch_bounds__(256) k_gather_words(uint32_t m, const uint32_t* __restrict__ idx,
                                                      const uint32_t* __restrict__ src, uint32_t* __restrict__ dst) {
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < m * W; t += gridDim.x * blockDim.x)
        dst[t] = src[(size_t)idx[t / W] * W + t % W];
}

template <int W, typename T>
static int compact_array(System* S, DevBuf<T>& buf, uint32_t m, const uint32_t* idx_dev, uint32_t* scratch) {
    static_assert(sizeof(T) % 4 == 0 || W * 4 % sizeof(T) == 0, "record made of 4-byte words");
    if (!buf.p || m == 0)
        return CB200_OK;
    k_gather_words<W><<<grid_for(S, (size_t)m * W, 256), 256, 0, S->stream>>>(
        m, idx_dev, reinterpret_cast<const uint32_t*>(buf.p), scratch);
    CB_CUDA(cudaMemcpyAsync(buf.p, scratch, (size_t)m * W * 4, cudaMemcpyDeviceToDevice, S->stream));
    S->launches++;
    return CB200_OK;
}

}  // namespace cb200

extern "C" {

int cb200_set_settings(cb200_system* sys, const cb200_settings* s) {
    System* S = SYS(sys);
    if (!S || !s)
        return set_error(CB200_ERR_ARG, "null argument");
    {
        cb200_settings eff = *s;
        if (eff.system_type == 1)
            eff.collision_envelope = 0;
        if (memcmp(&eff, &S->set, sizeof(eff)) != 0)
            invalidate_step_graph(S);  // captured launches carry the settings by value
    }
    if (!(s->step_size > 0))
        return set_error(CB200_ERR_ARG, "step_size must be positive");
    if (s->broadphase_grid < 0 || s->broadphase_grid > 2)
        return set_error(CB200_ERR_ARG, "broadphase_grid must be 0, 1 or 2");
    if (s->narrowphase_algorithm < 0 || s->narrowphase_algorithm > 2)
        return set_error(CB200_ERR_ARG, "narrowphase_algorithm must be 0, 1 or 2");
    S->set = *s;
    S->sphere_aabb_valid = false;
    // ChSystemMulticoreSMC forces a zero envelope every step (MC/physics/ChSystemMulticoreSMC.cpp:27,56-67)
    if (S->set.system_type == 1)
        S->set.collision_envelope = 0;
    return CB200_OK;
}

int cb200_append_bodies(cb200_system* sys, int n, const double* mass, const double* inertia_xx,
                        const double* inv_inertia_xx, const char* active, const char* collide, int* first) {
    System* S = SYS(sys);
    if (S)
        invalidate_step_graph(S);
    if (!S || n < 0 || !mass || !inertia_xx)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_append_bodies");
    CB_CUDA(cudaSetDevice(S->device));
    const size_t nb0 = S->nb, nb1 = nb0 + (size_t)n;
    cudaStream_t st = S->stream;
    CB_TRY(S->pos.reserve(3 * nb1, true, st));
    CB_TRY(S->rot.reserve(4 * nb1, true, st));
    CB_TRY(S->vel.reserve(6 * nb1, true, st));
    CB_TRY(S->hf.reserve(6 * nb1, true, st));
    CB_TRY(S->fapp.reserve(6 * nb1, true, st));
    CB_TRY(S->mass.reserve(nb1, true, st));
    CB_TRY(S->inertia.reserve(3 * nb1, true, st));
    CB_TRY(S->inv_inertia.reserve(3 * nb1, true, st));
    CB_TRY(S->bflags.reserve(nb1, true, st));
    CB_TRY(S->heavy.reserve(nb1, true, st));
    CB_TRY(S->vlim.reserve(2 * nb1, true, st));
    std::vector<double> inv(3 * (size_t)n), rot0(4 * (size_t)n, 0.0);
    std::vector<uint8_t> fl(n);
    for (int i = 0; i < n; i++) {
        for (int k = 0; k < 3; k++)
            inv[3 * i + k] = inv_inertia_xx ? inv_inertia_xx[3 * i + k] : 1.0 / inertia_xx[3 * i + k];
        rot0[4 * i] = 1.0;
        fl[i] = (uint8_t)(((!active || active[i]) ? 1 : 0) | ((!collide || collide[i]) ? 2 : 0));
    }
    CB_TRY(upload(S, S->mass, nb0, mass, (size_t)n));
    CB_TRY(upload(S, S->inertia, 3 * nb0, inertia_xx, 3 * (size_t)n));
    CB_TRY(upload(S, S->inv_inertia, 3 * nb0, inv.data(), 3 * (size_t)n));
    CB_TRY(upload(S, S->rot, 4 * nb0, rot0.data(), 4 * (size_t)n));
    CB_TRY(upload(S, S->bflags, nb0, fl.data(), (size_t)n));
    CB_CUDA(cudaMemsetAsync(S->pos.p + 3 * nb0, 0, 3 * (size_t)n * 8, st));
    CB_CUDA(cudaMemsetAsync(S->vel.p + 6 * nb0, 0, 6 * (size_t)n * 8, st));
    CB_CUDA(cudaMemsetAsync(S->hf.p + 6 * nb0, 0, 6 * (size_t)n * 8, st));
    CB_CUDA(cudaMemsetAsync(S->fapp.p + 6 * nb0, 0, 6 * (size_t)n * 8, st));
    CB_CUDA(cudaMemsetAsync(S->heavy.p + nb0, 0, (size_t)n, st));
    CB_CUDA(cudaMemsetAsync(S->vlim.p + 2 * nb0, 0, 2 * (size_t)n * 8, st));  // no speed limit
    if (S->inertia_off.p) {  // some earlier body has products of inertia: the new ones have none
        CB_TRY(S->inertia_off.reserve(3 * nb1, true, st));
        CB_TRY(S->inv_inertia_off.reserve(3 * nb1, true, st));
        CB_CUDA(cudaMemsetAsync(S->inertia_off.p + 3 * nb0, 0, 3 * (size_t)n * 8, st));
        CB_CUDA(cudaMemsetAsync(S->inv_inertia_off.p + 3 * nb0, 0, 3 * (size_t)n * 8, st));
    }
    S->h_b_sphere.resize(nb1, -1);
    S->h_b_rad.resize(nb1, 0.0);
    CB_TRY(S->b_sphere.reserve(nb1, true, st));
    CB_TRY(S->b_rad.reserve(nb1, true, st));
    CB_TRY(upload(S, S->b_sphere, nb0, S->h_b_sphere.data() + nb0, (size_t)n));
    CB_TRY(upload(S, S->b_rad, nb0, S->h_b_rad.data() + nb0, (size_t)n));
    CB_CUDA(cudaStreamSynchronize(st));  // host staging buffers go out of scope
    if (first)
        *first = (int)nb0;
    S->nb = (uint32_t)nb1;
    S->sphere_aabb_valid = false;
    return CB200_OK;
}

int cb200_append_shapes(cb200_system* sys, int n, const int32_t* type, const uint32_t* body_id, const int16_t* fam,
                        const double* pos_local, const double* rot_local, const double* dims,
                        const int32_t* material_id) {
    System* S = SYS(sys);
    if (S)
        invalidate_step_graph(S);
    if (!S || n < 0 || !type || !body_id || !fam || !pos_local || !rot_local || !dims)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_append_shapes");
    CB_CUDA(cudaSetDevice(S->device));
    for (int i = 0; i < n; i++) {
        switch (type[i]) {
            case CB200_SPHERE: case CB200_ELLIPSOID: case CB200_BOX: case CB200_CYLINDER: case CB200_CYLSHELL:
            case CB200_CAPSULE: case CB200_CONE: case CB200_ROUNDEDBOX: case CB200_ROUNDEDCYL:
                break;
            case CB200_TRIANGLE: case CB200_CONVEXHULL: {
                const double first = dims[4 * i], count = type[i] == CB200_TRIANGLE ? 3.0 : dims[4 * i + 1];
                if (!(first >= 0) || !(count >= 1) || first + count > (double)S->n_verts)
                    return set_error(CB200_ERR_ARG, "triangle / convex hull refers to vertices that were not added "
                                                    "(cb200_append_shape_vertices)");
                break;
            }
            default:
                return set_error(CB200_ERR_UNSUPPORTED, "shape type " + std::to_string(type[i]) + " is not supported");
        }
        if (body_id[i] != UINT_MAX && body_id[i] >= S->nb)
            return set_error(CB200_ERR_ARG, "shape refers to a body that was not added");
        if (material_id && (material_id[i] < 0 || material_id[i] > 0xffff))
            return set_error(CB200_ERR_ARG, "material id out of range");
    }
    const size_t ns0 = S->ns, ns1 = ns0 + (size_t)n;
    cudaStream_t st = S->stream;
    CB_TRY(S->s_type.reserve(ns1, true, st));
    CB_TRY(S->s_body.reserve(ns1, true, st));
    CB_TRY(S->s_fam.reserve(ns1, true, st));
    CB_TRY(S->s_mat.reserve(ns1, true, st));
    CB_TRY(S->s_lpos.reserve(3 * ns1, true, st));
    CB_TRY(S->s_lrot.reserve(4 * ns1, true, st));
    CB_TRY(S->s_dims.reserve(4 * ns1, true, st));
    CB_TRY(S->s_wpos.reserve(4 * ns1, true, st));
    CB_TRY(S->s_wpos_alt.reserve(4 * ns1, true, st));
    CB_TRY(S->s_wrot.reserve(4 * ns1, true, st));
    CB_TRY(S->s_box.reserve(ns1, true, st));
    CB_TRY(S->s_meta.reserve(ns1, true, st));
    std::vector<uint32_t> f(n);
    std::vector<int32_t> m(n, 0);
    for (int i = 0; i < n; i++) {
        f[i] = (uint32_t)(uint16_t)fam[2 * i] | ((uint32_t)(uint16_t)fam[2 * i + 1] << 16);
        if (material_id)
            m[i] = material_id[i];
    }
    CB_TRY(upload(S, S->s_type, ns0, type, (size_t)n));
    CB_TRY(upload(S, S->s_body, ns0, body_id, (size_t)n));
    CB_TRY(upload(S, S->s_fam, ns0, f.data(), (size_t)n));
    CB_TRY(upload(S, S->s_mat, ns0, m.data(), (size_t)n));
    CB_TRY(upload(S, S->s_lpos, 3 * ns0, pos_local, 3 * (size_t)n));
    CB_TRY(upload(S, S->s_lrot, 4 * ns0, rot_local, 4 * (size_t)n));
    CB_TRY(upload(S, S->s_dims, 4 * ns0, dims, 4 * (size_t)n));
    // A body's first centred sphere is "its" sphere (b_sphere): k_aabb<false> / the integration kernel write its world
    // record.  Everything else is visited by k_aabb<true> through this list (same rule as k_build_meta's bit 2).
    std::vector<uint32_t> gen;
    bool own_changed = false;
    for (int i = 0; i < n; i++) {
        const bool centred = type[i] == CB200_SPHERE && pos_local[3 * i] == 0 && pos_local[3 * i + 1] == 0 &&
                             pos_local[3 * i + 2] == 0 && body_id[i] != UINT_MAX;
        if (centred && S->h_b_sphere[body_id[i]] < 0) {
            S->h_b_sphere[body_id[i]] = (int32_t)(ns0 + i);
            S->h_b_rad[body_id[i]] = dims[4 * i];
            own_changed = true;
        } else {
            gen.push_back((uint32_t)(ns0 + i));
        }
    }
    if (own_changed) {
        CB_TRY(upload(S, S->b_sphere, 0, S->h_b_sphere.data(), (size_t)S->nb));
        CB_TRY(upload(S, S->b_rad, 0, S->h_b_rad.data(), (size_t)S->nb));
    }
    S->sphere_aabb_valid = false;
    if (!gen.empty()) {
        CB_TRY(S->general_ids.reserve(S->n_general + gen.size(), true, st));
        CB_TRY(upload(S, S->general_ids, S->n_general, gen.data(), gen.size()));
        S->n_general += (uint32_t)gen.size();
    }
    CB_CUDA(cudaStreamSynchronize(st));
    S->ns = (uint32_t)ns1;
    S->meta_dirty = true;
    for (int i = 0; i < n; i++) {
        if (type[i] == CB200_SPHERE)
            continue;
        S->n_nonsphere++;
        // non-sphere shapes on two different bodies: pairs without a sphere can occur (shapes of one body never pair)
        if (S->nonsphere_bodies == 0) {
            S->nonsphere_bodies = 1;
            S->nonsphere_body0 = body_id[i];
        } else if (body_id[i] != S->nonsphere_body0) {
            S->nonsphere_bodies = 2;
        }
        // no analytic test against a sphere either: such pairs go to MPR, i.e. the complex path
        if (type[i] == CB200_ELLIPSOID || type[i] == CB200_CONE || type[i] == CB200_CYLSHELL ||
            type[i] == CB200_CONVEXHULL)
            S->nonsphere_bodies = 2;
    }
    return CB200_OK;
}

int cb200_append_shape_vertices(cb200_system* sys, int n, const double* xyz, int* first) {
    System* S = SYS(sys);
    if (S)
        invalidate_step_graph(S);
    if (!S || n < 0 || (n > 0 && !xyz))
        return set_error(CB200_ERR_ARG, "bad argument to cb200_append_shape_vertices");
    CB_CUDA(cudaSetDevice(S->device));
    const size_t n0 = S->n_verts, n1 = n0 + (size_t)n;
    CB_TRY(S->s_verts.reserve(3 * n1, true, S->stream));
    CB_TRY(upload(S, S->s_verts, 3 * n0, xyz, 3 * (size_t)n));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    S->n_verts = (uint32_t)n1;
    if (first)
        *first = (int)n0;
    return CB200_OK;
}

int cb200_set_inertia_full(cb200_system* sys, int first, int n, const double* inertia9, const double* inv_inertia9) {
    System* S = SYS(sys);
    if (S)
        invalidate_step_graph(S);  // the kernels of a captured step were handed null off-diagonal pointers
    if (!S || first < 0 || n < 0 || (n > 0 && (!inertia9 || !inv_inertia9)) || (size_t)first + (size_t)n > S->nb)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_set_inertia_full");
    if (S->slab_on)
        return set_error(CB200_ERR_UNSUPPORTED, "products of inertia do not travel with the slab records yet");
    if (S->n_bil > 0)
        return set_error(CB200_ERR_UNSUPPORTED, "bilateral rows use the diagonal of the inverse inertia only");
    if (n == 0)
        return CB200_OK;
    CB_CUDA(cudaSetDevice(S->device));
    cudaStream_t st = S->stream;
    const size_t had = S->inertia_off.p ? S->inertia_off.cap : 0;
    CB_TRY(S->inertia_off.reserve(3 * (size_t)S->nb, true, st));
    CB_TRY(S->inv_inertia_off.reserve(3 * (size_t)S->nb, true, st));
    if (had < S->inertia_off.cap) {  // bodies never given products of inertia have none
        CB_CUDA(cudaMemsetAsync(S->inertia_off.p + had, 0, (S->inertia_off.cap - had) * sizeof(double), st));
        CB_CUDA(cudaMemsetAsync(S->inv_inertia_off.p + had, 0, (S->inv_inertia_off.cap - had) * sizeof(double), st));
    }
    std::vector<double> d(3 * (size_t)n), o(3 * (size_t)n), di(3 * (size_t)n), oi(3 * (size_t)n);
    for (int i = 0; i < n; i++) {
        const double* J = inertia9 + 9 * (size_t)i;
        const double* K = inv_inertia9 + 9 * (size_t)i;
        if (J[1] != J[3] || J[2] != J[6] || J[5] != J[7] || K[1] != K[3] || K[2] != K[6] || K[5] != K[7])
            return set_error(CB200_ERR_ARG, "cb200_set_inertia_full: the matrices must be symmetric");
        d[3 * i] = J[0], d[3 * i + 1] = J[4], d[3 * i + 2] = J[8];
        o[3 * i] = J[1], o[3 * i + 1] = J[2], o[3 * i + 2] = J[5];
        di[3 * i] = K[0], di[3 * i + 1] = K[4], di[3 * i + 2] = K[8];
        oi[3 * i] = K[1], oi[3 * i + 1] = K[2], oi[3 * i + 2] = K[5];
    }
    CB_TRY(upload(S, S->inertia, 3 * (size_t)first, d.data(), d.size()));
    CB_TRY(upload(S, S->inv_inertia, 3 * (size_t)first, di.data(), di.size()));
    CB_TRY(upload(S, S->inertia_off, 3 * (size_t)first, o.data(), o.size()));
    CB_TRY(upload(S, S->inv_inertia_off, 3 * (size_t)first, oi.data(), oi.size()));
    CB_CUDA(cudaStreamSynchronize(st));
    return CB200_OK;
}

int cb200_update_shape_vertices(cb200_system* sys, int first, int n, const double* xyz) {
    System* S = SYS(sys);
    if (!S || first < 0 || n < 0 || (n > 0 && !xyz))
        return set_error(CB200_ERR_ARG, "bad argument to cb200_update_shape_vertices");
    if ((size_t)first + (size_t)n > S->n_verts)
        return set_error(CB200_ERR_ARG, "cb200_update_shape_vertices: range beyond the vertex pool");
    if (n == 0)
        return CB200_OK;
    CB_CUDA(cudaSetDevice(S->device));
    // same addresses, new values: the captured step graph stays valid
    CB_TRY(upload(S, S->s_verts, 3 * (size_t)first, xyz, 3 * (size_t)n));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}

// Composite of two SMC materials in float, term for term as ChContactMaterialCompositeSMC's constructor
// (src/chrono/physics/ChContactMaterialSMC.cpp:107-130) with the default composition strategy
// (src/chrono/physics/ChContactMaterial.h:157-169).
int cb200_set_materials_smc(cb200_system* sys, int n, const float* t) {
    System* S = SYS(sys);
    if (S)
        invalidate_step_graph(S);
    if (!S || n <= 0 || !t)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_set_materials_smc");
    if (n > 1024)
        return set_error(CB200_ERR_UNSUPPORTED, "more than 1024 distinct materials");
    CB_CUDA(cudaSetDevice(S->device));
    std::vector<double> tab((size_t)n * n * SMC_MAT_STRIDE);
    for (int a = 0; a < n; a++)
        for (int b = 0; b < n; b++) {
            const float* m1 = t + 14 * a;
            const float* m2 = t + 14 * b;
            const float E1 = m1[0], nu1 = m1[1], E2 = m2[0], nu2 = m2[1];
            float inv_E = (1 - nu1 * nu1) / E1 + (1 - nu2 * nu2) / E2;
            float inv_G = 2 * (2 - nu1) * (1 + nu1) / E1 + 2 * (2 - nu2) * (1 + nu2) / E2;
            float E_eff = 1 / inv_E;
            float G_eff = 1 / inv_G;
            double* o = tab.data() + ((size_t)a * n + b) * SMC_MAT_STRIDE;
            o[0] = std::min<float>(m1[2], m2[2]);            // static friction
            o[1] = std::min<float>(m1[4], m2[4]);            // rolling
            o[2] = std::min<float>(m1[5], m2[5]);            // spinning
            o[3] = E_eff;
            o[4] = G_eff;
            o[5] = std::min<float>(m1[7], m2[7]);            // constant adhesion
            o[6] = std::min<float>(m1[8], m2[8]);            // DMT multiplier
            o[7] = std::min<float>(m1[9], m2[9]);            // Perko multiplier
            o[8] = std::min<float>(m1[6], m2[6]);            // restitution
            o[9] = (float)((m1[10] + m2[10]) / 2);           // kn
            o[10] = (float)((m1[11] + m2[11]) / 2);          // kt
            o[11] = (float)((m1[12] + m2[12]) / 2);          // gn
            o[12] = (float)((m1[13] + m2[13]) / 2);          // gt
            // Hertz damping ratio beta (MC/solver/ChIterativeSolverMulticoreSMC.cpp:289-290), constant per pair
            {
                const double cr = o[8], eps = std::numeric_limits<double>::epsilon();
                const double loge = (cr < eps) ? std::log(eps) : std::log(cr);
                o[13] = loge / std::sqrt(loge * loge + M_PI * M_PI);
            }
        }
    CB_TRY(S->mat_smc.reserve(tab.size()));
    CB_TRY(upload(S, S->mat_smc, 0, tab.data(), tab.size()));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    S->nm_smc = n;
    return CB200_OK;
}

// NSC composite (src/chrono/physics/ChContactMaterialNSC.cpp composite ctor): friction/cohesion/restitution by
// min, compliances by sum; 8 doubles: mu, mu_roll, mu_spin, cohesion, compliance, complianceT, cRoll, cSpin.
int cb200_set_materials_nsc(cb200_system* sys, int n, const float* t) {
    System* S = SYS(sys);
    if (S)
        invalidate_step_graph(S);
    if (!S || n <= 0 || !t)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_set_materials_nsc");
    if (n > 1024)
        return set_error(CB200_ERR_UNSUPPORTED, "more than 1024 distinct materials");
    CB_CUDA(cudaSetDevice(S->device));
    std::vector<double> tab((size_t)n * n * NSC_MAT_STRIDE);
    bool has_aux = false;
    for (int a = 0; a < n; a++)
        for (int b = 0; b < n; b++) {
            const float* m1 = t + 11 * a;
            const float* m2 = t + 11 * b;
            double* o = tab.data() + ((size_t)a * n + b) * NSC_MAT_STRIDE;
            o[0] = std::min<float>(m1[1], m2[1]);  // sliding friction is the mu of the cone (…NSC.cpp:174)
            o[1] = std::min<float>(m1[2], m2[2]);
            o[2] = std::min<float>(m1[3], m2[3]);
            o[3] = std::min<float>(m1[5], m2[5]);
            o[4] = (float)(m1[7] + m2[7]);
            o[5] = (float)(m1[8] + m2[8]);
            o[6] = (float)(m1[9] + m2[9]);
            o[7] = (float)(m1[10] + m2[10]);
            for (int k = 1; k < 8; k++)
                has_aux |= o[k] != 0;
        }
    S->nsc_has_aux = has_aux ? 1 : 0;
    CB_TRY(S->mat_nsc.reserve(tab.size()));
    CB_TRY(upload(S, S->mat_nsc, 0, tab.data(), tab.size()));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    S->nm_nsc = n;
    return CB200_OK;
}

int cb200_upload_state(cb200_system* sys, int first, int n, const double* pos, const double* rot, const double* vel) {
    System* S = SYS(sys);
    if (!S || first < 0 || n < 0 || (uint32_t)(first + n) > S->nb)
        return set_error(CB200_ERR_ARG, "body range out of bounds");
    CB_CUDA(cudaSetDevice(S->device));
    if (pos)
        S->sphere_aabb_valid = false;
    if (pos)
        CB_TRY(upload(S, S->pos, 3 * (size_t)first, pos, 3 * (size_t)n));
    if (rot)
        CB_TRY(upload(S, S->rot, 4 * (size_t)first, rot, 4 * (size_t)n));
    if (vel)
        CB_TRY(upload(S, S->vel, 6 * (size_t)first, vel, 6 * (size_t)n));
    return CB200_OK;
}

int cb200_upload_applied_forces(cb200_system* sys, int first, int n, const double* ft) {
    System* S = SYS(sys);
    if (S)
        invalidate_step_graph(S);
    if (!S || first < 0 || n < 0 || (uint32_t)(first + n) > S->nb)
        return set_error(CB200_ERR_ARG, "body range out of bounds");
    CB_CUDA(cudaSetDevice(S->device));
    if (!ft) {
        CB_CUDA(cudaMemsetAsync(S->fapp.p + 6 * (size_t)first, 0, 6 * (size_t)n * 8, S->stream));
        return CB200_OK;
    }
    S->has_fapp = true;
    return upload(S, S->fapp, 6 * (size_t)first, ft, 6 * (size_t)n);
}

// ChCollisionSystemMulticore::SetActiveBoundingBox / collision_settings.use_aabb_active (MC/ChSettings.h:60-66)
int cb200_set_active_aabb(cb200_system* sys, int enabled, const double* aabb_min, const double* aabb_max) {
    System* S = SYS(sys);
    if (!S || (enabled && (!aabb_min || !aabb_max)))
        return set_error(CB200_ERR_ARG, "bad argument to cb200_set_active_aabb");
    invalidate_step_graph(S);
    S->use_aabb_active = enabled != 0;
    for (int a = 0; a < 3 && enabled; a++)
        S->active_min[a] = aabb_min[a], S->active_max[a] = aabb_max[a];
    return CB200_OK;
}

// ChBody::SetLimitSpeed / SetMaxLinVel / SetMaxAngVel of bodies [first, first+n); NULL switches the clamp off everywhere
int cb200_set_speed_limits(cb200_system* sys, int first, int n, const double* max_speed, const double* max_wvel) {
    System* S = SYS(sys);
    if (!S || first < 0 || n < 0 || (uint32_t)(first + n) > S->nb)
        return set_error(CB200_ERR_ARG, "body range out of bounds");
    CB_CUDA(cudaSetDevice(S->device));
    invalidate_step_graph(S);
    if (!max_speed || !max_wvel) {
        S->has_vlim = false;
        return CB200_OK;
    }
    std::vector<double> v(2 * (size_t)n);
    for (int i = 0; i < n; i++)
        v[2 * i] = max_speed[i], v[2 * i + 1] = max_wvel[i];
    CB_TRY(upload(S, S->vlim, 2 * (size_t)first, v.data(), v.size()));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    S->has_vlim = true;
    return CB200_OK;
}

// host_data.active_rigid / collide_rigid as the device holds them (the active box deactivates bodies on the device)
int cb200_download_body_flags(cb200_system* sys, int first, int n, char* active, char* collide) {
    System* S = SYS(sys);
    if (!S || first < 0 || n < 0 || (uint32_t)(first + n) > S->nb)
        return set_error(CB200_ERR_ARG, "body range out of bounds");
    CB_CUDA(cudaSetDevice(S->device));
    std::vector<uint8_t> fl((size_t)n);
    if (n)
        CB_CUDA(cudaMemcpyAsync(fl.data(), S->bflags.p + first, (size_t)n, cudaMemcpyDeviceToHost, S->stream));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    for (int i = 0; i < n; i++) {
        if (active)
            active[i] = (char)(fl[i] & 1u);
        if (collide)
            collide[i] = (char)((fl[i] >> 1) & 1u);
    }
    return CB200_OK;
}

int cb200_set_body_flags(cb200_system* sys, int first, int n, const char* active, const char* collide) {
    System* S = SYS(sys);
    if (S)
        invalidate_step_graph(S);
    if (!S || first < 0 || n < 0 || (uint32_t)(first + n) > S->nb || !active || !collide)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_set_body_flags");
    CB_CUDA(cudaSetDevice(S->device));
    std::vector<uint8_t> fl(n);
    for (int i = 0; i < n; i++)
        fl[i] = (uint8_t)((active[i] ? 1 : 0) | (collide[i] ? 2 : 0));
    CB_TRY(upload(S, S->bflags, (size_t)first, fl.data(), (size_t)n));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    S->meta_dirty = true;
    S->sphere_aabb_valid = false;
    return CB200_OK;
}

int cb200_update_forces(cb200_system* sys) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    return run_stage(S, [&] { return enqueue_update_forces(S); });
}

int cb200_run_collision(cb200_system* sys, int stages) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(clear_unsupported(S));
    return run_stage(S, [&]() -> int {
        // a retry after a capacity overflow restarts from the AABBs (OffsetAABB rewrites the boxes in place)
        CB_TRY(timer_start(S, "collision_broad"));
        if ((stages & 1) || ((stages & 2) && S->box_relative))
            CB_TRY(enqueue_aabb(S));
        if (stages & 2) {
            if (!S->pairs.p)
                CB_TRY(alloc_pairs(S, 8 * S->ns + 1024));
            CB_TRY(enqueue_bins(S));
            S->body_scratch_clean = false;  // the heavy flags are set again and no integration follows in this call
        }
        CB_TRY(timer_stop(S, "collision_broad"));
        CB_TRY(timer_start(S, "collision_narrow"));
        if (stages & 2)
            CB_TRY(enqueue_pairs_only(S, true, (stages & 4) != 0));
        else if (stages & 4)
            CB_TRY(enqueue_pairs_only(S, false, true));
        CB_TRY(timer_stop(S, "collision_narrow"));
        return CB200_OK;
    });
}

int cb200_run_solver_smc(cb200_system* sys) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    return run_stage(S, [&] { return enqueue_smc(S); });
}

int cb200_run_solver_nsc(cb200_system* sys) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    return run_stage(S, [&] { return enqueue_nsc(S); });
}

int cb200_integrate(cb200_system* sys) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    return run_stage(S, [&] { return enqueue_integrate(S); });
}

static int enqueue_one_step_(System* S);
static int enqueue_one_step(System* S) {
    S->whole_step = true;
    const int rc = enqueue_one_step_(S);
    S->whole_step = false;
    return rc;
}
static int enqueue_one_step_(System* S) {
    S->aux_ev_next = 0;
    CB_TRY(timer_start(S, "step"));
    CB_TRY(timer_start(S, "update"));
    if (S->set.system_type != 1)  // SMC: formed in k_body_reduce<true> (defer_update_forces below)
        CB_TRY(enqueue_update_forces(S));
    CB_TRY(timer_stop(S, "update"));
    CB_TRY(timer_start(S, "collision"));
    CB_TRY(timer_start(S, "collision_broad"));
    CB_TRY(enqueue_aabb(S));
    CB_TRY(enqueue_bins(S));
    CB_TRY(timer_stop(S, "collision_broad"));
    CB_TRY(timer_start(S, "collision_narrow"));  // in-bin pair tests fused with the single-contact narrowphase
    CB_TRY(enqueue_pairs_only(S, false, true, /*defer_geometry=*/S->set.system_type == 1));
    CB_TRY(timer_stop(S, "collision_narrow"));
    CB_TRY(timer_stop(S, "collision"));
    CB_TRY(timer_start(S, "advance"));
    S->defer_vel_update = S->defer_update_forces = S->set.system_type == 1;  // SMC: done in k_integrate<true> / k_body_reduce<true>
    S->fuse_aabb = !S->slab_on && !getenv("CB200_NO_FUSE_AABB");  // this integration also prepares the next step's sphere boxes
    int rc = S->set.system_type == 1 ? enqueue_smc(S) : enqueue_nsc(S);
    if (rc == CB200_OK) {
        CB_TRY(timer_stop(S, "advance"));
        CB_TRY(timer_start(S, "integrate"));
        rc = enqueue_integrate(S);
    }
    S->fuse_aabb = false;
    S->bbox_reset_done = false;
    S->defer_vel_update = S->defer_update_forces = false;
    CB_TRY(rc);
    CB_TRY(timer_stop(S, "integrate"));
    CB_TRY(timer_stop(S, "step"));
    return CB200_OK;
}

// One step captured into an executable graph (stage timers are skipped while capturing; every allocation a step can
// need was made by the directly enqueued step before it).
static int capture_step_graph(System* S, int par) {
    cudaGraph_t g = nullptr;
    const int64_t l0 = S->launches;
    if (S->step_graph[par]) {
        cudaGraphExecDestroy(S->step_graph[par]);
        S->step_graph[par] = nullptr;
    }
    S->step_graph_meta[par] = S->meta_dirty;  // the captured step rebuilds the per-shape records
    CB_CUDA(cudaStreamBeginCapture(S->stream, cudaStreamCaptureModeThreadLocal));
    S->capturing = true;
    const int rc = enqueue_one_step(S);
    S->capturing = false;
    const cudaError_t e = cudaStreamEndCapture(S->stream, &g);
    S->step_graph_launches[par] = (int)(S->launches - l0);
    S->launches = l0;
    if (rc != CB200_OK) {
        if (g)
            cudaGraphDestroy(g);
        return rc;
    }
    if (e != cudaSuccess)
        return set_error(CB200_ERR_CUDA, std::string("cudaStreamEndCapture: ") + cudaGetErrorString(e));
    const cudaError_t e2 = cudaGraphInstantiate(&S->step_graph[par], g, 0);
    cudaGraphDestroy(g);
    if (e2 != cudaSuccess) {
        S->step_graph[par] = nullptr;
        return set_error(CB200_ERR_CUDA, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e2));
    }
    return CB200_OK;
}

// the legacy / per-thread default streams cannot be captured
static bool graph_capable(const System* S) {
    return S->stream != nullptr && S->stream != cudaStreamLegacy && S->stream != cudaStreamPerThread &&
           !getenv("CB200_NO_GRAPH");
}

// A step through the cached graph (captured once a directly enqueued step has made every lazy allocation).
static int replay_or_enqueue_step(System* S) {
    // Single systems: the captured step relies on the previous step's integration for the spheres' records (and swaps
    // the double-buffered records), so it needs them valid and clean shape records.  Slab systems rebuild the shape
    // records every step (bodies come and go): their graph holds k_build_meta and k_aabb<false> and has no such needs.
    const bool fused = !S->slab_on && !getenv("CB200_NO_FUSE_AABB");
    if (S->direct_steps < 1 || (fused && (!S->sphere_aabb_valid || S->meta_dirty))) {
        S->direct_steps++;
        return enqueue_one_step(S);
    }
    const int par = fused ? S->wpos_parity : 0;
    if (!S->step_graph[par] || (S->meta_dirty && !S->step_graph_meta[par])) {
        CB_TRY(capture_step_graph(S, par));  // leaves the host-side state as if the step had been enqueued
    } else if (fused) {
        // what enqueue_one_step does on the host side for such a step
        std::swap(S->s_wpos, S->s_wpos_alt);
        S->wpos_parity ^= 1;
    }
    CB_CUDA(cudaGraphLaunch(S->step_graph[par], S->stream));
    S->launches += S->step_graph_launches[par];
    S->pairs_valid = false;
    S->contacts_geom_valid = false;  // the replayed step made a new key list; the records are those of an older one
    S->box_relative = true;
    S->sphere_aabb_valid = fused;
    if (!fused)
        S->meta_dirty = false;  // (the slab graph was captured with k_build_meta in it: meta is dirty at every capture)
    return CB200_OK;
}

// ChSystem::DoStepDynamics x nsteps.  Steps are enqueued back to back with no host synchronisation in between
// (all extents live on the device); the host syncs once per chunk of 64 steps to resolve timers and to find out
// whether a capacity was exceeded, in which case the steps that did not run (the device counts them) are redone.
int cb200_do_steps(cb200_system* sys, int nsteps) {
    System* S = SYS(sys);
    if (!S || nsteps < 0)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_do_steps");
    CB_CUDA(cudaSetDevice(S->device));
    if (S->nb == 0)
        return CB200_OK;
    for (auto& kv : S->timers) {
        kv.second.total_ms = 0;
        kv.second.steps = 0;
    }
    S->timed_steps = nsteps;
    CB_TRY(ensure_capacity(S));
    int remaining = nsteps, stalls = 0;
    // Launch-bound inner loop: with every extent on the device a step is the same launch sequence each time, so
    // after one directly enqueued step (which also samples the stage timers) the rest of a chunk replays a CUDA
    // graph of that step.  The graph is rebuilt per call and after every capacity growth (pointers change).
    const bool use_graph = S->set.system_type == 1 && nsteps >= 8 && graph_capable(S);
    while (remaining > 0) {
        const int chunk = remaining < 64 ? remaining : 64;
        CB_TRY(timer_start(S, "do_steps"));
        CB_CUDA(cudaMemsetAsync(&S->d_cnt->steps_done, 0, 4, S->stream));
        CB_TRY(clear_unsupported(S));
        CB_TRY(enqueue_one_step(S));
        S->direct_steps++;
        for (int s = 1; s < chunk; s++)
            CB_TRY(use_graph ? replay_or_enqueue_step(S) : enqueue_one_step(S));
        CB_TRY(timer_stop(S, "do_steps", 0));
        CB_TRY(flush_timers(S));
        CB_TRY(fetch_counters(S));
        const int done = (int)S->h_cnt->steps_done;
        S->timers["do_steps"].steps += done;
        remaining -= done;
        if (S->h_cnt->overflow)
            invalidate_step_graph(S);
        if (S->h_cnt->overflow) {
            CB_TRY(grow_after_overflow(S));
            if (++stalls > 16 && done == 0)
                return set_error(CB200_ERR_STATE, "capacity growth did not converge");
        } else if (done != chunk) {
            return set_error(CB200_ERR_STATE, "step accounting mismatch");
        }
        if (S->h_cnt->unsupported)
            return set_error(CB200_ERR_UNSUPPORTED,
                             std::to_string(S->h_cnt->unsupported) +
                                 " candidate pairs have a shape combination this backend cannot test yet");
    }
    return CB200_OK;
}

// DoStepDynamics for a caller whose state lives in HOST buffers (the reference's per-step gather from / scatter to the
// ChBody objects, MC/physics/ChSystemMulticore.cpp:329-377 and :134-153): pose up, step, pose and velocity down, as one
// pipelined call.  The collision stages need positions and rotations only, so the velocity upload (48 of the 104
// bytes per body) travels on a second stream while the main stream already runs AABBs, binning and the pair kernels;
// the force stage waits for it.  pos/rot/vel are in-out and should be pinned (cb200_host_alloc).
int cb200_step_host(cb200_system* sys, double* pos, double* rot, double* vel) {
    System* S = SYS(sys);
    if (!S || !pos || !rot || !vel)
        return set_error(CB200_ERR_ARG, "null argument to cb200_step_host");
    CB_CUDA(cudaSetDevice(S->device));
    if (S->nb == 0)
        return CB200_OK;
    if (S->slab_on)
        return set_error(CB200_ERR_STATE, "cb200_step_host is not available on a slab system");
    if (!S->copy_stream) {
        CB_CUDA(cudaStreamCreateWithFlags(&S->copy_stream, cudaStreamNonBlocking));
        CB_CUDA(cudaEventCreateWithFlags(&S->ev_copy, cudaEventDisableTiming));
        CB_CUDA(cudaEventCreateWithFlags(&S->ev_main, cudaEventDisableTiming));
    }
    CB_TRY(ensure_capacity(S));
    const size_t nb = S->nb;
    for (int attempt = 0; attempt < 8; attempt++) {
        S->sphere_aabb_valid = false;  // the pose comes from the host
        CB_CUDA(cudaMemsetAsync(&S->d_cnt->steps_done, 0, 4, S->stream));
        CB_TRY(clear_unsupported(S));
        CB_CUDA(cudaMemcpyAsync(S->pos.p, pos, 3 * nb * sizeof(double), cudaMemcpyHostToDevice, S->stream));
        CB_CUDA(cudaMemcpyAsync(S->rot.p, rot, 4 * nb * sizeof(double), cudaMemcpyHostToDevice, S->stream));
        // the velocities follow the pose over the link (not beside it: both would share the bandwidth and the
        // collision stages would start later), on the second stream, while the main stream computes
        CB_CUDA(cudaEventRecord(S->ev_main, S->stream));
        CB_CUDA(cudaStreamWaitEvent(S->copy_stream, S->ev_main, 0));
        CB_CUDA(cudaMemcpyAsync(S->vel.p, vel, 6 * nb * sizeof(double), cudaMemcpyHostToDevice, S->copy_stream));
        CB_CUDA(cudaEventRecord(S->ev_copy, S->copy_stream));
        CB_TRY(enqueue_aabb(S));
        CB_TRY(enqueue_bins(S));
        CB_TRY(enqueue_pairs_only(S, false, true, /*defer_geometry=*/S->set.system_type == 1));
        CB_CUDA(cudaStreamWaitEvent(S->stream, S->ev_copy, 0));
        if (S->set.system_type != 1)
            CB_TRY(enqueue_update_forces(S));
        S->defer_update_forces = S->set.system_type == 1;
        S->defer_vel_update = S->set.system_type == 1;
        int rc = S->set.system_type == 1 ? enqueue_smc(S) : enqueue_nsc(S);
        if (rc == CB200_OK)
            rc = enqueue_integrate(S);
        S->defer_vel_update = S->defer_update_forces = false;
        CB_TRY(rc);
        CB_CUDA(cudaMemcpyAsync(pos, S->pos.p, 3 * nb * sizeof(double), cudaMemcpyDeviceToHost, S->stream));
        CB_CUDA(cudaMemcpyAsync(rot, S->rot.p, 4 * nb * sizeof(double), cudaMemcpyDeviceToHost, S->stream));
        CB_CUDA(cudaMemcpyAsync(vel, S->vel.p, 6 * nb * sizeof(double), cudaMemcpyDeviceToHost, S->stream));
        S->direct_steps++;
        CB_TRY(fetch_counters(S));
        CB_TRY(flush_timers(S));
        if (S->h_cnt->unsupported)
            return set_error(CB200_ERR_UNSUPPORTED, std::to_string(S->h_cnt->unsupported) +
                                                        " candidate pairs have a shape combination this backend cannot test yet");
        if (!S->h_cnt->overflow)
            return S->h_cnt->steps_done == 1 ? CB200_OK : set_error(CB200_ERR_STATE, "step accounting mismatch");
        // a capacity was exceeded: every kernel after it was a no-op, the host buffers got the uploaded state back
        CB_TRY(grow_after_overflow(S));
    }
    return set_error(CB200_ERR_STATE, "capacity growth did not converge");
}

// Asynchronous stepping for callers that interleave their own work on the stream (slab exchange): no host
// synchronisation and no capacity recovery here; cb200_check() reports a capacity overflow as an error.
int cb200_enqueue_steps(cb200_system* sys, int nsteps) {
    System* S = SYS(sys);
    if (!S || nsteps < 0)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_enqueue_steps");
    CB_CUDA(cudaSetDevice(S->device));
    if (S->nb == 0)
        return CB200_OK;
    CB_TRY(ensure_capacity(S));
    S->timed_steps += nsteps;
    // the first step of a call is enqueued directly (it samples the stage timers), the others replay the graph;
    // single-step callers (slab stepping) sample one step in 16
    const bool use_graph = S->set.system_type == 1 && graph_capable(S);
    for (int s = 0; s < nsteps; s++) {
        if (use_graph && (s > 0 || (S->async_calls++ & 15)))
            CB_TRY(replay_or_enqueue_step(S));
        else {
            CB_TRY(enqueue_one_step(S));
            S->direct_steps++;
        }
    }
    return CB200_OK;
}

int cb200_check(cb200_system* sys) {
    System* S = SYS(sys);
    if (!S)
        return set_error(CB200_ERR_ARG, "null system");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(flush_timers(S));
    CB_TRY(fetch_counters(S));
    if (S->h_cnt->overflow)
        return set_error(CB200_ERR_STATE, "a device capacity was exceeded during asynchronous stepping (overflow mask " +
                                              std::to_string(S->h_cnt->overflow) + "); steps after it did not run");
    if (S->h_cnt->unsupported) {
        CB_TRY(clear_unsupported(S));  // read and reported: the next check starts clean
        return set_error(CB200_ERR_UNSUPPORTED, std::to_string(S->h_cnt->unsupported) +
                                                    " candidate pairs have a shape combination this backend cannot test yet");
    }
    return CB200_OK;
}

int cb200_reset_timers(cb200_system* sys) {
    System* S = SYS(sys);
    if (!S)
        return set_error(CB200_ERR_ARG, "null system");
    CB_TRY(flush_timers(S));
    for (auto& kv : S->timers)
        kv.second.total_ms = 0, kv.second.steps = 0;
    S->timed_steps = 0;
    return CB200_OK;
}

int cb200_set_stream(cb200_system* sys, void* cuda_stream, int use_own) {
    System* S = SYS(sys);
    if (S)
        invalidate_step_graph(S);
    if (!S)
        return set_error(CB200_ERR_ARG, "null system");
    CB_CUDA(cudaSetDevice(S->device));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    S->stream = use_own ? S->own_stream : (cudaStream_t)cuda_stream;  // 0 is a valid handle: the legacy default stream
    return CB200_OK;
}

int cb200_synchronize(cb200_system* sys) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}

int cb200_get_counts(cb200_system* sys, cb200_counts* out) {
    System* S = SYS(sys);
    if (!S || !out)
        return set_error(CB200_ERR_ARG, "null argument");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(fetch_counters(S));
    const DevCounters& c = *S->h_cnt;
    if (c.unsupported)
        return set_error(CB200_ERR_UNSUPPORTED,
                         std::to_string(c.unsupported) +
                             " candidate pairs have a shape combination this backend cannot test yet");
    memset(out, 0, sizeof(*out));
    out->num_bodies = S->nb;
    out->num_shapes = S->ns;
    for (int a = 0; a < 3; a++) {
        out->bins_per_axis[a] = (uint32_t)c.bins[a];
        out->min_bounding_point[a] = c.origin[a];
        out->max_bounding_point[a] = c.maxpt[a];
        out->bin_size[a] = c.bin_size[a];
    }
    out->num_bins = c.num_bins;
    out->num_bin_aabb_intersections = c.n_intersections;
    out->num_active_bins = c.n_active_bins;
    out->num_possible_collisions = c.n_pairs;
    out->num_potential_contacts = c.n_pairs;
    out->num_rigid_contacts = c.n_contacts;
    out->ct_body_count = c.ct_body_count;
    out->solver_iterations = c.nsc_iters;
    out->solver_slots = c.nsc_slots;
    out->solver_residual = c.nsc_residual;
    out->solver_objective = c.nsc_objective;
    return CB200_OK;
}

int cb200_download_state(cb200_system* sys, int first, int n, double* pos, double* rot, double* vel) {
    System* S = SYS(sys);
    if (!S || first < 0 || n < 0 || (uint32_t)(first + n) > S->nb)
        return set_error(CB200_ERR_ARG, "body range out of bounds");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(download(S, S->pos.p + 3 * (size_t)first, pos, 3 * (size_t)n));
    CB_TRY(download(S, S->rot.p + 4 * (size_t)first, rot, 4 * (size_t)n));
    CB_TRY(download(S, S->vel.p + 6 * (size_t)first, vel, 6 * (size_t)n));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}

int cb200_download_aabb(cb200_system* sys, double* aabb_min, double* aabb_max) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    const size_t ns = S->ns;
    if (ns == 0)
        return CB200_OK;
    std::vector<ShapeBox> box(ns);
    std::vector<double> wpos(4 * ns);
    std::vector<int32_t> type(ns);
    CB_TRY(download(S, S->s_box.p, box.data(), ns));
    CB_TRY(download(S, S->s_wpos.p, wpos.data(), 4 * ns));
    CB_TRY(download(S, S->s_type.p, type.data(), ns));
    CB_TRY(fetch_counters(S));
    const double env = S->set.collision_envelope;
    const double* org = S->h_cnt->origin;
    for (size_t i = 0; i < ns; i++) {
        double mn[3], mx[3];
        if (type[i] == CB200_SPHERE) {
            // spheres keep no box record on the device; same operations as k_aabb + OffsetAABB
            const double r = wpos[4 * i + 3] + env;
            for (int a = 0; a < 3; a++) {
                mn[a] = wpos[4 * i + a] - r;
                mx[a] = wpos[4 * i + a] + r;
                if (S->box_relative) {
                    mn[a] = mn[a] - org[a];
                    mx[a] = mx[a] - org[a];
                }
            }
        } else {
            for (int a = 0; a < 3; a++)
                mn[a] = box[i].mn[a], mx[a] = box[i].mx[a];
        }
        for (int a = 0; a < 3; a++) {
            if (aabb_min)
                aabb_min[3 * i + a] = mn[a];
            if (aabb_max)
                aabb_max[3 * i + a] = mx[a];
        }
    }
    return CB200_OK;
}

int cb200_download_bins(cb200_system* sys, uint32_t* bin_number, uint32_t* bin_aabb_number, uint32_t* bin_active,
                        uint32_t* bin_start_index) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(fetch_counters(S));
    const uint32_t ni = S->h_cnt->n_intersections, nbins = S->h_cnt->num_bins;
    if (ni == 0)
        return CB200_OK;
    std::vector<uint64_t> packed((size_t)nbins + 1);
    CB_TRY(download(S, S->bin_start.p, packed.data(), packed.size()));
    CB_TRY(download(S, S->ent_shape.p, bin_aabb_number, (size_t)ni));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    uint32_t a = 0;
    for (uint32_t b = 0; b < nbins; b++) {
        const uint32_t s = (uint32_t)packed[b], e = (uint32_t)packed[b + 1];
        if (e == s)
            continue;
        if (bin_active)
            bin_active[a] = b;
        if (bin_start_index)
            bin_start_index[a] = s;
        if (bin_number)
            for (uint32_t k = s; k < e; k++)
                bin_number[k] = b;
        a++;
    }
    if (bin_start_index)
        bin_start_index[a] = ni;
    return CB200_OK;
}

// The candidate pair list is not part of the step (pairs are resolved into contacts on the fly); it is
// materialised here, from the bins of the last collision stage, when somebody asks for it.
int cb200_download_pairs(cb200_system* sys, int64_t* pairs) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    if (S->ns == 0)
        return CB200_OK;
    if (!S->pairs_valid) {
        if (!S->pairs.p)
            CB_TRY(alloc_pairs(S, 8 * S->ns + 1024));
        CB_TRY(run_stage(S, [&] { return enqueue_pairs_only(S, true, false); }));
    } else {
        CB_TRY(fetch_counters(S));
    }
    CB_TRY(download(S, S->pairs.p, pairs, (size_t)S->h_cnt->n_pairs));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}

int cb200_download_contacts(cb200_system* sys, double* normal, double* pt_a, double* pt_b, double* depth,
                            double* eff_radius, int32_t* body_ids, int64_t* shape_ids) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(ensure_contact_geometry(S));  // an SMC step leaves the records unwritten until somebody reads them
    CB_TRY(fetch_counters(S));
    const size_t nc = S->h_cnt->n_contacts;
    CB_TRY(download(S, S->c_norm.p, normal, 3 * nc));
    CB_TRY(download(S, S->c_pta.p, pt_a, 3 * nc));
    CB_TRY(download(S, S->c_ptb.p, pt_b, 3 * nc));
    CB_TRY(download(S, S->c_depth.p, depth, nc));
    CB_TRY(download(S, S->c_erad.p, eff_radius, nc));
    CB_TRY(download(S, S->c_bids.p, body_ids, 2 * nc));
    CB_TRY(download(S, S->c_shapes.p, shape_ids, nc));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}

int cb200_download_contact_forces(cb200_system* sys, double* ct_force, double* ct_torque) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(fetch_counters(S));
    const size_t nc = S->h_cnt->n_contacts;
    if (nc == 0)
        return CB200_OK;
    std::vector<double> r(CT_STRIDE * nc);
    CB_TRY(download(S, S->ct_force.p, r.data(), CT_STRIDE * nc));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    for (size_t i = 0; i < nc; i++)
        for (int a = 0; a < 3; a++) {
            if (ct_force) {
                ct_force[6 * i + a] = -r[CT_STRIDE * i + a];
                ct_force[6 * i + 3 + a] = r[CT_STRIDE * i + a];
            }
            if (ct_torque) {
                ct_torque[6 * i + a] = r[CT_STRIDE * i + 4 + a];
                ct_torque[6 * i + 3 + a] = r[CT_STRIDE * i + 8 + a];
            }
        }
    return CB200_OK;
}

// ---- contact callbacks: the host walks the contact list between the collision stage and the solver ----

// ChCollisionSystem::NarrowphaseCallback (OnNarrowphase returning false drops the contact,
// src/chrono/collision/multicore/ChCollisionSystemMulticore.cpp:361-368): keep[i] != 0 keeps contact i; the list is
// compacted in place, order preserved.
int cb200_filter_contacts(cb200_system* sys, int n, const uint8_t* keep) {
    System* S = SYS(sys);
    if (!S || n < 0 || (n && !keep))
        return set_error(CB200_ERR_ARG, "bad argument to cb200_filter_contacts");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(fetch_counters(S));
    if ((uint32_t)n != S->h_cnt->n_contacts)
        return set_error(CB200_ERR_ARG, "keep[] must have one entry per contact of the current list");
    std::vector<uint32_t> idx;
    idx.reserve(n);
    for (int i = 0; i < n; i++)
        if (keep[i])
            idx.push_back((uint32_t)i);
    const uint32_t m = (uint32_t)idx.size();
    if (m == (uint32_t)n)
        return CB200_OK;
    uint32_t *d_idx = nullptr, *scratch = nullptr;
    CB_CUDA(cudaMalloc(&d_idx, (size_t)std::max(m, 1u) * 4));
    CB_CUDA(cudaMalloc(&scratch, (size_t)std::max(m, 1u) * 24));  // largest record: 3 doubles
    CB_CUDA(cudaMemcpyAsync(d_idx, idx.data(), (size_t)m * 4, cudaMemcpyHostToDevice, S->stream));
    int rc = CB200_OK;
#define COMPACT(W, buf) if (rc == CB200_OK) rc = compact_array<W>(S, S->buf, m, d_idx, scratch)
    COMPACT(2, c_shapes);
    COMPACT(2, c_bids);
    COMPACT(1, c_mats);
    if (use_complex(S))
        COMPACT(1, c_src);
    if (S->contacts_geom_valid) {
        COMPACT(6, c_norm);
        COMPACT(6, c_pta);
        COMPACT(6, c_ptb);
        COMPACT(2, c_depth);
        COMPACT(2, c_erad);
    }
#undef COMPACT
    if (rc == CB200_OK) {
        CB_CUDA(cudaMemcpyAsync(&S->d_cnt->n_contacts, &m, 4, cudaMemcpyHostToDevice, S->stream));
        CB_CUDA(cudaStreamSynchronize(S->stream));
        S->h_cnt->n_contacts = m;
        S->mat_ovr_valid = false;
    }
    cudaFree(d_idx), cudaFree(scratch);
    return rc;
}

// The composite material of every contact as ChContactContainerMulticoreSMC/NSC::AddContact forms it before the
// AddContactCallback sees it (MC/collision/ChContactContainerMulticoreSMC.cpp:123-174, ...NSC.cpp:126-187).
// SMC, 13 per contact: mu, muRoll, muSpin, E, G, adhesion, adhesionMultDMT, adhesionSPerko, cr, kn, kt, gn, gt.
// NSC, 8 per contact: mu (sliding), muRoll, muSpin, cohesion, compliance, complianceT, complianceRoll, complianceSpin.
int cb200_download_contact_materials(cb200_system* sys, int capacity, double* composite, int* stride) {
    System* S = SYS(sys);
    if (!S || !stride)
        return set_error(CB200_ERR_ARG, "null argument");
    CB_CUDA(cudaSetDevice(S->device));
    const bool smc = S->set.system_type == 1;
    const int W = smc ? 13 : 8, TW = smc ? (int)SMC_MAT_STRIDE : (int)NSC_MAT_STRIDE;
    *stride = W;
    CB_TRY(fetch_counters(S));
    const size_t nc = S->h_cnt->n_contacts;
    if (nc == 0 || !composite)
        return CB200_OK;
    if ((size_t)capacity < nc)
        return set_error(CB200_ERR_ARG, "output capacity smaller than num_rigid_contacts");
    const int nm = smc ? S->nm_smc : S->nm_nsc;
    DevBuf<double>& tab = smc ? S->mat_smc : S->mat_nsc;
    if (nm == 0 || !tab.p)
        return set_error(CB200_ERR_STATE, "no materials set");
    std::vector<double> t((size_t)nm * nm * TW);
    std::vector<uint32_t> mm(nc);
    CB_TRY(download(S, tab.p, t.data(), t.size()));
    CB_TRY(download(S, S->c_mats.p, mm.data(), nc));
    std::vector<double> ovr;
    if (S->mat_ovr_valid) {
        ovr.resize(nc * TW);
        CB_TRY(download(S, S->mat_ovr.p, ovr.data(), ovr.size()));
    }
    CB_CUDA(cudaStreamSynchronize(S->stream));
    for (size_t i = 0; i < nc; i++) {
        const double* M = S->mat_ovr_valid ? ovr.data() + i * TW
                                           : t.data() + ((size_t)(mm[i] & 0xffff) * nm + (mm[i] >> 16)) * TW;
        for (int k = 0; k < W; k++)
            composite[W * i + k] = M[k];
    }
    return CB200_OK;
}

// ... and the records the callback handed back: they replace the table lookup for the current contact list (until the
// next collision pass).
int cb200_upload_contact_materials(cb200_system* sys, int n, const double* composite) {
    System* S = SYS(sys);
    if (!S || n < 0 || (n && !composite))
        return set_error(CB200_ERR_ARG, "bad argument to cb200_upload_contact_materials");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(fetch_counters(S));
    if ((uint32_t)n != S->h_cnt->n_contacts)
        return set_error(CB200_ERR_ARG, "one record per contact of the current list expected");
    if (n == 0)
        return CB200_OK;
    const bool smc = S->set.system_type == 1;
    const int W = smc ? 13 : 8, TW = smc ? (int)SMC_MAT_STRIDE : (int)NSC_MAT_STRIDE;
    std::vector<double> h((size_t)n * TW);
    bool aux = false;
    for (int i = 0; i < n; i++) {
        double* o = h.data() + (size_t)i * TW;
        for (int k = 0; k < W; k++)
            o[k] = composite[(size_t)W * i + k];
        if (smc) {  // the tabulated Hertz damping ratio follows the (possibly changed) restitution
            const double cr = o[8], eps = std::numeric_limits<double>::epsilon();
            const double loge = (cr < eps) ? std::log(eps) : std::log(cr);
            o[13] = loge / std::sqrt(loge * loge + M_PI * M_PI);
        } else {
            for (int k = 1; k < 8; k++)
                aux |= o[k] != 0;
        }
    }
    if (S->mat_ovr.cap < h.size())
        invalidate_step_graph(S);
    CB_TRY(S->mat_ovr.reserve(h.size()));
    CB_TRY(upload(S, S->mat_ovr, 0, h.data(), h.size()));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    if (aux && !S->nsc_has_aux) {
        S->nsc_has_aux = 1;
        invalidate_step_graph(S);
    }
    S->mat_ovr_valid = true;
    return CB200_OK;
}

// ChContactContainerMulticore::ReportAllContacts, batched: one row per contact with the arguments of OnReportContact
int cb200_report_contacts(cb200_system* sys, int capacity, double* pA, double* pB, double* plane, double* distance,
                          double* eff_radius, double* force, double* torque, int32_t* body_ids, int* count) {
    System* S = SYS(sys);
    if (!S || !count)
        return set_error(CB200_ERR_ARG, "null argument");
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(fetch_counters(S));
    const size_t nc = S->h_cnt->n_contacts;
    *count = (int)nc;
    if (nc == 0 || capacity <= 0)
        return CB200_OK;
    if ((size_t)capacity < nc)
        return set_error(CB200_ERR_ARG, "output capacity smaller than num_rigid_contacts");
    CB_TRY(enqueue_report(S));
    const int W = 21;
    std::vector<double> h(W * nc);
    CB_TRY(download(S, S->report.p, h.data(), W * nc));
    CB_TRY(download(S, S->c_depth.p, distance, nc));
    CB_TRY(download(S, S->c_erad.p, eff_radius, nc));
    CB_TRY(download(S, S->c_bids.p, body_ids, 2 * nc));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    for (size_t i = 0; i < nc; i++) {
        const double* r = h.data() + W * i;
        for (int a = 0; a < 3; a++) {
            if (pA) pA[3 * i + a] = r[a];
            if (pB) pB[3 * i + a] = r[3 + a];
            if (force) force[3 * i + a] = r[15 + a];
            if (torque) torque[3 * i + a] = r[18 + a];
        }
        if (plane)
            for (int a = 0; a < 9; a++)
                plane[9 * i + a] = r[6 + a];
    }
    return CB200_OK;
}

int cb200_download_body_contact_forces(cb200_system* sys, double* force, double* torque) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    const size_t nb = S->nb;
    if (nb == 0)
        return CB200_OK;
    if (S->b_force.cap < 6 * nb)
        return set_error(CB200_ERR_STATE, "no solve has run yet");
    if (S->set.system_type == 0)
        CB_TRY(nsc_contact_forces_inactive(S));
    std::vector<double> h(6 * nb);
    CB_TRY(download(S, S->b_force.p, h.data(), 6 * nb));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    for (size_t b = 0; b < nb; b++)
        for (int a = 0; a < 3; a++) {
            if (force)
                force[3 * b + a] = h[6 * b + a];
            if (torque)
                torque[3 * b + a] = h[6 * b + 3 + a];
        }
    return CB200_OK;
}

int cb200_download_hf(cb200_system* sys, double* hf) {
    System* S = SYS(sys);
    CB_CUDA(cudaSetDevice(S->device));
    CB_TRY(download(S, S->hf.p, hf, 6 * (size_t)S->nb));
    CB_CUDA(cudaStreamSynchronize(S->stream));
    return CB200_OK;
}

int cb200_get_timer(cb200_system* sys, const char* name, double* seconds) {
    System* S = SYS(sys);
    if (!S || !name || !seconds)
        return set_error(CB200_ERR_ARG, "null argument");
    CB_TRY(flush_timers(S));  // resolve intervals recorded since the last synchronising call
    auto it = S->timers.find(name);
    if (it == S->timers.end())
        return set_error(CB200_ERR_ARG, std::string("unknown timer '") + name + "'");
    *seconds = it->second.total_ms * 1e-3 / (it->second.steps > 0 ? it->second.steps : 1);
    return CB200_OK;
}

int64_t cb200_kernel_launches(cb200_system* sys) { return SYS(sys)->launches; }

// Self-test of the one-pass scan every list of the step is built with: exclusive scan of n pseudo-random counts, run
// `repeats` times back to back (the scratch is not cleared between scans), compared with a host scan.
int cb200_selftest_scan(cb200_system* sys, int64_t n, uint32_t seed, int repeats, int64_t* mismatches) {
    System* S = SYS(sys);
    if (!S || n < 0 || n > (int64_t)1 << 31 || !mismatches || repeats < 1)
        return set_error(CB200_ERR_ARG, "bad argument to cb200_selftest_scan");
    CB_CUDA(cudaSetDevice(S->device));
    std::vector<uint32_t> h((size_t)n + 1), ref((size_t)n + 1), got((size_t)n + 1);
    uint32_t x = seed * 2654435761u + 1u, acc = 0;
    for (int64_t i = 0; i < n; i++) {
        x = x * 1664525u + 1013904223u;
        h[i] = (x >> 24) % 7u;
        ref[i] = acc;
        acc += h[i];
    }
    ref[n] = acc;
    uint32_t *d_in = nullptr, *d_out = nullptr, *d_n = nullptr;
    CB_CUDA(cudaMalloc(&d_in, ((size_t)n + 4) * 4));
    CB_CUDA(cudaMalloc(&d_out, ((size_t)n + 4) * 4));
    CB_CUDA(cudaMalloc(&d_n, 4));
    const uint32_t n32 = (uint32_t)n;
    CB_CUDA(cudaMemcpyAsync(d_in, h.data(), (size_t)n * 4, cudaMemcpyHostToDevice, S->stream));
    CB_CUDA(cudaMemcpyAsync(d_n, &n32, 4, cudaMemcpyHostToDevice, S->stream));
    *mismatches = 0;
    int rc = CB200_OK;
    for (int r = 0; r < repeats && rc == CB200_OK; r++) {
        CB_CUDA(cudaMemsetAsync(d_out, 0xff, ((size_t)n + 1) * 4, S->stream));
        rc = scan_u32(S, d_in, d_out, d_n, (size_t)n + (size_t)(r % 3) * 1000);  // capacity above the extent, as in a step
        if (rc != CB200_OK)
            break;
        CB_CUDA(cudaMemcpyAsync(got.data(), d_out, ((size_t)n + 1) * 4, cudaMemcpyDeviceToHost, S->stream));
        CB_CUDA(cudaStreamSynchronize(S->stream));
        for (int64_t i = 0; i <= n; i++)
            *mismatches += got[i] != ref[i];
    }
    cudaFree(d_in), cudaFree(d_out), cudaFree(d_n);
    return rc;
}

int cb200_device_array(cb200_system* sys, const char* name, void** dev_ptr, int64_t* count) {
    System* S = SYS(sys);
    if (!S || !name || !dev_ptr || !count)
        return set_error(CB200_ERR_ARG, "null argument");
    const std::string n(name);
    if (n.rfind("contact_", 0) == 0)
        CB_TRY(ensure_contact_geometry(S));
    struct E { const char* k; void* p; int64_t c; };
    const E tab[] = {
        {"pos", S->pos.p, 3 * (int64_t)S->nb},           {"rot", S->rot.p, 4 * (int64_t)S->nb},
        {"vel", S->vel.p, 6 * (int64_t)S->nb},           {"hf", S->hf.p, 6 * (int64_t)S->nb},
        {"mass", S->mass.p, (int64_t)S->nb},             {"contact_normal", S->c_norm.p, 3 * (int64_t)S->h_cnt->n_contacts},
        {"contact_pt_a", S->c_pta.p, 3 * (int64_t)S->h_cnt->n_contacts},
        {"contact_pt_b", S->c_ptb.p, 3 * (int64_t)S->h_cnt->n_contacts},
        {"contact_depth", S->c_depth.p, (int64_t)S->h_cnt->n_contacts},
        {"contact_body_ids", S->c_bids.p, 2 * (int64_t)S->h_cnt->n_contacts},
        {"pair_shape_ids", S->pairs.p, (int64_t)S->h_cnt->n_pairs},
        {"gamma", S->gamma.p, 3 * (int64_t)S->h_cnt->n_contacts},
    };
    for (const E& e : tab)
        if (n == e.k) {
            *dev_ptr = e.p;
            *count = e.c;
            return CB200_OK;
        }
    return set_error(CB200_ERR_ARG, "unknown device array '" + n + "'");
}

}  // extern "C"
