/* chrono_b200.h -- C ABI of libchrono_b200: the B200 (sm_100a) backend for the Chrono::Multicore granular-contact
 * hot path (collision pipeline, SMC force accumulation, NSC APGD solve, body integration).
 *
 * The reference has no C ABI for this path; its seam is C++ virtual dispatch over two plain data structs
 * (SURVEY.md 8b).  Every entry point below names the reference interface it replaces (paths relative to the
 * reference root, MC = src/chrono_multicore, CMC = src/chrono/collision/multicore).  INTEGRATION.md shows the
 * C++ shim a Chrono maintainer adds on top (ChCollisionSystemB200 / ChIterativeSolverB200).
 *
 * Conventions
 *   - plain pointers and sizes, no C++ or torch types; all arrays are HOST pointers unless a name ends in _dev
 *   - 3-vectors are packed (x,y,z) doubles (the reference pads real3 to 4 doubles, MM/real3.h:85-90; the shim
 *     strips the pad), quaternions are (w,x,y,z), velocities 6 per body (vx,vy,vz, wx,wy,wz) with w in the body
 *     frame, as in host_data.v (MC/physics/ChSystemMulticore.cpp:350-362)
 *   - every function returns CB200_OK or a negative status; cb200_last_error() gives the message
 *     (the reference throws std::runtime_error / asserts; the shim rethrows)
 *   - one caller thread per system (same rule as the reference, SURVEY.md 8b "Threading")
 *   - there is no CPU fallback: cb200_create fails if no CUDA device is usable.
 */
#ifndef CHRONO_B200_H
#define CHRONO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct cb200_system cb200_system; /* opaque handle: owns one CUDA stream and all device SoA state */

enum {
    CB200_OK = 0,
    CB200_ERR_CUDA = -1,      /* a CUDA runtime call failed */
    CB200_ERR_ARG = -2,       /* bad argument */
    CB200_ERR_UNSUPPORTED = -3, /* feature of the reference not covered by this backend */
    CB200_ERR_STATE = -4      /* call sequence error (e.g. step before bodies were added) */
};

/* Shape types: values of chrono::ChCollisionShape::Type (src/chrono/collision/ChCollisionShape.h:31-53). */
enum {
    CB200_SPHERE = 0, CB200_ELLIPSOID = 1, CB200_BOX = 2, CB200_CYLINDER = 3, CB200_CYLSHELL = 4,
    CB200_CONVEXHULL = 5, CB200_TRIANGLE = 10, CB200_CAPSULE = 12, CB200_CONE = 13, CB200_ROUNDEDBOX = 14,
    CB200_ROUNDEDCYL = 15
};

/* Mirror of collision_settings + solver_settings + step/gravity (MC/ChSettings.h:39-246).  Re-read every step,
 * like ChCollisionSystemChronoMulticore::PreProcess does (MC/collision/ChCollisionSystemChronoMulticore.cpp:67-77). */
typedef struct cb200_settings {
    /* --- collision_settings (MC/ChSettings.h:39-89) */
    double collision_envelope;
    int32_t broadphase_grid;       /* ChBroadphase::GridType: 0 FIXED_RESOLUTION, 1 FIXED_BIN_SIZE, 2 FIXED_DENSITY */
    int32_t bins_per_axis[3];
    double bin_size[3];
    double grid_density;
    int32_t narrowphase_algorithm; /* ChNarrowphase::Algorithm: 0 MPR, 1 PRIMS, 2 HYBRID */
    /* --- system */
    int32_t system_type;           /* 0 = NSC (ChSystemMulticoreNSC), 1 = SMC (ChSystemMulticoreSMC) */
    double step_size;
    double gravity[3];
    /* --- solver_settings, SMC part (MC/ChSettings.h:93-218) */
    int32_t contact_force_model;   /* ChSystemSMC::ContactForceModel: 0 Hooke, 1 Hertz, 2 PlainCoulomb, 3 Flores */
    int32_t adhesion_force_model;  /* 0 Constant, 1 DMT, 2 Perko */
    int32_t tangential_displ_mode; /* 0 None, 1 OneStep, 2 MultiStep */
    int32_t use_material_properties;
    double characteristic_vel;
    double min_slip_vel;
    double min_roll_vel;
    double min_spin_vel;
    /* --- solver_settings, NSC part */
    int32_t solver_type;           /* SolverType: 0 APGD, 1 APGDREF (subset of MC/ChMulticoreDefines.h) */
    int32_t solver_mode;           /* SolverMode: 0 NORMAL, 1 SLIDING, 2 SPINNING */
    int32_t max_iteration_normal;
    int32_t max_iteration_sliding;
    int32_t max_iteration_spinning;
    int32_t cache_step_length;
    double alpha;
    double contact_recovery_speed;
    double tolerance;
    /* --- backend-only knobs (no reference counterpart) */
    int32_t deterministic;         /* 1: fixed summation order everywhere (default); 0: allow FP64 atomics */
    /* --- solver_settings, NSC part, continued (carved out of the reserved words: the struct keeps its size) */
    int32_t test_objective;        /* APGD stops on objective <= tolerance_objective instead of residual < tol_speed
                                      (MC/solver/ChSolverMulticoreAPGD.cpp:184-192, MC/ChSettings.h:100; default 0) */
    double tolerance_objective;    /* MC/ChSettings.h:98, default 1e-6 */
    int32_t reserved[4];
} cb200_settings;

/* Counters of the last step: measures_container (MC/ChMeasures.h:34-120) + cd_data counters
 * (CMC/ChCollisionData.h:185-221). */
typedef struct cb200_counts {
    uint32_t num_bodies;
    uint32_t num_shapes;
    uint32_t bins_per_axis[3];
    uint32_t num_bins;
    uint32_t num_bin_aabb_intersections; /* N_i */
    uint32_t num_active_bins;            /* N_a */
    uint32_t num_possible_collisions;    /* N_p */
    uint32_t num_potential_contacts;     /* pre-compaction slots, CMC/ChNarrowphase.cpp:136 */
    uint32_t num_rigid_contacts;         /* N_c */
    uint32_t ct_body_count;              /* bodies with >= 1 contact, MC/solver/ChIterativeSolverMulticoreSMC.cpp:726 */
    uint32_t solver_iterations;          /* NSC: iterations of the last Solve */
    uint32_t solver_slots;               /* NSC: ... plus its back-tracking retries (:148-154), summed over the phases */
    uint32_t reserved[2];
    double min_bounding_point[3];        /* global_origin */
    double max_bounding_point[3];
    double bin_size[3];
    double solver_residual;
    double solver_objective;
} cb200_counts;

const char* cb200_last_error(void);
int cb200_device_count(void);

/* ChSystemMulticore ctor/dtor (MC/physics/ChSystemMulticore.cpp:36-71): creates the data manager equivalent on
 * CUDA device `device`. */
int cb200_create(int device, cb200_system** out);
void cb200_destroy(cb200_system* sys);

/* Page-locked host buffers for the upload/download calls below (the reference's host_data vectors are plain
 * std::vector, MC/ChDataManager.h:204-324; pinning them lets the copies run at PCIe speed).  Optional: every
 * entry point also accepts pageable memory. */
int cb200_host_alloc(uint64_t bytes, void** out);
int cb200_host_free(void* p);

/* sys.GetSettings() write-back (MC/physics/ChSystemMulticore.cpp:684-686). */
void cb200_default_settings(cb200_settings* s); /* defaults of MC/ChSettings.h:41-48,96-131 */
int cb200_set_settings(cb200_system* sys, const cb200_settings* s);

/* ChSystemMulticore::AddBody (MC/physics/ChSystemMulticore.cpp:205-226) for n bodies at once: mass, body-frame
 * inertia diagonal (ChBody::GetInertiaXX), the diagonal of ChBody::GetInvInertia() as Chrono computed it (NULL:
 * 1/inertia_xx is used) and flags.  Products of inertia: cb200_set_inertia_full.
 * Returns the index of the first new body in *first. */
int cb200_append_bodies(cb200_system* sys, int n, const double* mass, const double* inertia_xx,
                        const double* inv_inertia_xx, const char* active, const char* collide, int* first);

/* use_full_inertia_tensor (MC/ChSettings.h:107, the reference's default; consumed by ComputeInvMassMatrix,
 * MC/solver/ChIterativeSolverMulticore.cpp:61-147, and by ChBody::ComputeGyro): inertia matrix and its inverse, both in
 * the body frame, 9 doubles each (row-major, symmetric), for bodies first .. first+n-1 -- ChBody::GetInertia() and
 * GetInvInertia().  Bodies never named here keep the principal moments of cb200_append_bodies.  Not combined with
 * bilateral rows or slabs yet (CB200_ERR_UNSUPPORTED). */
int cb200_set_inertia_full(cb200_system* sys, int first, int n, const double* inertia9, const double* inv_inertia9);

/* ChCollisionSystemMulticore::Add (CMC/ChCollisionSystemMulticore.cpp:86-187): append n shapes as shape_container
 * slices.  dims is 4 doubles per shape: sphere (r,-,-,-); ellipsoid/box/cylinder/cylshell/cone (B.x,B.y,B.z,-);
 * capsule (radius, hlen-as-stored,-,-) [sic, SURVEY Appendix A.1]; rounded box/cyl (x,y,z,sweep radius).
 * triangle / convex hull (first vertex, vertex count, -, -) into the pool of cb200_append_shape_vertices.
 * fam is (group, mask) int16 pairs.  material_id indexes the table given to cb200_set_materials_*. */
int cb200_append_shapes(cb200_system* sys, int n, const int32_t* type, const uint32_t* body_id, const int16_t* fam,
                        const double* pos_local, const double* rot_local, const double* dims,
                        const int32_t* material_id);

/* shape_container::triangle_rigid and ::convex_rigid (CMC/ChCollisionData.h:35-59) as one vertex pool: append n
 * points (3 doubles each) and get the index of the first in *first.  A triangle's three vertices are in the BODY
 * frame (GenerateAABB, CMC/ChCollisionSystemMulticore.cpp:524-535), a hull's points in the SHAPE frame (:424-456).
 * Add the vertices before the shapes that refer to them. */
int cb200_append_shape_vertices(cb200_system* sys, int n, const double* xyz, int* first);

/* Overwrite n points of the vertex pool from index first.  The reference's co-simulation terrain node writes the faces
 * of a deformable tire straight into shape_data.triangle_rigid at every exchange
 * (src/chrono_vehicle/cosim/terrain/ChVehicleCosimTerrainNodeGranularOMP.cpp:858-911, UpdateMeshProxy); this is that
 * write.  Addresses do not change, so a captured step graph stays valid. */
int cb200_update_shape_vertices(cb200_system* sys, int first, int n, const double* xyz);

/* Per-shape contact materials replacing the per-contact pointer chase of
 * ChContactContainerMulticoreSMC::AddContact (MC/collision/ChContactContainerMulticoreSMC.cpp:123-174).
 * 14 floats per material in ChContactMaterialSMC member order: young_modulus, poisson_ratio, static_friction,
 * sliding_friction, rolling_friction, spinning_friction, restitution, constant_adhesion, adhesionMultDMT,
 * adhesionSPerko, kn, kt, gn, gt. Composition is done on device in float, as the reference does
 * (src/chrono/physics/ChContactMaterialSMC.cpp:107-130). */
int cb200_set_materials_smc(cb200_system* sys, int n, const float* table);
/* NSC (MC/collision/ChContactContainerMulticoreNSC.cpp:130-180): 11 floats per material: static_friction,
 * sliding_friction, rolling_friction, spinning_friction, restitution, cohesion, dampingf, compliance, complianceT,
 * complianceRoll, complianceSpin. */
int cb200_set_materials_nsc(cb200_system* sys, int n, const float* table);

/* ChSystemMulticore::UpdateRigidBodies gather (MC/physics/ChSystemMulticore.cpp:329-377): upload pose and velocity of
 * bodies [first, first+n).  Any of pos/rot/vel may be NULL to leave it unchanged. */
int cb200_upload_state(cb200_system* sys, int first, int n, const double* pos, const double* rot, const double* vel);
/* Per-body applied force/torque other than gravity (accumulators / force list of ChBody::UpdateForces,
 * src/chrono/physics/ChBody.cpp:552-574): 6 doubles per body (F global, T local).  NULL clears. */
int cb200_upload_applied_forces(cb200_system* sys, int first, int n, const double* force_torque);
int cb200_set_body_flags(cb200_system* sys, int first, int n, const char* active, const char* collide);
/* host_data.active_rigid / collide_rigid as the device holds them now (either pointer may be NULL) */
int cb200_download_body_flags(cb200_system* sys, int first, int n, char* active, char* collide);
/* ChCollisionSystemMulticore::SetActiveBoundingBox (collision_settings.use_aabb_active,
 * src/chrono/collision/multicore/ChCollisionSystemMulticore.cpp:289-304 and GetOverlappingAABB :547-562): after the
 * AABB stage of every collision pass, an active colliding body none of whose shape boxes touches [min, max] is
 * deactivated for good (the reference puts it to sleep).  enabled = 0 switches the test off. */
int cb200_set_active_aabb(cb200_system* sys, int enabled, const double* aabb_min, const double* aabb_max);
/* ChBody::SetLimitSpeed / SetMaxLinVel / SetMaxAngVel (src/chrono/physics/ChBody.cpp:319-329, applied in
 * VariablesQbSetSpeed :268-286 to the new velocities, before the position increment) for bodies [first, first+n);
 * max_speed[i] <= 0: no limit on body i.  NULL arrays switch the clamp off for all bodies. */
int cb200_set_speed_limits(cb200_system* sys, int first, int n, const double* max_speed, const double* max_wvel);

/* ChSystem::DoStepDynamics for a caller that keeps the body state in host memory (what the reference does: gather from
 * the ChBody objects before, scatter to them after every step, MC/physics/ChSystemMulticore.cpp:329-377, :134-153):
 * uploads pos[3nb], rot[4nb], vel[6nb], runs one step, downloads the three arrays into the same buffers.  The
 * velocity upload overlaps the collision stages.  Buffers should come from cb200_host_alloc (pinned). */
int cb200_step_host(cb200_system* sys, double* pos, double* rot, double* vel);

/* --- the step, split at the reference's own seams ------------------------------------------------------------ */
/* ChSystemMulticore::Update -> UpdateRigidBodies force part (ChBody::Update/VariablesFbLoadForces): hf = h*(F,T-gyro). */
int cb200_update_forces(cb200_system* sys);
/* ChCollisionSystem::Run (CMC/ChCollisionSystemMulticore.cpp:286-322): GenerateAABB + ChBroadphase::Process +
 * ChNarrowphase::Process.  stages: bit0 AABB, bit1 broadphase, bit2 narrowphase (7 = all). */
int cb200_run_collision(cb200_system* sys, int stages);
/* ChIterativeSolverMulticoreSMC::RunTimeStep (MC/solver/ChIterativeSolverMulticoreSMC.cpp:792-852). */
int cb200_run_solver_smc(cb200_system* sys);
/* ChIterativeSolverMulticoreNSC::RunTimeStep (MC/solver/ChIterativeSolverMulticoreNSC.cpp:22-139). */
int cb200_run_solver_nsc(cb200_system* sys);
/* Scatter + integrate (MC/physics/ChSystemMulticore.cpp:134-153, ChBody::VariablesQbIncrementPosition). */
int cb200_integrate(cb200_system* sys);
/* ChSystem::DoStepDynamics x nsteps = the five calls above in sequence, state resident on device throughout. */
int cb200_do_steps(cb200_system* sys, int nsteps);
/* Block until all queued work of this system is done. */
int cb200_synchronize(cb200_system* sys);

/* --- read-backs (lazy downloads; each synchronizes) ---------------------------------------------------------- */
int cb200_get_counts(cb200_system* sys, cb200_counts* out);
int cb200_download_state(cb200_system* sys, int first, int n, double* pos, double* rot, double* vel);
/* cd_data->aabb_min/max after OffsetAABB (relative to global_origin, CMC/ChBroadphase.cpp:168-176) */
int cb200_download_aabb(cb200_system* sys, double* aabb_min, double* aabb_max);
/* bin_number/bin_aabb_number [N_i] sorted by (bin, shape); bin_active [N_a]; bin_start_index [N_a+1] */
int cb200_download_bins(cb200_system* sys, uint32_t* bin_number, uint32_t* bin_aabb_number, uint32_t* bin_active,
                        uint32_t* bin_start_index);
/* cd_data->pair_shapeIDs [N_p] */
int cb200_download_pairs(cb200_system* sys, int64_t* pair_shape_ids);
/* cd_data->{norm,cpta,cptb,dpth,erad,bids}_rigid_rigid and contact_shapeIDs, [N_c] */
int cb200_download_contacts(cb200_system* sys, double* normal, double* pt_a, double* pt_b, double* depth,
                            double* eff_radius, int32_t* body_ids, int64_t* shape_ids);
/* host_data.ct_force / ct_torque [2*N_c*3] in the reference's duplicated layout
 * (MC/solver/ChIterativeSolverMulticoreSMC.cpp:680-685) */
int cb200_download_contact_forces(cb200_system* sys, double* ct_force, double* ct_torque);
/* ChSystemMulticoreSMC::GetBodyContactForce/Torque for all bodies (MC/physics/ChSystemMulticoreSMC.cpp:74-90):
 * 3 doubles per body each, zero for bodies without contacts. */
/* Bilateral constraints (joints, motors) between bodies -- ChSystemMulticore::UpdateBilaterals +
 * ChConstraintBilateral::Build_D / Build_b, BODY_BODY rows (src/chrono_multicore/constraints/ChConstraintBilateral.cpp:
 * 34-78).  Row i couples bodies body_ids[2i], body_ids[2i+1] with the Jacobian blocks cq_a[6i..6i+5], cq_b[6i..6i+5]
 * of its ChConstraintTwoBodies (linear part in the absolute frame, angular part in the body frame) and right-hand side
 * rhs[i]: cq_a.v_a + cq_b.v_b + rhs = 0.  SMC systems solve the rows with MINRES on D_b^T M^-1 D_b after the contact
 * forces are known (ChIterativeSolverMulticoreSMC::RunTimeStep :816-842, ChSolverMulticoreMinRes::Solve;
 * settings.solver.max_iteration_bilateral, tol_speed) and add M^-1 D_b gamma_b to the velocities.  The rows stay in
 * force until replaced; n = 0 removes them.  Shaft rows and NSC systems: CB200_ERR_UNSUPPORTED. */
int cb200_set_bilaterals(cb200_system* sys, int n, const int32_t* body_ids, const double* cq_a, const double* cq_b,
                         const double* rhs, int max_iteration_bilateral, double tol_speed);
/* gamma_b of the last solve (impulses: ConstraintsFetch_react divides by the step), MINRES iterations, residual */
int cb200_download_bilateral_gamma(cb200_system* sys, double* gamma, int* iterations, double* residual);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The MINRES solve of the
 * bilateral stage (the kernel behind cb200_set_bilaterals) on caller-owned DEVICE buffers: N_dev n x n symmetric,
 * row-major; b_dev n; x_dev n (out, x0 = 0); info_dev 2 doubles (out: iterations run, residual).  Enqueued on the
 * system's stream; allocates nothing: the 7 n work doubles live in a static device array of the library, so
 * n <= cb200_minres_noalloc_max_rows() (2048; larger n: CB200_ERR_ARG).  Every call on a GPU shares that array: calls
 * that overlap on the GPU race, calls on one stream are safe; the part a call uses is cleared first. */
int cb200_minres_noalloc(cb200_system* sys, int n, const double* N_dev, const double* b_dev, double* x_dev,
                         int max_iter, double tol, double* info_dev);
int cb200_minres_noalloc_max_rows(void);
/* Contact callbacks.  The reference calls them on the host, one contact at a time, between the narrowphase and the
 * solver (ChCollisionSystemMulticore::ReportContacts, src/chrono/collision/multicore/ChCollisionSystemMulticore.cpp:
 * 326-372 -> ChContactContainerMulticoreSMC/NSC::AddContact).  Here the caller runs its loop over the downloaded list
 * (cb200_download_contacts) after cb200_run_collision and hands the outcome back before cb200_run_solver_*:
 *   - NarrowphaseCallback::OnNarrowphase == false  -> keep[i] = 0 in cb200_filter_contacts (list compacted in place);
 *   - AddContactCallback::OnAddContact(cinfo, composite material) -> cb200_download_contact_materials gives the
 *     composite records the container formed (SMC 13 doubles per contact: mu, muRoll, muSpin, E, G, adhesion, DMT,
 *     Perko, cr, kn, kt, gn, gt; NSC 8: mu, muRoll, muSpin, cohesion, compliance, complianceT,
 *     complianceRoll, complianceSpin), cb200_upload_contact_materials the modified ones; they hold for the current list. */
int cb200_filter_contacts(cb200_system* sys, int n, const uint8_t* keep);
int cb200_download_contact_materials(cb200_system* sys, int capacity, double* composite, int* stride);
int cb200_upload_contact_materials(cb200_system* sys, int n, const double* composite);
/* ChContactContainerMulticore::ReportAllContacts (src/chrono_multicore/collision/ChContactContainerMulticore.cpp:65-149),
 * batched: row i holds the arguments the reference passes to ReportContactCallback::OnReportContact for contact i --
 * pA, pB (absolute), plane (3x3 row-major, X axis = contact normal, SetFromAxisX(normal, VECT_Y)), distance,
 * eff_radius, react_forces and react_torques in the contact frame (NSC: multipliers / step, SMC: force and torque on
 * body B, the torque moved to the contact point), the two body indices.  Any output pointer may be NULL; capacity in
 * contacts; *count = num_rigid_contacts.  The caller runs its callback loop over the batch. */
int cb200_report_contacts(cb200_system* sys, int capacity, double* pA, double* pB, double* plane, double* distance,
                          double* eff_radius, double* force, double* torque, int32_t* body_ids, int* count);
int cb200_download_body_contact_forces(cb200_system* sys, double* force, double* torque);
/* host_data.hf (h * generalized force, 6 per body) as consumed by ComputeInvMassMatrix */
int cb200_download_hf(cb200_system* sys, double* hf);
/* NSC multipliers host_data.gamma in the reference layout [n | u v | tn tu tv] (MC/ChDataManager.h:307-317);
 * rows per contact = 1 / 3 / 6 for solver_mode NORMAL / SLIDING / SPINNING */
int cb200_download_gamma(cb200_system* sys, double* gamma, int capacity);
/* host_data.R_full = -b - D_T (v + M_inv hf) of the last NSC solve (MC/solver/ChIterativeSolverMulticoreNSC.cpp:80-82),
 * reference layout */
int cb200_download_nsc_rhs(cb200_system* sys, double* r_full);
/* measures.solver.maxd_hist / objective per APGD iteration of the last solve (AtIterationEnd,
 * MC/solver/ChSolverMulticoreAPGD.cpp:182); *count = iterations recorded */
int cb200_download_nsc_history(cb200_system* sys, double* residual, double* objective, int capacity, int* count);
/* Rows of host_data.D_T for the rigid contacts (ChConstraintRigidRigid::Build_D,
 * MC/constraints/ChConstraintRigidRigid.cpp:377-452), reference row order: 12 values per row = the 6 entries at the
 * columns of body A (bids.x) then the 6 of body B; body_ids (may be NULL) gets the two body ids per row.  The solver
 * itself is matrix-free; this is the read-back the reference's Jacobian unit test (utest_MCORE_jacobians.cpp) needs. */
int cb200_download_nsc_jacobian(cb200_system* sys, double* rows12, int32_t* body_ids);

/* --- functor seams of the inner NSC solver (ChSolverMulticore::Solve arguments, MC/solver/ChSolverMulticore.h:107-113)
 * on the system assembled by the last NSC solve; host vectors in the reference layout.  local_mode = the reference's
 * settings.solver.local_solver_mode (0 NORMAL, 1 SLIDING, 2 SPINNING), not above solver_mode. */
/* ChSchurProduct::operator()(x, out): out = D_T M_inv D x + E x (MC/solver/ChSchurProduct.cpp:58-173) */
int cb200_nsc_schur_product(cb200_system* sys, int local_mode, const double* x, double* out);
/* ChProjectConstraints::operator()(gamma) (MC/solver/ChSolverMulticore.cpp:25-30 ->
 * ChConstraintRigidRigid::Project), in place */
int cb200_nsc_project(cb200_system* sys, int local_mode, double* gamma);

/* Timers with the reference's names (MC/ChTimerMulticore.h): seconds of GPU time measured with CUDA
 * events over the last cb200_do_steps call, averaged per step.  Unknown name -> CB200_ERR_ARG. */
int cb200_get_timer(cb200_system* sys, const char* name, double* seconds);
/* Number of kernels this library launched since creation (for bench.py's gpu_launches). */
int64_t cb200_kernel_launches(cb200_system* sys);
/* Self-test of the device scan (the replacement of the thrust::exclusive_scan / inclusive_scan calls of
 * src/chrono/collision/multicore/ChBroadphase.cpp:271-301 and ChNarrowphase.cpp:141-150): n pseudo-random counts,
 * `repeats` scans back to back, every output compared with a host scan; *mismatches = entries that differ. */
int cb200_selftest_scan(cb200_system* sys, int64_t n, uint32_t seed, int repeats, int64_t* mismatches);

/* --- device-pointer access for zero-copy tooling and multi-GPU halo exchange ---------------------- */
/* names: "pos","rot","vel","hf","aabb_min","aabb_max","contact_normal",... ; returns device pointer + element count */
int cb200_device_array(cb200_system* sys, const char* name, void** dev_ptr, int64_t* count);
/* --- asynchronous stepping (callers that interleave their own work on the stream, e.g. the slab exchange) ----- */
/* Use the caller's CUDA stream (a cudaStream_t, e.g. torch.cuda.current_stream().cuda_stream; 0 = the legacy default
 * stream) for everything the library enqueues; use_own != 0 returns to the stream cb200_create made. */
int cb200_set_stream(cb200_system* sys, void* cuda_stream, int use_own);
/* cb200_do_steps without the final synchronisation and without capacity recovery. */
int cb200_enqueue_steps(cb200_system* sys, int nsteps);
/* Synchronise, resolve timers, and fail if a capacity overflowed since the last check (the steps after an overflow
 * did not run). */
int cb200_check(cb200_system* sys);
int cb200_reset_timers(cb200_system* sys);

/* --- slab decomposition across GPUs (no reference counterpart: Chrono::Multicore is single-node) ---- */
/* Global body ids (index in the undecomposed system) of bodies [first, first+n). */
int cb200_upload_body_ids(cb200_system* sys, int first, int n, const int64_t* ids);
/* ids and flags (bit0 owned/active, bit1 collide, bit2 vacant slot) of bodies [first, first+n); either may be NULL */
int cb200_download_body_ids(cb200_system* sys, int first, int n, int64_t* ids, uint8_t* flags);
/* Make this system one slab: bodies [nb_static, nb) are its owned bodies (one shape each, in order); appends vacant
 * owned slots up to owned_cap and two ghost regions of ghost_cap slots.  [lo, hi) is the owned interval along axis,
 * width the ghost-layer width (>= 2 r_max + 2 envelope). */
int cb200_slab_configure(cb200_system* sys, int nb_static, int ns_static, int axis, double lo, double hi, double width,
                         int has_left, int has_right, int owned_cap, int ghost_cap);
/* doubles in one exchange buffer: 2 (count, reserved) + 33 per record * ghost_cap */
int64_t cb200_slab_buffer_doubles(cb200_system* sys);
/* After a step: write the bodies within `width` of (or past) each cut plane into the neighbour's DEVICE send buffer
 * and release the slots of bodies that left [lo, hi).  Asynchronous on the system's stream. */
int cb200_slab_pack_dev(cb200_system* sys, void* send_left_dev, void* send_right_dev);
/* Rebuild the ghost regions from the two received DEVICE buffers; records inside [lo, hi) are adopted as owned. */
int cb200_slab_unpack_dev(cb200_system* sys, const void* recv_left_dev, const void* recv_right_dev);

/* --- the data plane of a decomposed system, in the library (csrc/comm.cu): ghost exchange every step, and for NSC the
 * halo of the body-space vector inside every Schur product plus one all-reduce of the APGD scalars per iteration
 * (ChSolverMulticoreAPGD.cpp:145,147,160,173).  Slab i is rank i; neighbours are rank +- 1. ---------------------- */
/* ncclGetUniqueId from the NCCL library at nccl_path (NULL or "": libnccl.so.2); 128 bytes, rank 0 shares them */
int cb200_comm_unique_id(const char* nccl_path, char* out128);
/* One slab per process: join an NCCL communicator (the library is resolved with dlopen, so the process keeps one NCCL) */
int cb200_comm_init_nccl(cb200_system* sys, const char* nccl_path, int rank, int world, const char* id128);
/* All slabs in one process, in rank order, on one stream (tests, single-GPU emulation): transfers are device copies */
int cb200_comm_init_local(cb200_system** systems, int n);
/* Ghost-layer exchange of the listed slabs (n = 1 over NCCL).  NSC systems exchange twice and fetch the record counts. */
int cb200_comm_exchange(cb200_system** systems, int n);
/* nsteps x (step of every listed slab + ghost exchange), enqueued by the library.  SMC: asynchronous (cb200_check). */
int cb200_comm_steps(cb200_system** systems, int n, int nsteps);

#ifdef __cplusplus
}
#endif
#endif /* CHRONO_B200_H */
