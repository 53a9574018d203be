// See ChCollisionSystemB200.h.  Every method cites the reference code whose role it takes.
#include "ChCollisionSystemB200.h"

#include <cstring>

#include "chrono/collision/ChCollisionModel.h"

namespace chrono {

ChCollisionSystemB200::ChCollisionSystemB200(std::shared_ptr<ChCollisionData> data, int device) : cd_data(data) {
    Check(cb200_create(device, &sys));
    cb200_default_settings(&settings);
    settings.system_type = 0;  // the envelope is the caller's (ChSystemMulticoreSMC sets it to zero itself)
}

ChCollisionSystemB200::~ChCollisionSystemB200() {
    cb200_destroy(sys);
}

void ChCollisionSystemB200::SetBroadphaseGridResolution(const vec3& bins) {
    settings.broadphase_grid = 0;
    settings.bins_per_axis[0] = bins.x, settings.bins_per_axis[1] = bins.y, settings.bins_per_axis[2] = bins.z;
}
void ChCollisionSystemB200::SetBroadphaseGridSize(const real3& bin_size) {
    settings.broadphase_grid = 1;
    settings.bin_size[0] = bin_size.x, settings.bin_size[1] = bin_size.y, settings.bin_size[2] = bin_size.z;
}
void ChCollisionSystemB200::SetBroadphaseGridDensity(real density) {
    settings.broadphase_grid = 2;
    settings.grid_density = density;
}
void ChCollisionSystemB200::SetNarrowphaseAlgorithm(int algorithm) {
    settings.narrowphase_algorithm = algorithm;
}

// collision_settings of the data manager -> the backend (ChCollisionSystemChronoMulticore::PreProcess,
// src/chrono_multicore/collision/ChCollisionSystemChronoMulticore.cpp:44-78 copies the same fields into cd_data)
void ChCollisionSystemB200::PushSettings() {
    settings.collision_envelope = cd_data->collision_envelope;
    Check(cb200_set_settings(sys, &settings));
}

// Shapes added since the last call (ChCollisionSystemMulticore::Add, ChCollisionSystemMulticore.cpp:86-187, appends to
// shape_data; nothing is ever removed: Remove throws).  Bodies follow the state arrays.
void ChCollisionSystemB200::PushNewBodiesAndShapes() {
    const auto& st = cd_data->state_data;
    const size_t nb = st.pos_rigid->size();
    if (nb > bodies_on_device) {
        const size_t n = nb - bodies_on_device;
        std::vector<double> mass(n, 1.0), inertia(3 * n, 1.0);  // not used by the collision stages
        std::vector<char> act(n), col(n);
        for (size_t i = 0; i < n; i++) {
            act[i] = (*st.active_rigid)[bodies_on_device + i];
            col[i] = (*st.collide_rigid)[bodies_on_device + i];
        }
        Check(cb200_append_bodies(sys, (int)n, mass.data(), inertia.data(), nullptr, act.data(), col.data(), nullptr));
        bodies_on_device = nb;
    }
    const auto& sh = cd_data->shape_data;
    const size_t ns = sh.typ_rigid.size();
    if (ns == shapes_on_device)
        return;
    const size_t n = ns - shapes_on_device;
    std::vector<int32_t> type(n);
    std::vector<uint32_t> body(n);
    std::vector<int16_t> fam(2 * n);
    std::vector<double> lpos(3 * n), lrot(4 * n), dims(4 * n, 0.0);
    for (size_t k = 0; k < n; k++) {
        const size_t i = shapes_on_device + k;
        const int start = sh.start_rigid[i];
        type[k] = sh.typ_rigid[i];
        body[k] = sh.id_rigid[i];
        fam[2 * k] = sh.fam_rigid[i].x, fam[2 * k + 1] = sh.fam_rigid[i].y;
        lpos[3 * k] = sh.ObA_rigid[i].x, lpos[3 * k + 1] = sh.ObA_rigid[i].y, lpos[3 * k + 2] = sh.ObA_rigid[i].z;
        lrot[4 * k] = sh.ObR_rigid[i].w, lrot[4 * k + 1] = sh.ObR_rigid[i].x, lrot[4 * k + 2] = sh.ObR_rigid[i].y;
        lrot[4 * k + 3] = sh.ObR_rigid[i].z;
        double* d = &dims[4 * k];
        switch (sh.typ_rigid[i]) {  // where ChConvexShape reads the dimensions from (ChConvexShape.h:70-160)
            case ChCollisionShape::SPHERE:
                d[0] = sh.sphere_rigid[start];
                break;
            case ChCollisionShape::ELLIPSOID:
            case ChCollisionShape::BOX:
            case ChCollisionShape::CYLINDER:
            case ChCollisionShape::CYLSHELL:
            case ChCollisionShape::CONE:
                d[0] = sh.box_like_rigid[start].x, d[1] = sh.box_like_rigid[start].y, d[2] = sh.box_like_rigid[start].z;
                break;
            case ChCollisionShape::CAPSULE:
                d[0] = sh.capsule_rigid[start].x, d[1] = sh.capsule_rigid[start].y;
                break;
            case ChCollisionShape::ROUNDEDBOX:
            case ChCollisionShape::ROUNDEDCYL:
                d[0] = sh.rbox_like_rigid[start].x, d[1] = sh.rbox_like_rigid[start].y;
                d[2] = sh.rbox_like_rigid[start].z, d[3] = sh.rbox_like_rigid[start].w;
                break;
            case ChCollisionShape::TRIANGLE:
            case ChCollisionShape::CONVEXHULL: {
                const bool tri = sh.typ_rigid[i] == ChCollisionShape::TRIANGLE;
                const int count = tri ? 3 : sh.length_rigid[i];
                const std::vector<real3>& pool = tri ? sh.triangle_rigid : sh.convex_rigid;
                std::vector<double> xyz(3 * (size_t)count);
                for (int v = 0; v < count; v++)
                    xyz[3 * v] = pool[start + v].x, xyz[3 * v + 1] = pool[start + v].y, xyz[3 * v + 2] = pool[start + v].z;
                int first = 0;
                Check(cb200_append_shape_vertices(sys, count, xyz.data(), &first));
                d[0] = first, d[1] = count;
                break;
            }
            default:
                throw std::runtime_error("ChCollisionSystemB200: shape type without a device counterpart");
        }
    }
    Check(cb200_append_shapes(sys, (int)n, type.data(), body.data(), fam.data(), lpos.data(), lrot.data(), dims.data(),
                              nullptr));
    shapes_on_device = ns;
}

// real3 is padded to 4 doubles (src/chrono/multicore_math/real3.h:85-90): strip the pad on the way to the device
void ChCollisionSystemB200::UploadState() {
    const auto& st = cd_data->state_data;
    const size_t nb = st.pos_rigid->size();
    pos3.resize(3 * nb), rot4.resize(4 * nb);
    for (size_t i = 0; i < nb; i++) {
        const real3& p = (*st.pos_rigid)[i];
        const quaternion& q = (*st.rot_rigid)[i];
        pos3[3 * i] = p.x, pos3[3 * i + 1] = p.y, pos3[3 * i + 2] = p.z;
        rot4[4 * i] = q.w, rot4[4 * i + 1] = q.x, rot4[4 * i + 2] = q.y, rot4[4 * i + 3] = q.z;
    }
    Check(cb200_upload_state(sys, 0, (int)nb, pos3.data(), rot4.data(), nullptr));
}

void ChCollisionSystemB200::SetActiveBoundingBox(const real3& aabbmin, const real3& aabbmax) {
    const double lo[3] = {aabbmin.x, aabbmin.y, aabbmin.z}, hi[3] = {aabbmax.x, aabbmax.y, aabbmax.z};
    Check(cb200_set_active_aabb(sys, 1, lo, hi));
    use_aabb_active = true;
}

void ChCollisionSystemB200::PreProcess() {
    PushSettings();
    PushNewBodiesAndShapes();
    UploadState();
}

void ChCollisionSystemB200::Run() {
    Check(cb200_run_collision(sys, 7));
}

// ChCollisionSystemMulticore::Run's bookkeeping (:324-345) and what the measures read from cd_data
void ChCollisionSystemB200::PostProcess() {
    cb200_counts c;
    Check(cb200_get_counts(sys, &c));
    cd_data->num_rigid_shapes = c.num_shapes;
    cd_data->num_rigid_contacts = c.num_rigid_contacts;
    cd_data->num_bins = c.num_bins;
    cd_data->num_bin_aabb_intersections = c.num_bin_aabb_intersections;
    cd_data->num_active_bins = c.num_active_bins;
    cd_data->num_possible_collisions = c.num_possible_collisions;
    cd_data->bins_per_axis = vec3((int)c.bins_per_axis[0], (int)c.bins_per_axis[1], (int)c.bins_per_axis[2]);
    cd_data->min_bounding_point = real3(c.min_bounding_point[0], c.min_bounding_point[1], c.min_bounding_point[2]);
    cd_data->max_bounding_point = real3(c.max_bounding_point[0], c.max_bounding_point[1], c.max_bounding_point[2]);
    cd_data->global_origin = cd_data->min_bounding_point;
    cd_data->bin_size = real3(c.bin_size[0], c.bin_size[1], c.bin_size[2]);
    cd_data->inv_bin_size = real3(1.0 / c.bin_size[0], 1.0 / c.bin_size[1], 1.0 / c.bin_size[2]);
    if (use_aabb_active) {  // ChCollisionSystemMulticore.cpp:297-303 writes the flags of the data manager
        auto& act = *cd_data->state_data.active_rigid;
        Check(cb200_download_body_flags(sys, 0, (int)act.size(), act.data(), nullptr));
    }
}

void ChCollisionSystemB200::GetBoundingBox(real3& aabb_min, real3& aabb_max) const {
    aabb_min = cd_data->min_bounding_point;
    aabb_max = cd_data->max_bounding_point;
}

double ChCollisionSystemB200::GetTimerCollisionBroad() const {
    double s = 0;
    cb200_get_timer(sys, "collision_broad", &s);
    return s;
}
double ChCollisionSystemB200::GetTimerCollisionNarrow() const {
    double s = 0;
    cb200_get_timer(sys, "collision_narrow", &s);
    return s;
}
void ChCollisionSystemB200::ResetTimers() {
    cb200_reset_timers(sys);
}

void ChCollisionSystemB200::FetchContacts() {
    const size_t nc = cd_data->num_rigid_contacts;
    std::vector<double> n(3 * nc), pa(3 * nc), pb(3 * nc), depth(nc), erad(nc);
    std::vector<int32_t> bids(2 * nc);
    cd_data->contact_shapeIDs.resize(nc);
    Check(cb200_download_contacts(sys, n.data(), pa.data(), pb.data(), depth.data(), erad.data(), bids.data(),
                                  reinterpret_cast<int64_t*>(cd_data->contact_shapeIDs.data())));
    cd_data->norm_rigid_rigid.resize(nc), cd_data->cpta_rigid_rigid.resize(nc), cd_data->cptb_rigid_rigid.resize(nc);
    cd_data->dpth_rigid_rigid.resize(nc), cd_data->erad_rigid_rigid.resize(nc), cd_data->bids_rigid_rigid.resize(nc);
    for (size_t i = 0; i < nc; i++) {  // re-pad to real3
        cd_data->norm_rigid_rigid[i] = real3(n[3 * i], n[3 * i + 1], n[3 * i + 2]);
        cd_data->cpta_rigid_rigid[i] = real3(pa[3 * i], pa[3 * i + 1], pa[3 * i + 2]);
        cd_data->cptb_rigid_rigid[i] = real3(pb[3 * i], pb[3 * i + 1], pb[3 * i + 2]);
        cd_data->dpth_rigid_rigid[i] = depth[i];
        cd_data->erad_rigid_rigid[i] = erad[i];
        cd_data->bids_rigid_rigid[i] = I2(bids[2 * i], bids[2 * i + 1]);
    }
}

void ChCollisionSystemB200::FetchBroadphase() {
    const size_t ns = cd_data->num_rigid_shapes, ni = cd_data->num_bin_aabb_intersections, na = cd_data->num_active_bins;
    std::vector<double> mn(3 * ns), mx(3 * ns);
    Check(cb200_download_aabb(sys, mn.data(), mx.data()));
    cd_data->aabb_min.resize(ns), cd_data->aabb_max.resize(ns);
    for (size_t i = 0; i < ns; i++) {
        cd_data->aabb_min[i] = real3(mn[3 * i], mn[3 * i + 1], mn[3 * i + 2]);
        cd_data->aabb_max[i] = real3(mx[3 * i], mx[3 * i + 1], mx[3 * i + 2]);
    }
    cd_data->bin_number.resize(ni), cd_data->bin_aabb_number.resize(ni);
    cd_data->bin_active.resize(na), cd_data->bin_start_index.resize(na + 1);
    Check(cb200_download_bins(sys, cd_data->bin_number.data(), cd_data->bin_aabb_number.data(), cd_data->bin_active.data(),
                              cd_data->bin_start_index.data()));
    cd_data->pair_shapeIDs.resize(cd_data->num_possible_collisions);
    Check(cb200_download_pairs(sys, reinterpret_cast<int64_t*>(cd_data->pair_shapeIDs.data())));
}

}  // namespace chrono
