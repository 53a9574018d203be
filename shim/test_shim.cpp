// Drives a 1.5k-shape bed through ChCollisionSystemB200 (the seam a Chrono build would use) and compares what lands in
// the reference's own ChCollisionData with the reference's own collision code (oracle/_ref) on the same state:
// counters, candidate pair list and contact arrays, bit for bit and in order.  Needs a CUDA device.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <vector>

#include "ChCollisionSystemB200.h"
#include "chrono/collision/ChCollisionModel.h"

extern "C" {  // oracle/ref_driver.cpp
void* ref_create();
void ref_destroy(void*);
void ref_set_bodies(void*, int, const double*, const double*, const char*, const char*);
void ref_set_shapes(void*, int, const int*, const unsigned*, const short*, const double*, const double*, const double*,
                    const double*, int, const double*);
void ref_set_collision_settings(void*, double, int, const int*, const double*, double, int);
void ref_run_collision(void*, int);
void ref_get_counts(void*, unsigned*);
void ref_get_pairs(void*, long long*);
void ref_get_contacts(void*, double*, double*, double*, double*, double*, int*, long long*);
}

using namespace chrono;

static double lcg(uint64_t& s) {
    s = s * 6364136223846793005ull + 1442695040888963407ull;
    return (double)(s >> 11) / 9007199254740992.0;
}

int main() {
    const int nx = 14, ny = 12, nz = 8;
    const double r = 0.02, spacing = 1.98 * r;
    // ---- flat scene: body 0 = container (5 boxes), then one sphere per body (a few boxes and capsules mixed in)
    std::vector<double> pos, rot, lpos, lrot, dims;
    std::vector<char> act, col;
    std::vector<int> type;
    std::vector<unsigned> body;
    std::vector<short> fam;
    auto add_body = [&](double x, double y, double z, const double* q, char a) {
        pos.insert(pos.end(), {x, y, z});
        rot.insert(rot.end(), {q[0], q[1], q[2], q[3]});
        act.push_back(a), col.push_back(1);
        return (unsigned)(act.size() - 1);
    };
    auto add_shape = [&](int t, unsigned b, double px, double py, double pz, double d0, double d1, double d2, double d3) {
        type.push_back(t), body.push_back(b);
        fam.push_back(1), fam.push_back(0x7fff);
        lpos.insert(lpos.end(), {px, py, pz});
        lrot.insert(lrot.end(), {1.0, 0.0, 0.0, 0.0});
        dims.insert(dims.end(), {d0, d1, d2, d3});
    };
    const double qid[4] = {1, 0, 0, 0};
    const unsigned ground = add_body(0, 0, 0, qid, 0);
    const double sx = nx * spacing + r, sy = ny * spacing + r, sz = 0.5, ht = 0.1;
    add_shape(ChCollisionShape::BOX, ground, 0, 0, -ht, sx / 2, sy / 2, ht, 0);
    add_shape(ChCollisionShape::BOX, ground, -sx / 2 - ht, 0, sz / 2, ht, sy / 2, sz / 2, 0);
    add_shape(ChCollisionShape::BOX, ground, sx / 2 + ht, 0, sz / 2, ht, sy / 2, sz / 2, 0);
    add_shape(ChCollisionShape::BOX, ground, 0, -sy / 2 - ht, sz / 2, sx / 2, ht, sz / 2, 0);
    add_shape(ChCollisionShape::BOX, ground, 0, sy / 2 + ht, sz / 2, sx / 2, ht, sz / 2, 0);
    uint64_t seed = 12345;
    for (int k = 0; k < nz; k++)
        for (int j = 0; j < ny; j++)
            for (int i = 0; i < nx; i++) {
                double q[4] = {lcg(seed) - 0.5, lcg(seed) - 0.5, lcg(seed) - 0.5, lcg(seed) - 0.5};
                const double qn = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
                for (double& c : q)
                    c /= qn;
                const unsigned b = add_body((i - (nx - 1) / 2.0) * spacing + 0.03 * r * (lcg(seed) - 0.5),
                                            (j - (ny - 1) / 2.0) * spacing + 0.03 * r * (lcg(seed) - 0.5),
                                            0.99 * r + k * spacing + 0.03 * r * (lcg(seed) - 0.5), q, 1);
                const int kind = (i + 3 * j + 5 * k) % 11;
                if (kind == 0)
                    add_shape(ChCollisionShape::BOX, b, 0, 0, 0, 0.9 * r, 0.8 * r, 0.7 * r, 0);
                else if (kind == 1)
                    add_shape(ChCollisionShape::CAPSULE, b, 0, 0, 0, 0.7 * r, 0.7 * r, 0, 0);
                else
                    add_shape(ChCollisionShape::SPHERE, b, 0, 0, 0, r, 0, 0, 0);
            }
    const int nb = (int)act.size(), ns = (int)type.size();

    // ---- the same scene in the reference's data contract, as ChCollisionSystemMulticore::Add leaves it (:86-187)
    auto cd = std::make_shared<ChCollisionData>(true);
    for (int b = 0; b < nb; b++) {
        cd->state_data.pos_rigid->push_back(real3(pos[3 * b], pos[3 * b + 1], pos[3 * b + 2]));
        cd->state_data.rot_rigid->push_back(quaternion(rot[4 * b], rot[4 * b + 1], rot[4 * b + 2], rot[4 * b + 3]));
        cd->state_data.active_rigid->push_back(act[b]);
        cd->state_data.collide_rigid->push_back(col[b]);
    }
    auto& sh = cd->shape_data;
    for (int i = 0; i < ns; i++) {
        const double* d = &dims[4 * i];
        int start = -1;
        if (type[i] == ChCollisionShape::SPHERE) {
            start = (int)sh.sphere_rigid.size();
            sh.sphere_rigid.push_back(d[0]);
        } else if (type[i] == ChCollisionShape::BOX) {
            start = (int)sh.box_like_rigid.size();
            sh.box_like_rigid.push_back(real3(d[0], d[1], d[2]));
        } else {
            start = (int)sh.capsule_rigid.size();
            sh.capsule_rigid.push_back(real2(d[0], d[1]));
        }
        sh.ObA_rigid.push_back(real3(lpos[3 * i], lpos[3 * i + 1], lpos[3 * i + 2]));
        sh.ObR_rigid.push_back(quaternion(1, 0, 0, 0));
        sh.start_rigid.push_back(start), sh.length_rigid.push_back(1);
        sh.fam_rigid.push_back(S2(1, 0x7fff));
        sh.typ_rigid.push_back(type[i]), sh.id_rigid.push_back(body[i]), sh.local_rigid.push_back(0);
    }
    cd->num_rigid_shapes = ns;
    cd->collision_envelope = 0.001;

    // ---- through the seam
    ChCollisionSystemB200 cs(cd, 0);
    cs.SetBroadphaseGridDensity(5.0);
    cs.SetNarrowphaseAlgorithm(2);  // HYBRID
    cs.PreProcess();
    cs.Run();
    cs.PostProcess();
    cs.FetchBroadphase();
    cs.FetchContacts();

    // ---- the reference's own code on the same arrays
    void* ref = ref_create();
    ref_set_shapes(ref, ns, type.data(), body.data(), fam.data(), lpos.data(), lrot.data(), dims.data(), nullptr, 0, nullptr);
    ref_set_bodies(ref, nb, pos.data(), rot.data(), act.data(), col.data());
    const int res[3] = {10, 10, 10};
    const double bs[3] = {1, 1, 1};
    ref_set_collision_settings(ref, 0.001, 2, res, bs, 5.0, 2);
    ref_run_collision(ref, 7);
    unsigned c[8];
    ref_get_counts(ref, c);
    int bad = 0;
    auto expect = [&](bool ok, const char* what) {
        if (!ok) {
            std::printf("MISMATCH: %s\n", what);
            bad++;
        }
    };
    expect(cd->num_bin_aabb_intersections == c[0] && cd->num_active_bins == c[1], "bin counters");
    expect(cd->num_possible_collisions == c[2], "number of candidate pairs");
    expect(cd->num_rigid_contacts == c[3], "number of contacts");
    expect((unsigned)cd->bins_per_axis.x == c[4] && (unsigned)cd->bins_per_axis.y == c[5] && (unsigned)cd->bins_per_axis.z == c[6],
           "grid resolution");
    const size_t np = c[2], nc = c[3];
    std::vector<long long> pairs(np), ids(nc);
    std::vector<double> n(3 * nc), pa(3 * nc), pb(3 * nc), dp(nc), er(nc);
    std::vector<int> bids(2 * nc);
    ref_get_pairs(ref, pairs.data());
    ref_get_contacts(ref, n.data(), pa.data(), pb.data(), dp.data(), er.data(), bids.data(), ids.data());
    if (!bad) {
        expect(std::memcmp(pairs.data(), cd->pair_shapeIDs.data(), np * sizeof(long long)) == 0, "pair list");
        expect(std::memcmp(ids.data(), cd->contact_shapeIDs.data(), nc * sizeof(long long)) == 0, "contact shape ids");
        for (size_t i = 0; i < nc && !bad; i++) {
            const real3 N = cd->norm_rigid_rigid[i], A = cd->cpta_rigid_rigid[i], B = cd->cptb_rigid_rigid[i];
            expect(N.x == n[3 * i] && N.y == n[3 * i + 1] && N.z == n[3 * i + 2], "contact normal");
            expect(A.x == pa[3 * i] && A.y == pa[3 * i + 1] && A.z == pa[3 * i + 2], "contact point A");
            expect(B.x == pb[3 * i] && B.y == pb[3 * i + 1] && B.z == pb[3 * i + 2], "contact point B");
            expect(cd->dpth_rigid_rigid[i] == dp[i] && cd->erad_rigid_rigid[i] == er[i], "depth / effective radius");
            expect(cd->bids_rigid_rigid[i].x == bids[2 * i] && cd->bids_rigid_rigid[i].y == bids[2 * i + 1], "body ids");
        }
    }
    ref_destroy(ref);
    std::printf("%s: %d bodies, %d shapes, N_i=%u N_p=%zu N_c=%zu, broad %.3f ms narrow %.3f ms\n", bad ? "FAIL" : "PASS", nb, ns,
                c[0], np, nc, cs.GetTimerCollisionBroad() * 1e3, cs.GetTimerCollisionNarrow() * 1e3);
    return bad ? 1 : 0;
}
