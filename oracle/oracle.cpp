// oracle.cpp -- CPU restatement of the Chrono::Multicore granular-contact hot path.  TEST INFRASTRUCTURE.
//
// Plain scalar C++ (compile with -ffp-contract=off), own containers, std::stable_sort where the reference uses
// Thrust.  Each function names the reference lines it restates (paths relative to the reference root;
// CMC = src/chrono/collision/multicore, MC = src/chrono_multicore, MM = src/chrono/multicore_math).
// Pinned by tests/test_oracle_cpu.py against the reference's own code (oracle/_ref) and the golden vectors in
// tests/golden/.  Nothing in chrono_wisc_b200/ includes or links this file.
#include "oracle.h"

#include <algorithm>
#include <cfloat>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstring>
#include <vector>
#include <omp.h>

namespace {

struct V3 {
    double x, y, z;
};
struct Q {
    double w, x, y, z;
};
inline V3 v3(double x, double y, double z) { return V3{x, y, z}; }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, double s) { return v3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(double s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
inline V3 operator/(V3 a, double s) { return v3(a.x / s, a.y / s, a.z / s); }
inline V3 operator+(V3 a, double s) { return v3(a.x + s, a.y + s, a.z + s); }
inline V3 operator-(V3 a, double s) { return v3(a.x - s, a.y - s, a.z - s); }
// MM/simd_non.h:57-75
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline double len(V3 a) { return std::sqrt(dot(a, a)); }
inline Q conj(Q q) { return Q{q.w, -q.x, -q.y, -q.z}; }
// MM/real4.cpp:139-150
inline Q qmul(Q a, Q b) {
    return Q{a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w - a.z * b.y + a.y * b.z,
             a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z, a.w * b.z + a.z * b.w - a.y * b.x + a.x * b.y};
}
// MM/real4.cpp:158-165
inline V3 rot(V3 v, Q q) {
    V3 u = v3(q.x, q.y, q.z);
    V3 t = 2.0 * cross(u, v);
    return v + q.w * t + cross(u, t);
}
inline V3 rotT(V3 v, Q q) { return rot(v, conj(q)); }
// MM/real4.cpp:168-188
inline V3 absrot(Q q, V3 v) {
    double ww = q.w * q.w, xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
    double wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z, xy = q.x * q.y, xz = q.x * q.z, yz = q.y * q.z;
    return v3(std::abs((ww + xx) * 2 - 1) * v.x + std::abs((xy - wz) * 2) * v.y + std::abs((xz + wy) * 2) * v.z,
              std::abs((xy + wz) * 2) * v.x + std::abs((ww + yy) * 2 - 1) * v.y + std::abs((yz - wx) * 2) * v.z,
              std::abs((xz - wy) * 2) * v.x + std::abs((yz + wx) * 2) * v.y + std::abs((ww + zz) * 2 - 1) * v.z);
}
inline V3 toParent(V3 p, Q q, V3 l) { return p + rot(l, q); }    // MM/utility.h:48-50
inline V3 toLocal(V3 p, Q q, V3 g) { return rotT(g - p, q); }    // MM/utility.h:55-57
inline V3 colW(Q q) {                                            // AMatW, MM/real4.cpp:235-250
    return v3((q.x * q.z + q.w * q.y) * 2, (q.y * q.z - q.w * q.x) * 2, (q.w * q.w + q.z * q.z) * 2 - 1);
}

enum { T_SPHERE = 0, T_ELLIPSOID = 1, T_BOX = 2, T_CYLINDER = 3, T_CYLSHELL = 4, T_CAPSULE = 12, T_CONE = 13,
       T_RBOX = 14, T_RCYL = 15 };

struct Contact {
    V3 n, pa, pb;
    double depth, erad;
    int b1, b2;
    int64_t key;
};

double now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

}  // namespace

struct orc_nsc;
struct orc_system {
    cb200_settings st;
    orc_nsc* nsc = nullptr;  // NSC solver state (oracle_nsc.inc)
    int nb = 0, ns = 0, nm = 0;
    std::vector<double> mass, inertia;
    std::vector<char> active, collide;
    std::vector<V3> pos;
    std::vector<Q> rotq;
    std::vector<double> vel, hf;
    std::vector<int> typ, mat;
    std::vector<unsigned> sbody;
    std::vector<short> fam;
    std::vector<V3> lpos, wpos;
    std::vector<Q> lrot, wrot;
    std::vector<double> dims;
    std::vector<float> mats;
    // collision outputs
    std::vector<V3> amin, amax;
    V3 origin, maxpt, bsize, inv;
    int bins[3] = {1, 1, 1};
    unsigned N_i = 0, N_a = 0;
    std::vector<int64_t> pairs;
    std::vector<Contact> contacts;
    std::vector<V3> cforce, ctq1, ctq2;  // per contact: force on body 2, torques on bodies 1 and 2
    std::vector<V3> bforce, btorque;
    double timers[6] = {0, 0, 0, 0, 0, 0};
};

namespace {

const double EDGE_R = 0.1;                 // CMC/ChNarrowphasePRIMS.cpp:31
const double TINYF = (double)1e-12f;       // float literal thresholds in the reference

// ---- GenerateAABB, CMC/ChCollisionSystemMulticore.cpp:376-545 -------------------------------------------------
void gen_aabb(orc_system& S) {
    const double env = S.st.system_type == 1 ? 0.0 : S.st.collision_envelope;
    S.amin.assign(S.ns, v3(0, 0, 0));
    S.amax.assign(S.ns, v3(0, 0, 0));
    S.wpos.resize(S.ns);
    S.wrot.resize(S.ns);
#pragma omp parallel for
    for (int i = 0; i < S.ns; i++) {
        unsigned id = S.sbody[i];
        if (id == UINT_MAX)
            continue;
        const double* d = &S.dims[4 * i];
        Q brot = S.rotq[id];
        Q r = qmul(brot, S.lrot[i]);
        V3 p = rot(S.lpos[i], brot) + S.pos[id];
        S.wpos[i] = toParent(S.pos[id], brot, S.lpos[i]);  // PreprocessLocalToParent, CMC/ChNarrowphase.cpp:173-196
        S.wrot[i] = r;
        V3 h;
        switch (S.typ[i]) {
            case T_SPHERE:
                S.amin[i] = p - (d[0] + env);
                S.amax[i] = p + (d[0] + env);
                continue;
            case T_ELLIPSOID: case T_BOX: case T_CYLINDER: case T_CYLSHELL: case T_CONE:
                h = v3(d[0], d[1], d[2]) + env;
                break;
            case T_RBOX: case T_RCYL:
                h = v3(d[0], d[1], d[2]) + d[3] + env;
                break;
            case T_CAPSULE:
                h = v3(d[0], d[0] + d[1], d[0]) + env;
                break;
            default:
                continue;
        }
        V3 t = absrot(r, h);
        S.amin[i] = p - t;
        S.amax[i] = p + t;
    }
}

// ---- ChBroadphase::Process, CMC/ChBroadphase.cpp:85-345 -------------------------------------------------------
inline int hmin(double a, double inv) { return (int)std::floor(a * inv); }        // CMC/ChCollisionUtils.h:44-51
inline int hmax(double a, double inv) { return (int)std::ceil(a * inv) - 1; }     // :54-60

void broadphase(orc_system& S) {
    S.pairs.clear();
    S.N_i = S.N_a = 0;
    // RigidBoundingBox: union over shapes of colliding bodies (:85-100)
    V3 lo = v3(DBL_MAX, DBL_MAX, DBL_MAX), hi = v3(-DBL_MAX, -DBL_MAX, -DBL_MAX);
    bool any = false;
    for (int i = 0; i < S.ns; i++) {
        unsigned id = S.sbody[i];
        if (id == UINT_MAX || !S.collide[id])
            continue;
        any = true;
        lo = v3(std::min(lo.x, S.amin[i].x), std::min(lo.y, S.amin[i].y), std::min(lo.z, S.amin[i].z));
        hi = v3(std::max(hi.x, S.amax[i].x), std::max(hi.y, S.amax[i].y), std::max(hi.z, S.amax[i].z));
    }
    if (!any)
        lo = hi = v3(0, 0, 0);
    // DetermineBoundingBox (:142-166)
    const double fraction = 1e-3;
    V3 size = hi - lo;
    S.origin = lo - fraction * size;
    S.maxpt = hi + fraction * size;
    // OffsetAABB (:168-176)
    for (int i = 0; i < S.ns; i++) {
        S.amin[i] = S.amin[i] - S.origin;
        S.amax[i] = S.amax[i] - S.origin;
    }
    // ComputeTopLevelResolution (:179-208), Compute_Grid_Resolution (CMC/ChCollisionUtilsBroadphase.cpp:36-46)
    V3 diag = v3(std::abs(S.maxpt.x - S.origin.x), std::abs(S.maxpt.y - S.origin.y), std::abs(S.maxpt.z - S.origin.z));
    if (S.st.broadphase_grid == 0) {
        for (int a = 0; a < 3; a++)
            S.bins[a] = S.st.bins_per_axis[a];
    } else if (S.st.broadphase_grid == 1) {
        S.bins[0] = (int)std::ceil(diag.x / S.st.bin_size[0]);
        S.bins[1] = (int)std::ceil(diag.y / S.st.bin_size[1]);
        S.bins[2] = (int)std::ceil(diag.z / S.st.bin_size[2]);
    } else {
        double V = diag.x * diag.y * diag.z;
        double k = S.st.grid_density;
        unsigned n = (unsigned)S.ns;
        S.bins[0] = std::max(1, int(diag.x * std::pow(k * n / V, 1.0 / 3.0)));
        S.bins[1] = std::max(1, int(diag.y * std::pow(k * n / V, 1.0 / 3.0)));
        S.bins[2] = std::max(1, int(diag.z * std::pow(k * n / V, 1.0 / 3.0)));
    }
    S.bsize = v3(diag.x / S.bins[0], diag.y / S.bins[1], diag.z / S.bins[2]);
    S.inv = v3(1.0 / S.bsize.x, 1.0 / S.bsize.y, 1.0 / S.bsize.z);
    if (S.ns == 0)
        return;

    // OneLevelBroadphase (:228-345): (bin, shape) list, sorted by bin (stable => shapes ascending per bin)
    std::vector<std::pair<unsigned, unsigned>> ent;
    for (int i = 0; i < S.ns; i++) {
        if (S.sbody[i] == UINT_MAX)
            continue;
        int x0 = hmin(S.amin[i].x, S.inv.x), y0 = hmin(S.amin[i].y, S.inv.y), z0 = hmin(S.amin[i].z, S.inv.z);
        int x1 = hmax(S.amax[i].x, S.inv.x), y1 = hmax(S.amax[i].y, S.inv.y), z1 = hmax(S.amax[i].z, S.inv.z);
        for (int x = x0; x <= x1; x++)
            for (int y = y0; y <= y1; y++)
                for (int z = z0; z <= z1; z++)
                    ent.push_back({(unsigned)((z * S.bins[1]) * S.bins[0] + y * S.bins[0] + x), (unsigned)i});
    }
    S.N_i = (unsigned)ent.size();
    std::stable_sort(ent.begin(), ent.end(),
                     [](const std::pair<unsigned, unsigned>& a, const std::pair<unsigned, unsigned>& b) {
                         return a.first < b.first;
                     });
    // per-bin all-pairs with the filters of CMC/ChCollisionUtilsBroadphase.cpp:86-217
    size_t s = 0;
    while (s < ent.size()) {
        size_t e = s;
        while (e < ent.size() && ent[e].first == ent[s].first)
            e++;
        S.N_a++;
        const unsigned bin = ent[s].first;
        if (e - s > 1) {
            for (size_t i = s; i < e; i++) {
                unsigned sa = ent[i].second, ba = S.sbody[sa];
                if (ba == UINT_MAX || !S.collide[ba])
                    continue;
                for (size_t k = i + 1; k < e; k++) {
                    unsigned sb = ent[k].second, bb = S.sbody[sb];
                    if (bb == UINT_MAX || sa == sb || ba == bb || !S.collide[bb])
                        continue;
                    if (!S.active[ba] && !S.active[bb])
                        continue;
                    short ga = S.fam[2 * sa], ma = S.fam[2 * sa + 1], gb = S.fam[2 * sb], mb = S.fam[2 * sb + 1];
                    if (!((ma & gb) && (mb & ga)))  // collide(), CMC/ChCollisionUtils.h:84-88
                        continue;
                    const V3 &An = S.amin[sa], &Ax = S.amax[sa], &Bn = S.amin[sb], &Bx = S.amax[sb];
                    if (!((An.x <= Bx.x && Bn.x <= Ax.x) && (An.y <= Bx.y && Bn.y <= Ax.y) &&
                          (An.z <= Bx.z && Bn.z <= Ax.z)))  // overlap(), :91-95
                        continue;
                    // current_bin(), :97-110
                    int cx = hmin(std::max(An.x, Bn.x), S.inv.x), cy = hmin(std::max(An.y, Bn.y), S.inv.y),
                        cz = hmin(std::max(An.z, Bn.z), S.inv.z);
                    if ((unsigned)((cz * S.bins[1]) * S.bins[0] + cy * S.bins[0] + cx) != bin)
                        continue;
                    unsigned a = std::min(sa, sb), b = std::max(sa, sb);
                    S.pairs.push_back(((int64_t)a << 32) | (int64_t)b);
                }
            }
        }
        s = e;
    }
}

// ---- analytic pair tests, CMC/ChNarrowphasePRIMS.cpp -----------------------------------------------------------
struct Hit {
    V3 n, p1, p2;
    double depth, erad;
};

unsigned snapBox(V3 h, V3& l) {  // CMC/ChCollisionUtils.h:545-562
    unsigned code = 0;
    if (std::abs(l.x) > h.x) code |= 1, l.x = (l.x > 0) ? h.x : -h.x;
    if (std::abs(l.y) > h.y) code |= 2, l.y = (l.y > 0) ? h.y : -h.y;
    if (std::abs(l.z) > h.z) code |= 4, l.z = (l.z > 0) ? h.z : -h.z;
    return code;
}
unsigned snapCyl(double rad, double hlen, V3& l) {  // CMC/ChCollisionUtils.h:494-516
    unsigned code = 0;
    if (l.z > hlen) code |= 1, l.z = hlen;
    else if (l.z < -hlen) code |= 1, l.z = -hlen;
    double d2 = l.x * l.x + l.y * l.y;
    if (d2 > rad * rad) {
        code |= 2;
        double d = std::sqrt(d2);
        l.x *= (rad / d);
        l.y *= (rad / d);
    }
    return code;
}

bool sphereSphere(V3 p1, double r1, V3 p2, double r2, double sep, Hit& h) {  // :40-71
    V3 d = p2 - p1;
    double d2 = dot(d, d), rs = r1 + r2, rss = rs + sep;
    if (d2 >= rss * rss || d2 < 1e-12)
        return false;
    double dist = std::sqrt(d2);
    h.n = d / dist;
    h.p1 = p1 + h.n * r1;
    h.p2 = p2 - h.n * r2;
    h.depth = dist - rs;
    h.erad = r1 * r2 / rs;
    return true;
}
bool capsuleSphere(V3 p1, Q q1, double r1, double hl, V3 p2, double r2, double sep, Hit& h) {  // :81-124
    V3 W = colW(q1);
    double a = dot(p2 - p1, W);
    double lo = -hl, hi = hl;
    if (lo > hi) std::swap(lo, hi);
    a = a < lo ? lo : (a > hi ? hi : a);
    V3 loc = p1 + a * W;
    double rs = r1 + r2, rss = rs + sep;
    V3 d = p2 - loc;
    double d2 = dot(d, d);
    if (d2 >= rss * rss || d2 <= TINYF)
        return false;
    double dist = std::sqrt(d2);
    h.n = d / dist;
    h.p1 = loc + h.n * r1;
    h.p2 = p2 - h.n * r2;
    h.depth = dist - rs;
    h.erad = r1 * r2 / rs;
    return true;
}
bool cylinderSphere(V3 p1, Q q1, double r1, double hl, V3 p2, double r2, double sep, Hit& h) {  // :134-189
    V3 sp = toLocal(p1, q1, p2), cp = sp;
    unsigned code = snapCyl(r1, hl, cp);
    if (!code)
        return false;
    V3 d = sp - cp;
    double d2 = dot(d, d), r2s = r2 + sep;
    if (d2 >= r2s * r2s || d2 <= TINYF)
        return false;
    double dist = std::sqrt(d2);
    h.depth = dist - r2;
    h.n = rot(d / dist, q1);
    h.p1 = toParent(p1, q1, cp);
    h.p2 = p2 - h.n * r2;
    h.erad = code == 1 ? r2 : (code == 2 ? r1 * r2 / (r1 + r2) : r2 * EDGE_R / (r2 + EDGE_R));
    return true;
}
bool rcylSphere(V3 p1, Q q1, double r1, double hl, double sr, V3 p2, double r2, double sep, Hit& h) {  // :200-260
    V3 sp = toLocal(p1, q1, p2), cp = sp;
    unsigned code = snapCyl(r1, hl, cp);
    if (!code)
        return false;
    double rs = sr + r2, rss = rs + sep;
    V3 d = sp - cp;
    double d2 = dot(d, d);
    if (d2 >= rss * rss || d2 <= TINYF)
        return false;
    double dist = std::sqrt(d2);
    h.depth = dist - rs;
    h.n = rot(d / dist, q1);
    h.p2 = p2 - h.n * r2;
    h.p1 = h.p2 - h.depth * h.n;
    h.erad = code == 1 ? r2 : (code == 2 ? r1 * r2 / (r1 + r2) : sr * r2 / (sr + r2));
    return true;
}
bool boxSphere(V3 p1, Q q1, V3 hd, V3 p2, double r2, double sep, Hit& h) {  // :269-312
    V3 sp = toLocal(p1, q1, p2), bp = sp;
    unsigned code = snapBox(hd, bp);
    V3 d = sp - bp;
    double d2 = dot(d, d), r2s = r2 + sep;
    if (d2 >= r2s * r2s || d2 <= TINYF)
        return false;
    double dist = std::sqrt(d2);
    h.depth = dist - r2;
    h.n = rot(d / dist, q1);
    h.p1 = toParent(p1, q1, bp);
    h.p2 = p2 - h.n * r2;
    h.erad = (code != 1 && code != 2 && code != 4) ? r2 * EDGE_R / (r2 + EDGE_R) : r2;
    return true;
}
bool rboxSphere(V3 p1, Q q1, V3 hd, double sr, V3 p2, double r2, double sep, Hit& h) {  // :323-370
    V3 sp = toLocal(p1, q1, p2), bp = sp;
    unsigned code = snapBox(hd, bp);
    double rs = sr + r2, rss = rs + sep;
    V3 d = sp - bp;
    double d2 = dot(d, d);
    if (d2 >= rss * rss || d2 <= TINYF)
        return false;
    double dist = std::sqrt(d2);
    h.depth = dist - rs;
    h.n = rot(d / dist, q1);
    h.p2 = p2 - h.n * r2;
    h.p1 = h.p2 - h.depth * h.n;
    h.erad = (code != 1 && code != 2 && code != 4) ? r2 * sr / (r2 + sr) : r2;
    return true;
}

// one-contact subset of PRIMSCollision, CMC/ChNarrowphasePRIMS.cpp:1406-1607
int primsPair(const orc_system& S, unsigned a, unsigned b, double sep, Hit& h) {
    int ta = S.typ[a], tb = S.typ[b];
    const double *da = &S.dims[4 * a], *db = &S.dims[4 * b];
    if (ta == T_SPHERE && tb == T_SPHERE)
        return sphereSphere(S.wpos[a], da[0], S.wpos[b], db[0], sep, h);
    bool swap = false;
    if (ta == T_SPHERE) {
        std::swap(a, b), std::swap(ta, tb), std::swap(da, db);
        swap = true;
    } else if (tb != T_SPHERE) {
        return -1;
    }
    bool hit;
    switch (ta) {
        case T_CAPSULE: hit = capsuleSphere(S.wpos[a], S.wrot[a], da[0], da[1], S.wpos[b], db[0], sep, h); break;
        case T_CYLINDER: hit = cylinderSphere(S.wpos[a], S.wrot[a], da[0], da[1], S.wpos[b], db[0], sep, h); break;
        case T_RCYL: hit = rcylSphere(S.wpos[a], S.wrot[a], da[0], da[1], da[3], S.wpos[b], db[0], sep, h); break;
        case T_BOX: hit = boxSphere(S.wpos[a], S.wrot[a], v3(da[0], da[1], da[2]), S.wpos[b], db[0], sep, h); break;
        case T_RBOX: hit = rboxSphere(S.wpos[a], S.wrot[a], v3(da[0], da[1], da[2]), da[3], S.wpos[b], db[0], sep, h); break;
        default: return -1;
    }
    if (hit && swap) {
        std::swap(h.p1, h.p2);
        h.n = -h.n;
    }
    return hit ? 1 : 0;
}

// ChNarrowphase::ProcessRigidRigid, CMC/ChNarrowphase.cpp:325-386 (stable compaction keeps pair order)
void narrowphase(orc_system& S) {
    const double env = S.st.system_type == 1 ? 0.0 : S.st.collision_envelope;
    const size_t np = S.pairs.size();
    std::vector<Hit> hits(np);
    std::vector<char> ok(np, 0);
#pragma omp parallel for
    for (long i = 0; i < (long)np; i++) {
        unsigned a = (unsigned)(S.pairs[i] >> 32), b = (unsigned)(S.pairs[i] & 0xffffffff);
        ok[i] = primsPair(S, a, b, 2 * env, hits[i]) == 1;
    }
    S.contacts.clear();
    for (size_t i = 0; i < np; i++) {
        if (!ok[i])
            continue;
        unsigned a = (unsigned)(S.pairs[i] >> 32), b = (unsigned)(S.pairs[i] & 0xffffffff);
        S.contacts.push_back(Contact{hits[i].n, hits[i].p1, hits[i].p2, hits[i].depth, hits[i].erad, (int)S.sbody[a],
                                     (int)S.sbody[b], S.pairs[i]});
    }
}

// ---- composite material, src/chrono/physics/ChContactMaterialSMC.cpp:107-130 (float arithmetic) ----------------
struct Comp {
    double mu, muRoll, muSpin, E, G, adh, dmt, perko, cr, kn, kt, gn, gt;
};
Comp composite(const float* m1, const float* m2) {
    float inv_E = (1 - m1[1] * m1[1]) / m1[0] + (1 - m2[1] * m2[1]) / m2[0];
    float inv_G = 2 * (2 - m1[1]) * (1 + m1[1]) / m1[0] + 2 * (2 - m2[1]) * (1 + m2[1]) / m2[0];
    float E = 1 / inv_E, G = 1 / inv_G;
    Comp c;
    c.E = E, c.G = G;
    c.mu = std::min<float>(m1[2], m2[2]);
    c.muRoll = std::min<float>(m1[4], m2[4]);
    c.muSpin = std::min<float>(m1[5], m2[5]);
    c.cr = std::min<float>(m1[6], m2[6]);
    c.adh = std::min<float>(m1[7], m2[7]);
    c.dmt = std::min<float>(m1[8], m2[8]);
    c.perko = std::min<float>(m1[9], m2[9]);
    c.kn = (float)((m1[10] + m2[10]) / 2);
    c.kt = (float)((m1[11] + m2[11]) / 2);
    c.gn = (float)((m1[12] + m2[12]) / 2);
    c.gt = (float)((m1[13] + m2[13]) / 2);
    return c;
}

// ---- function_CalcContactForces, MC/solver/ChIterativeSolverMulticoreSMC.cpp:56-552 (None / OneStep history) ----
void contactForce(const orc_system& S, const Contact& c, V3& force, V3& tq1, V3& tq2) {
    force = tq1 = tq2 = v3(0, 0, 0);
    if (c.depth >= 0)
        return;
    const cb200_settings& st = S.st;
    const int b1 = c.b1, b2 = c.b2;
    const double PI = 3.141592653589793238462643383279;
    V3 s1 = toLocal(S.pos[b1], S.rotq[b1], c.pa), s2 = toLocal(S.pos[b2], S.rotq[b2], c.pb);
    V3 v1 = v3(S.vel[6 * b1], S.vel[6 * b1 + 1], S.vel[6 * b1 + 2]), w1 = v3(S.vel[6 * b1 + 3], S.vel[6 * b1 + 4], S.vel[6 * b1 + 5]);
    V3 v2 = v3(S.vel[6 * b2], S.vel[6 * b2 + 1], S.vel[6 * b2 + 2]), w2 = v3(S.vel[6 * b2 + 3], S.vel[6 * b2 + 4], S.vel[6 * b2 + 5]);
    V3 vp1 = v1 + rot(cross(w1, s1), S.rotq[b1]), vp2 = v2 + rot(cross(w2, s2), S.rotq[b2]);
    V3 rv = vp2 - vp1;
    double rvn = dot(rv, c.n);
    V3 rvt = rv - rvn * c.n;
    double rvt_mag = len(rvt);
    double m_eff = S.mass[b1] * S.mass[b2] / (S.mass[b1] + S.mass[b2]);
    unsigned sa = (unsigned)(c.key >> 32), sb = (unsigned)(c.key & 0xffffffff);
    Comp M = composite(&S.mats[14 * S.mat[sa]], &S.mats[14 * S.mat[sb]]);
    double muRoll = M.muRoll, muSpin = M.muSpin, cr = M.cr;
    double kn = 0, kt = 0, gn = 0, gt = 0, kns = 0, gns = 0;
    const double t_contact = 0, dn = -c.depth, eps = DBL_EPSILON;
    V3 dt = v3(0, 0, 0);
    if (st.tangential_displ_mode == 1)
        dt = rvt * st.step_size;
    double cv = st.characteristic_vel;
    bool coulomb = false;
    switch (st.contact_force_model) {
        case 0:
            if (st.use_material_properties) {
                double tk = (16.0 / 15) * std::sqrt(c.erad) * M.E, vv = cv * cv;
                double lg = (cr < eps) ? std::log(eps) : std::log(cr);
                lg = (cr > 1 - eps) ? std::log(1 - eps) : lg;
                double tg = 1 + std::pow(PI / lg, 2);
                kn = tk * std::pow(m_eff * vv / tk, 0.2);
                kt = kn;
                gn = std::sqrt(4 * m_eff * kn / tg);
                gt = gn;
            } else {
                kn = M.kn, kt = M.kt, gn = m_eff * M.gn, gt = m_eff * M.gt;
            }
            kns = kn, gns = gn;
            break;
        case 1:
            if (st.use_material_properties) {
                double sq = std::sqrt(c.erad * dn), Sn = 2 * M.E * sq, St = 8 * M.G * sq;
                double lg = (cr < eps) ? std::log(eps) : std::log(cr);
                double beta = lg / std::sqrt(lg * lg + PI * PI);
                kn = (2.0 / 3.0) * Sn;
                kt = St;
                gn = -2 * std::sqrt(5.0 / 6) * beta * std::sqrt(Sn * m_eff);
                gt = -2 * std::sqrt(5.0 / 6) * beta * std::sqrt(St * m_eff);
            } else {
                double t = c.erad * std::sqrt(dn);
                kn = t * M.kn, kt = t * M.kt, gn = t * m_eff * M.gn, gt = t * m_eff * M.gt;
            }
            kns = kn / std::sqrt(dn);
            gns = gn / std::pow(dn, 0.25);
            break;
        case 3:
            if (st.use_material_properties) {
                double sq = std::sqrt(c.erad * dn), Sn = 2 * M.E * sq, St = 8 * M.G * sq;
                cr = (cr < 0.01) ? 0.01 : cr;
                cr = (cr > 1.0 - eps) ? 1.0 - eps : cr;
                double lg = std::log(cr), beta = lg / std::sqrt(lg * lg + PI * PI);
                kn = (2.0 / 3.0) * Sn;
                kt = (2.0 / 3.0) * St;
                gn = 8.0 * (1.0 - cr) * kn * dn / (5.0 * cr * cv);
                gt = -2 * std::sqrt(5.0 / 6) * beta * std::sqrt(St * m_eff);
            } else {
                double t = c.erad * std::sqrt(dn);
                kn = t * M.kn, kt = t * M.kt, gn = t * m_eff * M.gn * dn, gt = t * m_eff * M.gt;
            }
            kns = kn / std::sqrt(dn);
            gns = gn / std::pow(dn, 1.5);
            break;
        default:
            if (st.use_material_properties) {
                double sq = std::sqrt(dn), Sn = 2 * M.E * sq;
                double lg = (cr < eps) ? std::log(eps) : std::log(cr);
                double beta = lg / std::sqrt(lg * lg + PI * PI);
                kn = (2.0 / 3.0) * Sn;
                gn = -2 * std::sqrt(5.0 / 6) * beta * std::sqrt(Sn * m_eff);
            } else {
                double t = std::sqrt(dn);
                kn = t * M.kn, gn = t * M.gn;
            }
            kns = kn / std::sqrt(dn);
            gns = gn / std::pow(dn, 0.25);
            coulomb = true;
            break;
    }
    double fN = kn * dn - gn * rvn;
    if (coulomb) {
        double fT = M.mu * std::tanh(5.0 * rvt_mag) * fN;
        force = fN * c.n;
        if (rvt_mag >= st.min_slip_vel)
            force = force - (fT / rvt_mag) * rvt;
    } else {
        V3 fd = gt * rvt;
        V3 fT = kt * dt + fd;
        double fT_mag = len(fT), dt_mag = len(dt), slide = M.mu * std::abs(fN);
        if (fT_mag > slide) {
            if (dt_mag > eps)
                fT = fT * (slide / fT_mag);
            else
                fT = v3(0, 0, 0);
        }
        force = fN * c.n - fT;
    }
    V3 t1 = cross(s1, rotT(force, S.rotq[b1])), t2 = cross(s2, rotT(force, S.rotq[b2]));
    double dco = gns / (2.0 * m_eff * std::sqrt(kns / m_eff));
    if (dco < 1.0) {
        double tc = PI * std::sqrt(m_eff / (kns * (1 - dco * dco)));
        if (t_contact <= tc)
            muRoll = muSpin = 0.0;
    }
    V3 vrot = rot(cross(w2, s2), S.rotq[b2]) - rot(cross(w1, s1), S.rotq[b1]);
    V3 relo = rot(w2, S.rotq[b2]) - rot(w1, S.rotq[b1]);
    V3 mr1 = v3(0, 0, 0), mr2 = mr1, ms1 = mr1, ms2 = mr1;
    if (len(vrot) > st.min_roll_vel && muRoll > eps) {
        mr1 = muRoll * cross(fN * s1, rotT(vrot, S.rotq[b1])) / len(vrot);
        mr2 = muRoll * cross(fN * s2, rotT(vrot, S.rotq[b2])) / len(vrot);
    }
    if (len(relo) > st.min_spin_vel && muSpin > eps) {
        double r1 = len(s1), r2 = len(s2);
        double xc = (r1 * r1 - r2 * r2) / (2 * (r1 + r2 - dn)) + 0.5 * (r1 + r2 - dn);
        double rc = r1 * r1 - xc * xc;
        rc = (rc < eps) ? eps : std::sqrt(rc);
        ms1 = muSpin * rc * rotT(dot(relo, fN * c.n) * c.n, S.rotq[b1]) / len(relo);
        ms2 = muSpin * rc * rotT(dot(relo, fN * c.n) * c.n, S.rotq[b2]) / len(relo);
    }
    if (st.adhesion_force_model == 0)
        force = force - M.adh * c.n;
    else if (st.adhesion_force_model == 1)
        force = force - M.dmt * std::sqrt(c.erad) * c.n;
    else
        force = force - M.perko * c.erad * c.n;
    tq1 = -t1 + mr1 + ms1;
    tq2 = t2 - mr2 - ms2;
}

// ---- UpdateRigidBodies force part, src/chrono/physics/ChBody.cpp:552-574,423-426,247-257 ------------------------
void updateForces(orc_system& S) {
    const double h = S.st.step_size;
    V3 g = v3(S.st.gravity[0], S.st.gravity[1], S.st.gravity[2]);
    S.hf.resize(6 * (size_t)S.nb);
#pragma omp parallel for
    for (int b = 0; b < S.nb; b++) {
        V3 w = v3(S.vel[6 * b + 3], S.vel[6 * b + 4], S.vel[6 * b + 5]);
        V3 Jw = v3(S.inertia[3 * b] * w.x, S.inertia[3 * b + 1] * w.y, S.inertia[3 * b + 2] * w.z);
        V3 gyro = cross(w, Jw);
        V3 F = h * (g * S.mass[b]);
        V3 T = h * (v3(0, 0, 0) - gyro);
        double* o = &S.hf[6 * b];
        o[0] = F.x, o[1] = F.y, o[2] = F.z, o[3] = T.x, o[4] = T.y, o[5] = T.z;
    }
}

// ---- ProcessContacts, MC/solver/ChIterativeSolverMulticoreSMC.cpp:653-738 ---------------------------------------
void processContacts(orc_system& S, double* tForce, double* tAcc) {
    const size_t nc = S.contacts.size();
    double t0 = now();
    S.cforce.resize(nc), S.ctq1.resize(nc), S.ctq2.resize(nc);
#pragma omp parallel for
    for (long i = 0; i < (long)nc; i++)
        contactForce(S, S.contacts[i], S.cforce[i], S.ctq1[i], S.ctq2[i]);
    double t1 = now();
    // sort (body, 2i+side) keys stable by body, reduce per body in that order
    std::vector<std::pair<int, unsigned>> keys(2 * nc);
    for (size_t i = 0; i < nc; i++) {
        keys[2 * i] = {S.contacts[i].b1, (unsigned)(2 * i)};
        keys[2 * i + 1] = {S.contacts[i].b2, (unsigned)(2 * i + 1)};
    }
    std::stable_sort(keys.begin(), keys.end(),
                     [](const std::pair<int, unsigned>& a, const std::pair<int, unsigned>& b) { return a.first < b.first; });
    S.bforce.assign(S.nb, v3(0, 0, 0));
    S.btorque.assign(S.nb, v3(0, 0, 0));
    const double h = S.st.step_size;
    size_t s = 0;
    while (s < keys.size()) {
        size_t e = s;
        V3 F = v3(0, 0, 0), T = F;
        while (e < keys.size() && keys[e].first == keys[s].first) {
            unsigned k = keys[e].second, ci = k >> 1;
            V3 f = (k & 1) ? S.cforce[ci] : -S.cforce[ci];
            V3 t = (k & 1) ? S.ctq2[ci] : S.ctq1[ci];
            if (e == s)
                F = f, T = t;
            else
                F = F + f, T = T + t;
            e++;
        }
        int b = keys[s].first;
        S.bforce[b] = F, S.btorque[b] = T;
        V3 hF = h * F, hT = h * T;  // host_AddContactForces, :611-626
        double* o = &S.hf[6 * (size_t)b];
        o[0] += hF.x, o[1] += hF.y, o[2] += hF.z, o[3] += hT.x, o[4] += hT.y, o[5] += hT.z;
        s = e;
    }
    *tForce += t1 - t0;
    *tAcc += now() - t1;
}

// ---- ComputeInvMassMatrix/ComputeImpulses + VariablesQbIncrementPosition ---------------------------------------
// MC/solver/ChIterativeSolverMulticore.cpp:61-147, ...SMC.cpp:854-868, src/chrono/physics/ChBody.cpp:288-310
void advance(orc_system& S) {
    const double h = S.st.step_size;
#pragma omp parallel for
    for (int b = 0; b < S.nb; b++) {
        if (!S.active[b])
            continue;
        double im = 1.0 / S.mass[b];
        double* v = &S.vel[6 * (size_t)b];
        const double* f = &S.hf[6 * (size_t)b];
        for (int k = 0; k < 3; k++)
            v[k] = v[k] + im * f[k];
        for (int k = 0; k < 3; k++)
            v[3 + k] = v[3 + k] + (1.0 / S.inertia[3 * b + k]) * f[3 + k];
        S.pos[b] = S.pos[b] + v3(v[0], v[1], v[2]) * h;
        Q q = S.rotq[b];
        double ww = q.w * q.w, xx = q.x * q.x, yy = q.y * q.y, zz = q.z * q.z;
        double wx = q.w * q.x, wy = q.w * q.y, wz = q.w * q.z, xy = q.x * q.y, xz = q.x * q.z, yz = q.y * q.z;
        double ax = ((ww + xx) * 2 - 1) * v[3] + ((xy - wz) * 2) * v[4] + ((xz + wy) * 2) * v[5];
        double ay = ((xy + wz) * 2) * v[3] + ((ww + yy) * 2 - 1) * v[4] + ((yz - wx) * 2) * v[5];
        double az = ((xz - wy) * 2) * v[3] + ((yz + wx) * 2) * v[4] + ((ww + zz) * 2 - 1) * v[5];
        double l = std::sqrt(ax * ax + ay * ay + az * az);
        double ang = l * h;
        if (l < DBL_MIN) {
            ax = 1, ay = 0, az = 0;
        } else {
            double s = 1 / l;
            ax *= s, ay *= s, az *= s;
        }
        double half = ang / 2, sh = std::sin(half);
        Q d = Q{std::cos(half), ax * sh, ay * sh, az * sh};
        S.rotq[b] = Q{d.w * q.w - d.x * q.x - d.y * q.y - d.z * q.z, d.w * q.x + d.x * q.w - d.z * q.y + d.y * q.z,
                      d.w * q.y + d.y * q.w + d.z * q.x - d.x * q.z, d.w * q.z + d.z * q.w - d.y * q.x + d.x * q.y};
    }
}

}  // namespace

#include "oracle_nsc.inc"

extern "C" {

orc_system* orc_create(void) {
    orc_system* S = new orc_system();
    memset(&S->st, 0, sizeof(S->st));
    return S;
}
void orc_destroy(orc_system* S) {
    delete S->nsc;
    delete S;
}
void orc_set_threads(int n) { omp_set_num_threads(n > 0 ? n : 1); }
int orc_max_threads(void) { return omp_get_max_threads(); }
void orc_set_settings(orc_system* S, const cb200_settings* s) { S->st = *s; }

// The reference's defaults (collision_settings / solver_settings constructors, src/chrono_multicore/ChSettings.h:39-246)
// for the fields the boundary struct carries; the checkers keep their own table (they never load the product library).
void orc_default_settings(cb200_settings* s) {
    memset(s, 0, sizeof(*s));
    s->collision_envelope = 0;
    s->broadphase_grid = 0;  // FIXED_RESOLUTION, 10 x 10 x 10
    s->bins_per_axis[0] = s->bins_per_axis[1] = s->bins_per_axis[2] = 10;
    s->bin_size[0] = s->bin_size[1] = s->bin_size[2] = 1;
    s->grid_density = 5;
    s->narrowphase_algorithm = 2;  // HYBRID
    s->system_type = 0;
    s->step_size = 0.01;
    s->contact_force_model = 1;    // Hertz
    s->adhesion_force_model = 0;   // Constant
    s->tangential_displ_mode = 1;  // OneStep
    s->use_material_properties = 1;
    s->characteristic_vel = 1;
    s->min_slip_vel = 1e-4;
    s->min_roll_vel = 1e-4;
    s->min_spin_vel = 1e-4;
    s->solver_type = 0;            // APGD
    s->solver_mode = 1;            // SLIDING
    s->max_iteration_normal = 0;
    s->max_iteration_sliding = 100;
    s->max_iteration_spinning = 0;
    s->cache_step_length = 0;
    s->alpha = .2;
    s->contact_recovery_speed = .6;
    s->tolerance = 1e-4;
    s->deterministic = 1;
    s->test_objective = 0;         // MC/ChSettings.h:100
    s->tolerance_objective = 1e-6;  // MC/ChSettings.h:98
}

void orc_set_bodies(orc_system* S, int nb, const double* mass, const double* inertia, const char* active,
                    const char* collide) {
    S->nb = nb;
    S->mass.assign(mass, mass + nb);
    S->inertia.assign(inertia, inertia + 3 * (size_t)nb);
    S->active.assign(active, active + nb);
    S->collide.assign(collide, collide + nb);
    S->pos.assign(nb, v3(0, 0, 0));
    S->rotq.assign(nb, Q{1, 0, 0, 0});
    S->vel.assign(6 * (size_t)nb, 0.0);
    S->hf.assign(6 * (size_t)nb, 0.0);
}

void orc_set_shapes(orc_system* S, int ns, const int32_t* type, const uint32_t* body, const int16_t* fam,
                    const double* lpos, const double* lrot, const double* dims, const int32_t* mat) {
    S->ns = ns;
    S->typ.assign(type, type + ns);
    S->sbody.assign(body, body + ns);
    S->fam.assign(fam, fam + 2 * (size_t)ns);
    S->dims.assign(dims, dims + 4 * (size_t)ns);
    S->mat.assign(ns, 0);
    S->lpos.resize(ns), S->lrot.resize(ns);
    for (int i = 0; i < ns; i++) {
        S->lpos[i] = v3(lpos[3 * i], lpos[3 * i + 1], lpos[3 * i + 2]);
        S->lrot[i] = Q{lrot[4 * i], lrot[4 * i + 1], lrot[4 * i + 2], lrot[4 * i + 3]};
        if (mat)
            S->mat[i] = mat[i];
    }
}

void orc_set_materials_smc(orc_system* S, int nm, const float* t) {
    S->nm = nm;
    S->mats.assign(t, t + 14 * (size_t)nm);
}

void orc_set_state(orc_system* S, const double* pos, const double* rot, const double* vel) {
    for (int b = 0; b < S->nb; b++) {
        if (pos)
            S->pos[b] = v3(pos[3 * b], pos[3 * b + 1], pos[3 * b + 2]);
        if (rot)
            S->rotq[b] = Q{rot[4 * b], rot[4 * b + 1], rot[4 * b + 2], rot[4 * b + 3]};
    }
    if (vel)
        S->vel.assign(vel, vel + 6 * (size_t)S->nb);
}

void orc_get_state(orc_system* S, double* pos, double* rot, double* vel) {
    for (int b = 0; b < S->nb; b++) {
        if (pos)
            pos[3 * b] = S->pos[b].x, pos[3 * b + 1] = S->pos[b].y, pos[3 * b + 2] = S->pos[b].z;
        if (rot)
            rot[4 * b] = S->rotq[b].w, rot[4 * b + 1] = S->rotq[b].x, rot[4 * b + 2] = S->rotq[b].y,
                    rot[4 * b + 3] = S->rotq[b].z;
    }
    if (vel)
        std::copy(S->vel.begin(), S->vel.end(), vel);
}

void orc_collide(orc_system* S) {
    gen_aabb(*S);
    broadphase(*S);
    narrowphase(*S);
}

void orc_smc_step(orc_system* S, int n) {
    for (int k = 0; k < 6; k++)
        S->timers[k] = 0;
    for (int it = 0; it < n; it++) {
        double t0 = now();
        updateForces(*S);
        double t1 = now();
        gen_aabb(*S);
        broadphase(*S);
        double t2 = now();
        narrowphase(*S);
        double t3 = now();
        S->bforce.assign(S->nb, v3(0, 0, 0));
        S->btorque.assign(S->nb, v3(0, 0, 0));
        if (!S->contacts.empty())
            processContacts(*S, &S->timers[3], &S->timers[4]);
        double t4 = now();
        advance(*S);
        double t5 = now();
        S->timers[0] += t1 - t0, S->timers[1] += t2 - t1, S->timers[2] += t3 - t2, S->timers[5] += t5 - t4;
    }
}

void orc_get_counts(orc_system* S, uint32_t* c) {
    c[0] = S->N_i, c[1] = S->N_a, c[2] = (uint32_t)S->pairs.size(), c[3] = (uint32_t)S->contacts.size();
    c[4] = S->bins[0], c[5] = S->bins[1], c[6] = S->bins[2];
    c[7] = (uint32_t)(S->bins[0] * S->bins[1] * S->bins[2]);
}
void orc_get_grid(orc_system* S, double* g) {
    const V3* v[4] = {&S->origin, &S->maxpt, &S->bsize, &S->inv};
    for (int k = 0; k < 4; k++)
        g[3 * k] = v[k]->x, g[3 * k + 1] = v[k]->y, g[3 * k + 2] = v[k]->z;
}
void orc_get_aabb(orc_system* S, double* mn, double* mx) {
    for (int i = 0; i < S->ns; i++) {
        mn[3 * i] = S->amin[i].x, mn[3 * i + 1] = S->amin[i].y, mn[3 * i + 2] = S->amin[i].z;
        mx[3 * i] = S->amax[i].x, mx[3 * i + 1] = S->amax[i].y, mx[3 * i + 2] = S->amax[i].z;
    }
}
void orc_get_pairs(orc_system* S, int64_t* p) { std::copy(S->pairs.begin(), S->pairs.end(), p); }
void orc_get_contacts(orc_system* S, double* n, double* pa, double* pb, double* d, double* e, int32_t* bids,
                      int64_t* ids) {
    for (size_t i = 0; i < S->contacts.size(); i++) {
        const Contact& c = S->contacts[i];
        n[3 * i] = c.n.x, n[3 * i + 1] = c.n.y, n[3 * i + 2] = c.n.z;
        pa[3 * i] = c.pa.x, pa[3 * i + 1] = c.pa.y, pa[3 * i + 2] = c.pa.z;
        pb[3 * i] = c.pb.x, pb[3 * i + 1] = c.pb.y, pb[3 * i + 2] = c.pb.z;
        d[i] = c.depth, e[i] = c.erad;
        bids[2 * i] = c.b1, bids[2 * i + 1] = c.b2;
        ids[i] = c.key;
    }
}
void orc_get_contact_forces(orc_system* S, double* f, double* t) {
    for (size_t i = 0; i < S->cforce.size(); i++) {
        const V3 &F = S->cforce[i], &A = S->ctq1[i], &B = S->ctq2[i];
        f[6 * i] = -F.x, f[6 * i + 1] = -F.y, f[6 * i + 2] = -F.z, f[6 * i + 3] = F.x, f[6 * i + 4] = F.y, f[6 * i + 5] = F.z;
        t[6 * i] = A.x, t[6 * i + 1] = A.y, t[6 * i + 2] = A.z, t[6 * i + 3] = B.x, t[6 * i + 4] = B.y, t[6 * i + 5] = B.z;
    }
}
void orc_get_body_forces(orc_system* S, double* f, double* t) {
    for (int b = 0; b < S->nb && b < (int)S->bforce.size(); b++) {
        f[3 * b] = S->bforce[b].x, f[3 * b + 1] = S->bforce[b].y, f[3 * b + 2] = S->bforce[b].z;
        t[3 * b] = S->btorque[b].x, t[3 * b + 1] = S->btorque[b].y, t[3 * b + 2] = S->btorque[b].z;
    }
}
void orc_get_hf(orc_system* S, double* hf) { std::copy(S->hf.begin(), S->hf.end(), hf); }
void orc_get_timers(orc_system* S, double* t) { std::copy(S->timers, S->timers + 6, t); }

}  // extern "C"
