// vmath.cuh -- FP64 3-vector / quaternion helpers for the device code.
//
// Operation order is chosen so that every function returns bit-for-bit what the reference's scalar math path
// returns (src/chrono/multicore_math/simd_non.h:57-75, real4.cpp:139-188) when the translation unit is compiled
// with -fmad=false (no FMA contraction; the reference's x86-64 build has no FMA either).
// Vectors are packed (24 B) instead of the reference's padded 32 B real3.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#define HD __host__ __device__ __forceinline__

struct d3 {
    double x, y, z;
};
struct q4 {
    double w, x, y, z;
};

HD d3 mk3(double x, double y, double z) {
    d3 r;
    r.x = x, r.y = y, r.z = z;
    return r;
}
HD d3 mk3(double s) { return mk3(s, s, s); }
HD q4 mkq(double w, double x, double y, double z) {
    q4 r;
    r.w = w, r.x = x, r.y = y, r.z = z;
    return r;
}

HD d3 operator+(const d3& a, const d3& b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
HD d3 operator-(const d3& a, const d3& b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
HD d3 operator-(const d3& a) { return mk3(-a.x, -a.y, -a.z); }
HD d3 operator*(const d3& a, double s) { return mk3(a.x * s, a.y * s, a.z * s); }
HD d3 operator*(double s, const d3& a) { return mk3(s * a.x, s * a.y, s * a.z); }
HD d3 operator*(const d3& a, const d3& b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
HD d3 operator/(const d3& a, double s) { return mk3(a.x / s, a.y / s, a.z / s); }
HD d3 operator/(const d3& a, const d3& b) { return mk3(a.x / b.x, a.y / b.y, a.z / b.z); }
HD d3 operator+(const d3& a, double s) { return mk3(a.x + s, a.y + s, a.z + s); }
HD d3 operator-(const d3& a, double s) { return mk3(a.x - s, a.y - s, a.z - s); }
HD d3& operator+=(d3& a, const d3& b) {
    a = a + b;
    return a;
}
HD d3& operator-=(d3& a, const d3& b) {
    a = a - b;
    return a;
}
HD d3& operator*=(d3& a, double s) {
    a = a * s;
    return a;
}

// simd_non.h:57-63 (left-to-right sum), :69-75
HD double dot3(const d3& a, const d3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
HD double dot3(const d3& a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
HD d3 cross3(const d3& a, const d3& b) {
    return mk3((a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x));
}
HD double length3(const d3& a) { return sqrt(dot3(a)); }
HD d3 min3(const d3& a, const d3& b) { return mk3(fmin(a.x, b.x), fmin(a.y, b.y), fmin(a.z, b.z)); }
HD d3 max3(const d3& a, const d3& b) { return mk3(fmax(a.x, b.x), fmax(a.y, b.y), fmax(a.z, b.z)); }
HD d3 abs3(const d3& a) { return mk3(fabs(a.x), fabs(a.y), fabs(a.z)); }

HD q4 conj(const q4& q) { return mkq(q.w, -q.x, -q.y, -q.z); }
HD d3 qvec(const q4& q) { return mk3(q.x, q.y, q.z); }

// real4.cpp:139-150 (non-AVX branch)
HD q4 qmul(const q4& a, const q4& b) {
    q4 t;
    t.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    t.x = a.w * b.x + a.x * b.w - a.z * b.y + a.y * b.z;
    t.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
    t.z = a.w * b.z + a.z * b.w - a.y * b.x + a.x * b.y;
    return t;
}

// real4.cpp:158-165:  t = 2 (q_v x v);  v' = v + w t + q_v x t
HD d3 rotate(const d3& v, const q4& q) {
    d3 qv = qvec(q);
    d3 t = 2.0 * cross3(qv, v);
    return v + q.w * t + cross3(qv, t);
}
HD d3 rotate_t(const d3& v, const q4& q) { return rotate(v, conj(q)); }

// real4.cpp:168-188: |R(q)| v
HD d3 abs_rotate(const q4& q, const d3& v) {
    double e0e0 = q.w * q.w, e1e1 = q.x * q.x, e2e2 = q.y * q.y, e3e3 = q.z * q.z;
    double e0e1 = q.w * q.x, e0e2 = q.w * q.y, e0e3 = q.w * q.z;
    double e1e2 = q.x * q.y, e1e3 = q.x * q.z, e2e3 = q.y * q.z;
    d3 r;
    r.x = fabs((e0e0 + e1e1) * 2 - 1) * v.x + fabs((e1e2 - e0e3) * 2) * v.y + fabs((e1e3 + e0e2) * 2) * v.z;
    r.y = fabs((e1e2 + e0e3) * 2) * v.x + fabs((e0e0 + e2e2) * 2 - 1) * v.y + fabs((e2e3 - e0e1) * 2) * v.z;
    r.z = fabs((e1e3 - e0e2) * 2) * v.x + fabs((e2e3 + e0e1) * 2) * v.y + fabs((e0e0 + e3e3) * 2 - 1) * v.z;
    return r;
}

// MM/utility.h:48-57
HD d3 local_to_parent(const d3& p, const q4& q, const d3& rl) { return p + rotate(rl, q); }
// y = S x for a symmetric 3 x 3 matrix kept as its diagonal (diag[3b..]) and, when the system has bodies with products of
// inertia, its off-diagonal part (off[3b..] = xy, xz, yz; null otherwise).  Row sums in column order, as a row-major
// sparse product forms them.
HD d3 sym3_apply(const double* __restrict__ diag, const double* __restrict__ off, size_t b, const d3& x) {
    const double d0 = diag[3 * b], d1 = diag[3 * b + 1], d2 = diag[3 * b + 2];
    if (!off)
        return mk3(d0 * x.x, d1 * x.y, d2 * x.z);
    const double xy = off[3 * b], xz = off[3 * b + 1], yz = off[3 * b + 2];
    return mk3((d0 * x.x + xy * x.y) + xz * x.z, (xy * x.x + d1 * x.y) + yz * x.z, (xz * x.x + yz * x.y) + d2 * x.z);
}

HD d3 parent_to_local(const d3& p, const q4& q, const d3& rp) { return rotate_t(rp - p, q); }

// columns of R(q): AMatU / AMatV / AMatW (real4.cpp:200-250)
HD d3 amat_u(const q4& q) {
    return mk3((q.w * q.w + q.x * q.x) * 2 - 1, (q.x * q.y + q.w * q.z) * 2, (q.x * q.z - q.w * q.y) * 2);
}
HD d3 amat_v(const q4& q) {
    return mk3((q.x * q.y - q.w * q.z) * 2, (q.w * q.w + q.y * q.y) * 2 - 1, (q.y * q.z + q.w * q.x) * 2);
}
HD d3 amat_w(const q4& q) {
    return mk3((q.x * q.z + q.w * q.y) * 2, (q.y * q.z - q.w * q.x) * 2, (q.w * q.w + q.z * q.z) * 2 - 1);
}

// MM/real.h:141-150
HD double clampd(double x, double lo, double hi) {
    if (lo > hi) {
        double t = lo;
        lo = hi;
        hi = t;
    }
    return x < lo ? lo : (x > hi ? hi : x);
}
