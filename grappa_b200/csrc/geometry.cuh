// Internal-coordinate geometry with analytic first derivatives (host + device, fp32).
// Follows the reference's definitions in src/grappa/models/internal_coordinates.py:
//   distance :134-136   angle :139-154 (atan2(|a x b|, a.b), a = x1-x0, b = x1-x2)
//   dihedral :162-194   (r01 = x1-x0, r21 = x1-x2, r23 = x3-x2, n1 = r01 x r21, n2 = r21 x r23,
//                        phi = atan2((n1 x n2).r21/|r21|, n1.n2)); the randn*1e-5 noise (:178-180) is NOT added:
//                        degenerate geometries are guarded explicitly instead (SURVEY.md §7 "Dihedral noise").
// The derivative vectors returned are d(coordinate)/d(x_i); the reference obtains the same quantities by autograd.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define GHD __host__ __device__ __forceinline__
#else
#define GHD inline
#endif

namespace grappa {

struct V3 {
    float x, y, z;
};

GHD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
GHD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
GHD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
GHD V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
GHD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
GHD V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }

// Device: single-instruction approximations (MUFU, <= 2 ulp), far inside the rtol 1e-4 budget of the path and free of
// the IEEE slow paths that the plain operators carry without --use_fast_math.  Host: exact, for unit tests.
GHD float rsqrt_(float x) {
#ifdef __CUDA_ARCH__
    float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
#else
    return 1.0f / sqrtf(x);
#endif
}
GHD float rcp_(float x) {
#ifdef __CUDA_ARCH__
    float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
#else
    return 1.0f / x;
#endif
}
GHD float sqrt_(float x) {
#ifdef __CUDA_ARCH__
    float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r;
#else
    return sqrtf(x);
#endif
}

constexpr float kTiny = 1e-30f;

// r = |x0 - x1|, g0 = dr/dx0 (dr/dx1 = -g0)
GHD void bond_geom(V3 p0, V3 p1, float& r, V3& g0) {
    V3 d = p0 - p1;
    float r2 = dot(d, d);
    float inv = r2 > kTiny ? rsqrt_(r2) : 0.f;
    r = r2 * inv;
    g0 = inv * d;
}

// atan2(y, x) for y >= 0 (result in [0, pi]), branch-free: atan(t) on t = min/max in [0, 1] by an odd minimax
// polynomial (degree 17, max abs error 1.0e-7 in fp32 evaluation), then the two reflections.  atan2(0, 0) = 0.
GHD float atan2_pos(float y, float x) {
    const float ax = fabsf(x);
    const float lo = fminf(y, ax), hi = fmaxf(y, ax);
    const float t = hi > 0.f ? lo * rcp_(hi) : 0.f;
    const float s = t * t;
    float p = 0.0024567246f;
    p = fmaf(p, s, -0.014401359f);
    p = fmaf(p, s, 0.039781228f);
    p = fmaf(p, s, -0.07234858f);
    p = fmaf(p, s, 0.10498947f);
    p = fmaf(p, s, -0.14161229f);
    p = fmaf(p, s, 0.19985907f);
    p = fmaf(p, s, -0.33332598f);
    p = fmaf(p, s, 0.99999988f);
    p *= t;
    p = y > ax ? 1.57079632679f - p : p;
    return x < 0.f ? 3.14159265359f - p : p;
}

// theta in [0, pi]; g0 = dtheta/dx0, g2 = dtheta/dx2 (dtheta/dx1 = -(g0 + g2))
GHD void angle_geom(V3 p0, V3 p1, V3 p2, float& theta, V3& g0, V3& g2) {
    V3 u = p0 - p1, v = p2 - p1;
    V3 c = cross(u, v);
    float c2 = dot(c, c), uv = dot(u, v);
    float icn = c2 > kTiny ? rsqrt_(c2) : 0.f;
    float cn = c2 * icn;
    theta = atan2_pos(cn, uv);
    float u2 = dot(u, u), v2 = dot(v, v);
    float su = u2 > kTiny ? icn * rcp_(u2) : 0.f;
    float sv = v2 > kTiny ? icn * rcp_(v2) : 0.f;
    g0 = su * cross(u, c);
    g2 = sv * cross(c, v);
}

// cos(phi), sin(phi) of the reference's dihedral and g[i] = dphi/dx_i.
GHD void dihedral_geom(V3 p0, V3 p1, V3 p2, V3 p3, float& cphi, float& sphi, V3 g[4]) {
    V3 F = p0 - p1, G = p1 - p2, H = p3 - p2;
    V3 A = cross(F, G), B = cross(H, G);
    float A2 = dot(A, A), B2 = dot(B, B), G2 = dot(G, G);
    float iG = G2 > kTiny ? rsqrt_(G2) : 0.f;
    float Gn = G2 * iG;
    float x = dot(A, B);
    float y = -dot(A, H) * Gn;      // (A x B).G / |G| with A x B = (F x G) x (H x G) = -(A.H) G
    float rho2 = x * x + y * y;
    float irho = rho2 > kTiny ? rsqrt_(rho2) : 0.f;
    cphi = rho2 > kTiny ? x * irho : 1.f;
    sphi = y * irho;
    float iA2 = A2 > kTiny ? rcp_(A2) : 0.f;
    float iB2 = B2 > kTiny ? rcp_(B2) : 0.f;
    // phi_ref = -phi_BK (Blondel & Karplus 1996 sign convention), hence the signs below.
    float fa = Gn * iA2, fb = Gn * iB2;
    float ta = dot(F, G) * iA2 * iG, tb = dot(H, G) * iB2 * iG;
    g[0] = fa * A;
    g[3] = (-fb) * B;
    V3 s = ta * A - tb * B;
    g[1] = (-fa) * A - s;
    g[2] = fb * B + s;
}

}  // namespace grappa
