// oracle/dxm_shim/DirectXMath.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A from-scratch, minimal stand-in for the parts of DirectXMath 3.11 (Windows SDK 10.0.16299.0, the version the
// reference pins in common-lib.vcxproj:17) that the reference's camera/light/material/quasi-random/ray-tracing/
// texture .cpp files call.  It exists so those files can be compiled UNMODIFIED, in place, by oracle/Makefile and
// used to check the scalar restatement in rt_oracle.cpp function by function.
//
// Every function follows the operation order of DirectXMath's _XM_SSE_INTRINSICS_ code path (the reference builds
// x64 /arch:SSE2: no FMA, no SSE4 dot products).  DirectXMath's source is NOT available in this container: the
// sequences below are restated from memory of the SSE2 paths, which is why DESIGN.md says "parity unpinned at the
// DirectXMath boundary".  Only what the six reference files use is provided.
#pragma once
#include <emmintrin.h>
#include <xmmintrin.h>

#include <cmath>
#include <cstdint>
#include <cstring>

namespace DirectX {

constexpr float XM_PI = 3.141592654f;

struct XMFLOAT2 { float x, y; XMFLOAT2() = default; constexpr XMFLOAT2(float _x, float _y) : x(_x), y(_y) {} };
struct XMFLOAT3 { float x, y, z; XMFLOAT3() = default; constexpr XMFLOAT3(float _x, float _y, float _z) : x(_x), y(_y), z(_z) {} };

// On MSVC XMVECTOR is __m128 itself (a union there, so brace-init and operator overloads are legal).  GCC's __m128 is
// a builtin vector type, so wrap it: the wrapper keeps MSVC's "{x, y, z}" initialisation (missing lanes are zero).
struct alignas(16) XMVECTOR {
    __m128 v;
    XMVECTOR() = default;
    XMVECTOR(__m128 m) : v(m) {}
    XMVECTOR(float x, float y, float z = 0.f, float w = 0.f) : v(_mm_set_ps(w, z, y, x)) {}
    operator __m128() const { return v; }
};
typedef const XMVECTOR FXMVECTOR;
typedef const XMVECTOR GXMVECTOR;

struct alignas(16) XMVECTORF32 {
    float f[4];
    operator XMVECTOR() const { return XMVECTOR(_mm_load_ps(f)); }
};

// ---- arithmetic operators (DirectXMathVector.inl: thin wrappers over *_ps) ----
inline XMVECTOR operator+(FXMVECTOR a, FXMVECTOR b) { return _mm_add_ps(a.v, b.v); }
inline XMVECTOR operator-(FXMVECTOR a, FXMVECTOR b) { return _mm_sub_ps(a.v, b.v); }
inline XMVECTOR operator*(FXMVECTOR a, FXMVECTOR b) { return _mm_mul_ps(a.v, b.v); }
inline XMVECTOR operator/(FXMVECTOR a, FXMVECTOR b) { return _mm_div_ps(a.v, b.v); }
inline XMVECTOR operator*(FXMVECTOR a, float s) { return _mm_mul_ps(_mm_set_ps1(s), a.v); }      // XMVectorScale
inline XMVECTOR operator*(float s, FXMVECTOR a) { return _mm_mul_ps(_mm_set_ps1(s), a.v); }
inline XMVECTOR operator/(FXMVECTOR a, float s) { return _mm_div_ps(a.v, _mm_set_ps1(s)); }      // XMVectorDivide(V, Replicate(S))
inline XMVECTOR operator-(FXMVECTOR a) { return _mm_sub_ps(_mm_setzero_ps(), a.v); }            // XMVectorNegate: 0 - v
inline XMVECTOR operator+(FXMVECTOR a) { return a; }
inline XMVECTOR& operator+=(XMVECTOR& a, FXMVECTOR b) { a.v = _mm_add_ps(a.v, b.v); return a; }
inline XMVECTOR& operator-=(XMVECTOR& a, FXMVECTOR b) { a.v = _mm_sub_ps(a.v, b.v); return a; }
inline XMVECTOR& operator*=(XMVECTOR& a, FXMVECTOR b) { a.v = _mm_mul_ps(a.v, b.v); return a; }

// ---- construction / access ----
inline XMVECTOR XMVectorZero() { return _mm_setzero_ps(); }
inline XMVECTOR XMVectorReplicate(float s) { return _mm_set_ps1(s); }
inline XMVECTOR XMVectorSet(float x, float y, float z, float w) { return _mm_set_ps(w, z, y, x); }
inline float XMVectorGetX(FXMVECTOR a) { return _mm_cvtss_f32(a.v); }
inline float XMVectorGetY(FXMVECTOR a) { return _mm_cvtss_f32(_mm_shuffle_ps(a.v, a.v, _MM_SHUFFLE(1, 1, 1, 1))); }
inline float XMVectorGetZ(FXMVECTOR a) { return _mm_cvtss_f32(_mm_shuffle_ps(a.v, a.v, _MM_SHUFFLE(2, 2, 2, 2))); }
inline float XMVectorGetW(FXMVECTOR a) { return _mm_cvtss_f32(_mm_shuffle_ps(a.v, a.v, _MM_SHUFFLE(3, 3, 3, 3))); }
inline void XMStoreFloat3(XMFLOAT3* dst, FXMVECTOR a) { dst->x = XMVectorGetX(a); dst->y = XMVectorGetY(a); dst->z = XMVectorGetZ(a); }
inline void XMStoreFloat2(XMFLOAT2* dst, FXMVECTOR a) { dst->x = XMVectorGetX(a); dst->y = XMVectorGetY(a); }
inline XMVECTOR XMLoadFloat3(const XMFLOAT3* src) { return _mm_set_ps(0.f, src->z, src->y, src->x); }

// ---- per-lane functions ----
inline XMVECTOR XMVectorSqrt(FXMVECTOR a) { return _mm_sqrt_ps(a.v); }
inline XMVECTOR XMVectorReciprocal(FXMVECTOR a) { return _mm_div_ps(_mm_set_ps1(1.0f), a.v); }
inline XMVECTOR XMVectorReciprocalEst(FXMVECTOR a) { return _mm_rcp_ps(a.v); }                   // rcpps, ~12 bit
inline XMVECTOR XMVectorMin(FXMVECTOR a, FXMVECTOR b) { return _mm_min_ps(a.v, b.v); }
inline XMVECTOR XMVectorMax(FXMVECTOR a, FXMVECTOR b) { return _mm_max_ps(a.v, b.v); }
inline XMVECTOR XMVectorSaturate(FXMVECTOR a) {                                                   // max(v,0) then min(.,1)
    return _mm_min_ps(_mm_max_ps(a.v, _mm_setzero_ps()), _mm_set_ps1(1.0f));
}
inline XMVECTOR XMVectorMultiplyAdd(FXMVECTOR a, FXMVECTOR b, FXMVECTOR c) {                     // mul then add (no FMA3 in an SSE2 build)
    return _mm_add_ps(_mm_mul_ps(a.v, b.v), c.v);
}
inline XMVECTOR XMVectorPow(FXMVECTOR a, FXMVECTOR b) {                                          // SSE path: powf lane by lane
    alignas(16) float x[4], y[4];
    _mm_store_ps(x, a.v); _mm_store_ps(y, b.v);
    return _mm_setr_ps(powf(x[0], y[0]), powf(x[1], y[1]), powf(x[2], y[2]), powf(x[3], y[3]));
}

// ---- comparisons ----
inline bool XMVector3Greater(FXMVECTOR a, FXMVECTOR b) { return (_mm_movemask_ps(_mm_cmpgt_ps(a.v, b.v)) & 7) == 7; }
inline bool XMVector3Less(FXMVECTOR a, FXMVECTOR b) { return (_mm_movemask_ps(_mm_cmplt_ps(a.v, b.v)) & 7) == 7; }
inline bool XMVector3NotEqual(FXMVECTOR a, FXMVECTOR b) { return (_mm_movemask_ps(_mm_cmpeq_ps(a.v, b.v)) & 7) != 7; }
constexpr uint32_t XM_CRMASK_CR6TRUE = 0x00000080, XM_CRMASK_CR6FALSE = 0x00000020;
inline XMVECTOR XMVectorGreaterR(uint32_t* cr, FXMVECTOR a, FXMVECTOR b) {                       // all FOUR lanes feed the record
    const __m128 m = _mm_cmpgt_ps(a.v, b.v);
    const int bits = _mm_movemask_ps(m);
    *cr = bits == 0xf ? XM_CRMASK_CR6TRUE : (bits == 0 ? XM_CRMASK_CR6FALSE : 0u);
    return m;
}
inline bool XMComparisonAnyTrue(uint32_t cr) { return (cr & XM_CRMASK_CR6FALSE) != XM_CRMASK_CR6FALSE; }

// ---- 3D vector operations (SSE2 sequences) ----
inline XMVECTOR XMVector3Dot(FXMVECTOR a, FXMVECTOR b) {                                         // (x + y) + z, splatted
    __m128 dot = _mm_mul_ps(a.v, b.v);
    __m128 t = _mm_shuffle_ps(dot, dot, _MM_SHUFFLE(2, 1, 2, 1));
    dot = _mm_add_ss(dot, t);
    t = _mm_shuffle_ps(t, t, _MM_SHUFFLE(1, 1, 1, 1));
    dot = _mm_add_ss(dot, t);
    return _mm_shuffle_ps(dot, dot, _MM_SHUFFLE(0, 0, 0, 0));
}
inline XMVECTOR XMVector3Cross(FXMVECTOR a, FXMVECTOR b) {
    __m128 t1 = _mm_shuffle_ps(a.v, a.v, _MM_SHUFFLE(3, 0, 2, 1));   // y z x
    __m128 t2 = _mm_shuffle_ps(b.v, b.v, _MM_SHUFFLE(3, 1, 0, 2));   // z x y
    __m128 r = _mm_mul_ps(t1, t2);
    t1 = _mm_shuffle_ps(t1, t1, _MM_SHUFFLE(3, 0, 2, 1));            // z x y
    t2 = _mm_shuffle_ps(t2, t2, _MM_SHUFFLE(3, 1, 0, 2));            // y z x
    r = _mm_sub_ps(r, _mm_mul_ps(t1, t2));
    const __m128 mask3 = _mm_castsi128_ps(_mm_set_epi32(0, -1, -1, -1));
    return _mm_and_ps(r, mask3);
}
inline XMVECTOR XMVector3LengthSq(FXMVECTOR a) { return XMVector3Dot(a, a); }
inline XMVECTOR XMVector3Length(FXMVECTOR a) { return _mm_sqrt_ps(XMVector3Dot(a, a).v); }
inline XMVECTOR XMVector3Normalize(FXMVECTOR a) {                                                // v / sqrt(dot); 0 -> 0; inf -> NaN
    const __m128 len_sq = XMVector3Dot(a, a).v;
    __m128 len = _mm_sqrt_ps(len_sq);
    const __m128 nonzero = _mm_cmpneq_ps(_mm_setzero_ps(), len);
    const __m128 finite = _mm_cmpneq_ps(len_sq, _mm_set_ps1(INFINITY));
    __m128 r = _mm_div_ps(a.v, len);
    r = _mm_and_ps(r, nonzero);
    const __m128 qnan = _mm_castsi128_ps(_mm_set1_epi32(0x7FC00000));
    return _mm_or_ps(_mm_andnot_ps(finite, qnan), _mm_and_ps(r, finite));
}
inline XMVECTOR XMVector3Reflect(FXMVECTOR incident, FXMVECTOR normal) {                         // I - (2 * dot(I,N)) * N
    __m128 r = XMVector3Dot(incident, normal).v;
    r = _mm_add_ps(r, r);
    return _mm_sub_ps(incident.v, _mm_mul_ps(r, normal.v));
}
inline XMVECTOR XMVector3RefractV(FXMVECTOR incident, FXMVECTOR normal, FXMVECTOR index) {
    const __m128 one = _mm_set_ps1(1.0f);
    const __m128 idn = XMVector3Dot(incident, normal).v;
    __m128 r = _mm_mul_ps(idn, idn);
    r = _mm_sub_ps(one, r);
    r = _mm_mul_ps(r, index.v);
    r = _mm_mul_ps(r, index.v);
    r = _mm_sub_ps(one, r);
    if (_mm_movemask_ps(_mm_cmple_ps(r, _mm_setzero_ps())) == 0xf) return _mm_setzero_ps();       // total internal reflection
    r = _mm_sqrt_ps(r);
    __m128 res = _mm_mul_ps(index.v, idn);
    r = _mm_add_ps(r, res);
    res = _mm_mul_ps(index.v, incident.v);
    r = _mm_mul_ps(r, normal.v);
    return _mm_sub_ps(res, r);
}
inline XMVECTOR XMFresnelTerm(FXMVECTOR cos_incident, FXMVECTOR index) {
    const __m128 one = _mm_set_ps1(1.0f), c = cos_incident.v;
    __m128 g = _mm_mul_ps(index.v, index.v);
    __m128 t = _mm_mul_ps(c, c);
    g = _mm_sub_ps(g, one);
    t = _mm_add_ps(t, g);
    g = _mm_max_ps(_mm_sub_ps(_mm_setzero_ps(), t), t);              // abs
    g = _mm_sqrt_ps(g);
    __m128 gac = _mm_add_ps(g, c), gsc = _mm_sub_ps(g, c);
    __m128 res = _mm_mul_ps(gsc, gsc);
    t = _mm_mul_ps(gac, gac);
    res = _mm_mul_ps(res, _mm_set_ps1(0.5f));
    res = _mm_div_ps(res, t);
    gac = _mm_mul_ps(gac, c);
    gsc = _mm_mul_ps(gsc, c);
    gac = _mm_sub_ps(gac, one);
    gsc = _mm_add_ps(gsc, one);
    gac = _mm_mul_ps(gac, gac);
    gsc = _mm_mul_ps(gsc, gsc);
    gac = _mm_div_ps(gac, gsc);
    gac = _mm_add_ps(gac, one);
    res = _mm_mul_ps(res, gac);
    return _mm_min_ps(_mm_max_ps(res, _mm_setzero_ps()), one);
}

// ---- DirectXPackedVector: XMCOLOR (0xAARRGGBB) ----
namespace PackedVector {
struct XMCOLOR {
    uint32_t c;
    XMCOLOR() = default;
    constexpr XMCOLOR(uint32_t color) : c(color) {}
    XMCOLOR(float r, float g, float b, float a);
};
inline void XMStoreColor(XMCOLOR* dst, FXMVECTOR v) {                                            // saturate, *255, cvtps2dq (nearest-even)
    __m128 r = _mm_max_ps(v.v, _mm_setzero_ps());
    r = _mm_min_ps(r, _mm_set_ps1(1.0f));
    r = _mm_mul_ps(r, _mm_set_ps1(255.0f));
    alignas(16) int32_t k[4];
    _mm_store_si128(reinterpret_cast<__m128i*>(k), _mm_cvtps_epi32(r));
    dst->c = (static_cast<uint32_t>(k[3]) << 24) | (static_cast<uint32_t>(k[0]) << 16) | (static_cast<uint32_t>(k[1]) << 8) | static_cast<uint32_t>(k[2]);
}
inline XMCOLOR::XMCOLOR(float r, float g, float b, float a) { XMStoreColor(this, XMVectorSet(r, g, b, a)); }
inline XMVECTOR XMLoadColor(const XMCOLOR* src) {
    // mask each channel in place, convert, then scale by 1/(255 * 2^shift): equals float(k) * (1.0f/255.0f) bit for bit
    const __m128i vi = _mm_and_si128(_mm_set1_epi32(static_cast<int>(src->c)), _mm_set_epi32(static_cast<int>(0xFF000000u), 0x000000FF, 0x0000FF00, 0x00FF0000));
    const __m128i flipped = _mm_xor_si128(vi, _mm_set_epi32(static_cast<int>(0x80000000u), 0, 0, 0));
    __m128 f = _mm_cvtepi32_ps(flipped);
    f = _mm_add_ps(f, _mm_set_ps(32768.0f * 65536.0f, 0.f, 0.f, 0.f));
    return _mm_mul_ps(f, _mm_set_ps(1.0f / (255.0f * 65536.0f * 256.0f), 1.0f / 255.0f, 1.0f / (255.0f * 256.0f), 1.0f / (255.0f * 65536.0f)));
}
}  // namespace PackedVector

}  // namespace DirectX
