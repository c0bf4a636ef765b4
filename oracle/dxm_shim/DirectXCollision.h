// oracle/dxm_shim/DirectXCollision.h -- TEST INFRASTRUCTURE.  BoundingBox as the reference's AABB / BvhNode use it
// (ray-tracing.cpp:96-105,159-161): constructor, ray test and CreateMerged, restated from memory of DirectXCollision.inl.
#pragma once
#include <cfloat>

#include "DirectXMath.h"

namespace DirectX {

struct BoundingBox {
    XMFLOAT3 Center;
    XMFLOAT3 Extents;
    BoundingBox() : Center(0.f, 0.f, 0.f), Extents(1.f, 1.f, 1.f) {}
    constexpr BoundingBox(const XMFLOAT3& center, const XMFLOAT3& extents) : Center(center), Extents(extents) {}

    // Slab test against the three axis-aligned slabs; Direction is assumed unit length.  No upper bound on t.
    bool Intersects(FXMVECTOR Origin, FXMVECTOR Direction, float& Dist) const {
        const __m128 center = XMLoadFloat3(&Center).v, extents = XMLoadFloat3(&Extents).v;
        const __m128 to_origin = _mm_sub_ps(center, Origin.v);                 // box centre relative to the ray origin
        const __m128 dir = Direction.v;
        const __m128 abs_dir = _mm_max_ps(_mm_sub_ps(_mm_setzero_ps(), dir), dir);
        const __m128 parallel = _mm_cmple_ps(abs_dir, _mm_set_ps1(1e-20f));    // g_RayEpsilon
        const __m128 inv = _mm_div_ps(_mm_set_ps1(1.0f), dir);                 // XMVectorReciprocal
        const __m128 t1 = _mm_mul_ps(_mm_sub_ps(to_origin, extents), inv);
        const __m128 t2 = _mm_mul_ps(_mm_add_ps(to_origin, extents), inv);
        const __m128 fmin = _mm_set_ps1(-FLT_MAX), fmax = _mm_set_ps1(FLT_MAX);
        __m128 t_min = _mm_or_ps(_mm_and_ps(parallel, fmin), _mm_andnot_ps(parallel, _mm_min_ps(t1, t2)));
        __m128 t_max = _mm_or_ps(_mm_and_ps(parallel, fmax), _mm_andnot_ps(parallel, _mm_max_ps(t1, t2)));
        // reduce over x, y, z into lane x
        t_min = _mm_max_ps(t_min, _mm_shuffle_ps(t_min, t_min, _MM_SHUFFLE(1, 1, 1, 1)));
        t_min = _mm_max_ps(t_min, _mm_shuffle_ps(t_min, t_min, _MM_SHUFFLE(2, 2, 2, 2)));
        t_max = _mm_min_ps(t_max, _mm_shuffle_ps(t_max, t_max, _MM_SHUFFLE(1, 1, 1, 1)));
        t_max = _mm_min_ps(t_max, _mm_shuffle_ps(t_max, t_max, _MM_SHUFFLE(2, 2, 2, 2)));
        const __m128 tmin_x = _mm_shuffle_ps(t_min, t_min, _MM_SHUFFLE(0, 0, 0, 0));
        const __m128 tmax_x = _mm_shuffle_ps(t_max, t_max, _MM_SHUFFLE(0, 0, 0, 0));
        __m128 miss = _mm_cmpgt_ps(tmin_x, tmax_x);                            // slabs do not overlap
        miss = _mm_or_ps(miss, _mm_cmplt_ps(tmax_x, _mm_setzero_ps()));        // box entirely behind the origin
        const __m128 inside = _mm_and_ps(_mm_cmple_ps(to_origin, extents), _mm_cmple_ps(_mm_sub_ps(_mm_setzero_ps(), extents), to_origin));
        miss = _mm_or_ps(miss, _mm_andnot_ps(inside, parallel));               // parallel to a slab and outside it
        if ((_mm_movemask_ps(miss) & 7) == 0) { Dist = _mm_cvtss_f32(tmin_x); return true; }
        Dist = 0.f;
        return false;
    }

    static void CreateMerged(BoundingBox& Out, const BoundingBox& b1, const BoundingBox& b2) {
        const __m128 c1 = XMLoadFloat3(&b1.Center).v, e1 = XMLoadFloat3(&b1.Extents).v;
        const __m128 c2 = XMLoadFloat3(&b2.Center).v, e2 = XMLoadFloat3(&b2.Extents).v;
        const __m128 mn = _mm_min_ps(_mm_sub_ps(c1, e1), _mm_sub_ps(c2, e2));
        const __m128 mx = _mm_max_ps(_mm_add_ps(c1, e1), _mm_add_ps(c2, e2));
        XMStoreFloat3(&Out.Center, XMVECTOR(_mm_mul_ps(_mm_add_ps(mn, mx), _mm_set_ps1(0.5f))));
        XMStoreFloat3(&Out.Extents, XMVECTOR(_mm_mul_ps(_mm_sub_ps(mx, mn), _mm_set_ps1(0.5f))));
    }
};

}  // namespace DirectX
