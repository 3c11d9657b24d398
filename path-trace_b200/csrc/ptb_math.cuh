// ptb_math.cuh — float3 algebra of the render path under two arithmetic policies.
//
//   Exact (EX = true) : every operation is a single correctly-rounded binary32 op in the
//                       reference's order — __fmul_rn/__fadd_rn are never contracted into FMA,
//                       __fdiv_rn/__fsqrt_rn are IEEE — so t, normals and hit points are
//                       bit-identical to the reference built for x86-64 without FMA
//                       (reference include/vector3d.h:77-120).
//   Fast  (EX = false): plain operators (ptxas contracts a*b+c into FFMA), MUFU-based
//                       reciprocal / rsqrt; relative error ~1e-7..1e-6, inside the 1e-5 gate of
//                       BASELINE.json.  This is the policy the renderer runs.
#pragma once
#if defined(PTB_HOST_EMU)
#include "ptb_host_emu.h" /* tests only: the evaluators compiled as host code, one lane at a time */
#elif !defined(__CUDACC_RTC__)
#include <cuda_runtime.h>
#include <stdint.h>
#endif

namespace ptb
{

#define PTB_EPS 1e-3f        /* reference include/misc.h:7 */
#define PTB_MAX_VALUE 1e20f  /* reference include/misc.h:8 */

struct V3
{
    float x, y, z;
};
__host__ __device__ __forceinline__ V3 mk(float x, float y, float z)
{
    V3 v;
    v.x = x, v.y = y, v.z = z;
    return v;
}
__device__ __forceinline__ V3 neg(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ bool isZero(V3 a) { return a.x == 0.0f && a.y == 0.0f && a.z == 0.0f; }

template <bool EX>
struct Ar
{
    static __device__ __forceinline__ float mul(float a, float b) { return EX ? __fmul_rn(a, b) : a * b; }
    static __device__ __forceinline__ float add(float a, float b) { return EX ? __fadd_rn(a, b) : a + b; }
    static __device__ __forceinline__ float sub(float a, float b) { return EX ? __fsub_rn(a, b) : a - b; }
    static __device__ __forceinline__ float div(float a, float b) { return EX ? __fdiv_rn(a, b) : __fdividef(a, b); }
    static __device__ __forceinline__ float sqrt(float a)
    {
        if(EX)
            return __fsqrt_rn(a);
#ifdef PTB_HOST_EMU
        return sqrtf(a);
#else
        float r;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
        return r;
#endif
    }
    static __device__ __forceinline__ V3 add(V3 a, V3 b) { return mk(add(a.x, b.x), add(a.y, b.y), add(a.z, b.z)); }
    static __device__ __forceinline__ V3 sub(V3 a, V3 b) { return mk(sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)); }
    static __device__ __forceinline__ V3 mul(V3 a, V3 b) { return mk(mul(a.x, b.x), mul(a.y, b.y), mul(a.z, b.z)); }
    static __device__ __forceinline__ V3 scale(V3 a, float s) { return mk(mul(a.x, s), mul(a.y, s), mul(a.z, s)); }
    // vector3d.h:102-106: the three products, then (x + y) + z
    static __device__ __forceinline__ float dot(V3 a, V3 b) { return add(add(mul(a.x, b.x), mul(a.y, b.y)), mul(a.z, b.z)); }
    static __device__ __forceinline__ float length(V3 a) { return sqrt(dot(a, a)); }
    // vector3d.h:115-120: v / |v| by three divisions, |v| == 0 -> divide by 1
    static __device__ __forceinline__ V3 normalize(V3 a)
    {
        if(EX)
        {
            float m = __fsqrt_rn(dot(a, a));
            if(m == 0.0f)
                m = 1.0f;
            return mk(__fdiv_rn(a.x, m), __fdiv_rn(a.y, m), __fdiv_rn(a.z, m));
        }
        float d2 = dot(a, a);
        float inv = d2 == 0.0f ? 1.0f : rsqrtf(d2);
        return mk(a.x * inv, a.y * inv, a.z * inv);
    }
    // ray.getPoint(t) = origin + t * dir (include/ray.h:19-22)
    static __device__ __forceinline__ V3 point(V3 o, V3 d, float t) { return mk(add(o.x, mul(t, d.x)), add(o.y, mul(t, d.y)), add(o.z, mul(t, d.z))); }
    // Matrix::apply / applyNoTranslate (include/transform.h:408-421); r0..r2 are the three
    // float4 rows (x00,x10,x20,x30), (x01,...), (x02,...) of the 12-float layout.
    static __device__ __forceinline__ float row3(float4 r, V3 v) { return add(add(mul(v.x, r.x), mul(v.y, r.y)), mul(v.z, r.z)); }
    static __device__ __forceinline__ V3 apply(float4 r0, float4 r1, float4 r2, V3 v) { return mk(add(row3(r0, v), r0.w), add(row3(r1, v), r1.w), add(row3(r2, v), r2.w)); }
    static __device__ __forceinline__ V3 applyNoTranslate(float4 r0, float4 r1, float4 r2, V3 v) { return mk(row3(r0, v), row3(r1, v), row3(r2, v)); }
};

// ---- counter-based RNG: Philox4x32-10 (Salmon et al., SC'11) -------------------------------
// Replaces the reference's shared 64-bit LCG (include/path-trace.h:21-54, a data race between
// its worker threads).  Stream = (pixel, global sample index); parity with the reference is
// statistical.
struct Philox
{
    uint32_t key0, key1; // seed
    uint32_t c0, c1;     // pixel, global sample index
    uint32_t block;      // next counter block
    uint4 buf;
    uint32_t have;       // unread words in buf
};

__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1)
{
#pragma unroll
    for(int r = 0; r < 10; r++)
    {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0, c1 = n1, c2 = n2, c3 = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ void rngInit(Philox &r, uint64_t seed, uint32_t pixel, uint32_t sample, uint32_t block)
{
    r.key0 = (uint32_t)seed;
    r.key1 = (uint32_t)(seed >> 32);
    r.c0 = pixel;
    r.c1 = sample;
    r.block = block;
    r.have = 0;
}
// Sub-pixel jitter of a camera sample: two 32-bit words from an integer hash of (seed, pixel, sample)
// (the 'lowbias32' finaliser twice).  The primary trace kernel and the primary shade kernel both
// regenerate the camera ray, so this runs twice per path; a Philox block here was 80 of the ~1100
// instructions of a camera ray.  Shading keeps Philox (blocks 0, 1, ... of the same stream).
__device__ __forceinline__ uint32_t mix32(uint32_t h)
{
    h ^= h >> 16, h *= 0x7FEB352Du;
    h ^= h >> 15, h *= 0x846CA68Bu;
    return h ^ (h >> 16);
}
__device__ __forceinline__ uint2 cameraJitter(uint64_t seed, uint32_t pixel, uint32_t sample)
{
    const uint32_t a = mix32(pixel ^ (uint32_t)seed) + sample * 0x9E3779B1u;
    const uint32_t x = mix32(a ^ (uint32_t)(seed >> 32));
    return make_uint2(x, mix32(x + 0x85EBCA77u + a));
}
__device__ __forceinline__ uint4 rngNext4(Philox &r) { return philox4x32_10(r.c0, r.c1, r.block++, 0u, r.key0, r.key1); }
// u32 -> closed [0,1]: float(u) / 2^32, the mapping of the reference's
// uniform_real_distribution<float>(0,1) over a 32-bit engine (include/vector3d.h:26-33).
__device__ __forceinline__ float u01(uint32_t u) { return __uint2float_rn(u) * 2.3283064365386963e-10f; }

} // namespace ptb
