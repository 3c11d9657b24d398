// ptb_scene_dev.cuh — device-side view of a flattened scene (all pointers are device memory).
#pragma once
#include "ptb200.h"
#include "ptb_csg.cuh"

namespace ptb
{

struct DevImage
{
    cudaTextureObject_t tex; // float4 2D array, point filter, border (= black outside) addressing
    cudaTextureObject_t texLinear; // same array, bilinear (PTB_FLAG_BILINEAR extra)
    int w, h;
};

// A texture chain folded into one 80-byte record when the scene is created (five independent 16-byte loads instead of
// the texops -> texdata -> texops ... chain of dependent ones, and no dispatch on op codes).  Chains of the form
// [affine] [mirror-ball | spherical map] image-or-sky-box leaf, filtered by at most one MultiplyTexture and one
// LogTexture in either order — every texture of the reference's scenes (src/test.cpp:93-105) — are folded; any
// other chain has flags = 0 and runs through the interpreter (ptb_texture.cuh).  Indexed like texops (by chain root).
#define PTB_TB_VALID 1u
#define PTB_TB_AFFINE 2u
#define PTB_TB_MIRRORBALL 4u
#define PTB_TB_SPHERICAL 8u
#define PTB_TB_SKYBOX 16u
#define PTB_TB_ALPHA 32u
#define PTB_TB_MUL 64u
#define PTB_TB_LOG 128u
#define PTB_TB_LOG_FIRST 256u /* the LogTexture lies below the MultiplyTexture: log first, then multiply */
struct DevTexBaked
{
    uint32_t flags;
    int image; // the leaf's image (the first of six for a sky box)
    float pad0, pad1;
    float4 r0, r1, r2; // rows of the TransformedTexture's matrix
    float4 mul;        // the MultiplyTexture's factor
};

#define PTB_MAT_TERMINAL (1 << 8) /* constMask: a path ends on this material after the emission term (no refraction, zero reflectance) */
struct DevMaterial
{
    int tex[5];  // PTB_MAT_* texture roots
    float ior;   // relative ior when leaving the material (include/path-trace.h:93)
    float iorIn; // float(1.0 / ior): relative ior when entering (include/path-trace.h:79)
    int constMask; // bit k: texture k is a bare ColorTexture whose value is folded into the fields below; PTB_MAT_TERMINAL
    // folded constants (rgb, getFloat = mean rgb): one 16-byte load instead of texops -> texdata chains
    float4 emissive, reflect, transmit;
    float scatterF, trcF, pad0, pad1;
};

struct DevScene
{
    const uint4 *program;     // compiled CSG stream (see ptb_csg.cuh), programWords 16-byte words
    int programWords;
    const float4 *primGeo;    // [n_prims] sphere: c.xyz, r^2; plane: n.xyz, d
    const int4 *primMeta;     // [n_prims] type, material, chain first, chain count
    const int *chainXf;       // xform indices, outermost first
    const float4 *xforms;     // [n_xforms][6]: rows of m, rows of inv
    const DevMaterial *materials;
    const int2 *texops;       // kind, arg
    const float *texdata;
    const DevTexBaked *texBaked; // [n_texops]
    const DevImage *images;
    int nPrims, nMaterials, nTexops, nImages;
    const uint4 *fastProgram; // the same program compiled for the fast evaluator (ptb_csg_fast.cuh)
    int fastWords;
    int maxStack;             // deepest list stack the program reaches
    int poolBound;            // worst-case pool slots (incl. merge scratch)
};

// Normal of primitive `tag>>1` at parameter t of the WORLD ray (o, d), brought back through its
// TransformedObject chain and sign-flipped per the tag: what the reference carries in
// Span::startNormal / endNormal (src/sphere.cpp:47-48, src/plane.cpp:29-30,
// include/span.h:122-127 once per wrapper, include/span.h:100-113 flips).
template <bool EX>
__device__ __forceinline__ V3 boundaryNormal(const DevScene &sc, uint32_t tag, float t, V3 o, V3 d)
{
    typedef Ar<EX> A;
    int prim = (int)(tag >> 1);
    int4 meta = __ldg(&sc.primMeta[prim]);
    float4 g = __ldg(&sc.primGeo[prim]);
    V3 n;
    if(meta.x == PTB_PRIM_SPHERE)
    {
        for(int k = 0; k < meta.w; k++)
        {
            const float4 *x = sc.xforms + 6 * (size_t)__ldg(&sc.chainXf[meta.z + k]);
            float4 r0 = __ldg(x), r1 = __ldg(x + 1), r2 = __ldg(x + 2);
            o = A::apply(r0, r1, r2, o);
            d = A::applyNoTranslate(r0, r1, r2, d);
        }
        n = A::normalize(A::sub(A::point(o, d, t), mk(g.x, g.y, g.z)));
    }
    else
        n = A::normalize(mk(g.x, g.y, g.z));
    for(int k = meta.w - 1; k >= 0; k--)
    {
        const float4 *x = sc.xforms + 6 * (size_t)__ldg(&sc.chainXf[meta.z + k]) + 3;
        n = A::normalize(A::applyNoTranslate(__ldg(x), __ldg(x + 1), __ldg(x + 2), n));
    }
    return (tag & 1u) ? neg(n) : n;
}

// ---- path state -----------------------------------------------------------------------------
struct PathState
{
    V3 o, d;         // current ray (direction NOT normalised)
    V3 thr;          // product of branch weights so far
    float strength;  // the reference's `strength` argument
    uint32_t pixel;  // y * width + x
    uint32_t sample; // global camera-sample index
    uint32_t depth;  // the reference's `depth` argument for this ray
    uint32_t rngBlock;
};

struct Hit
{
    float t;
    V3 n;      // faces against the ray
    int mat;
    float ior; // relative ior across the boundary
};

} // namespace ptb
