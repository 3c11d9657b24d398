// ptb_shade.cuh — one path vertex of traceRay (reference include/path-trace.h:101-163) as a
// NON-branching estimator with the same expectation (SURVEY.md §7 hard part 1):
//   * the reference fires N = (int)(10000*strength*addFactor*scatter) scatter rays, each
//     weighted addFactor/N; we fire ONE, weighted addFactor, and hand it the child strength
//     strength/N*addFactor*factor*|reflect| the reference would (path-trace.h:130-136,160-162),
//     so paths stop at the same vertices (`strength < eps`, path-trace.h:105);
//   * at a dielectric the reference traces the refracted AND the reflected branch
//     (path-trace.h:110-125); we pick the refracted one with probability refractFactor and
//     weight it `transmit`, else the reflected one weighted `factor*reflect` — the
//     expectation is the reference's sum; child strengths are the reference's deterministic ones.
#pragma once
#include "ptb_texture.cuh"

namespace ptb
{

// Vector3D::reflect / refractStrength / refract, reference include/vector3d.h:186-214.
__device__ __forceinline__ V3 vreflect(V3 d, V3 n)
{
    n = Ar<false>::normalize(n);
    float k = 2.f * Ar<false>::dot(d, n);
    return mk(d.x - k * n.x, d.y - k * n.y, d.z - k * n.z);
}
__device__ __forceinline__ float refractStrength(V3 d, float ior, V3 n)
{
    if(ior < PTB_EPS || ior > 1.f / PTB_EPS || isZero(n) || isZero(d))
        return 0.f;
    n = Ar<false>::normalize(n);
    float c = Ar<false>::dot(Ar<false>::normalize(d), n);
    float k = 1.f - ior * ior * (1.f - c * c);
    if(k <= 0.f)
        return 0.f;
    return sqrtf(sqrtf(k));
}
__device__ __forceinline__ V3 vrefract(V3 d, float ior, V3 n)
{
    if(ior < PTB_EPS || ior > 1.f / PTB_EPS || isZero(n) || isZero(d))
        return mk(0.f, 0.f, 0.f);
    n = Ar<false>::normalize(n);
    V3 in = Ar<false>::normalize(d);
    float c = Ar<false>::dot(in, n);
    float k = 1.f - ior * ior * (1.f - c * c);
    if(k < 0.f)
        return mk(0.f, 0.f, 0.f);
    float s = ior * c + sqrtf(k);
    return Ar<false>::normalize(mk(ior * in.x - s * n.x, ior * in.y - s * n.y, ior * in.z - s * n.z));
}

// ---- scatter lobe ---------------------------------------------------------------------------
// Reference (path-trace.h:142-158): d = ball() + (1/sc - 1)*r, resampled until n.d > eps
// (at most 1000 times, else the vertex returns what it has), then normalised.
//
// Literal form: nested rejection loops (cube -> ball, ball -> half-space).
__device__ __forceinline__ bool scatterRejection(Philox &rng, V3 n, V3 r, float sc, V3 &out)
{
    float k = 1.f / sc - 1.f;
    for(int tries = 0; tries < 1000; tries++)
    {
        V3 b;
        float mag;
        do
        { // Vector3D::rand(engine, 1, 0), include/vector3d.h:164-185
            uint4 u = rngNext4(rng);
            b = mk(u01(u.x) * 2.f - 1.f, u01(u.y) * 2.f - 1.f, u01(u.z) * 2.f - 1.f);
            mag = sqrtf(b.x * b.x + b.y * b.y + b.z * b.z);
        } while(mag > 1.f);
        V3 d = mk(b.x + k * r.x, b.y + k * r.y, b.z + k * r.z);
        if(Ar<false>::dot(n, d) > PTB_EPS)
        {
            out = Ar<false>::normalize(d);
            return true;
        }
    }
    return false;
}
// Divergence-free form: the accepted ball point is uniform in {|x|<=1, n.x > c} with
// c = eps - (1/sc-1) n.r, a spherical cap.  Sample its height h along n by inverting the cap's
// volume CDF (h - h^3/3, a cubic solved trigonometrically), then a uniform point on the disc
// at that height.  Same distribution, three uniforms, no loop.  The reference's 1000-try
// give-up is reproduced in expectation: with per-try acceptance p = cap/ball volume the vertex
// is dropped with probability (1-p)^1000.
__device__ __forceinline__ bool scatterCap(Philox &rng, V3 n, V3 r, float sc, V3 &out)
{
    float k = 1.f / sc - 1.f;
    float c = PTB_EPS - k * Ar<false>::dot(n, r);
    if(c >= 1.f)
        return false;
    c = fmaxf(c, -1.f);
    uint4 u = rngNext4(rng);
    float lo = c - c * c * c * (1.f / 3.f);     // G(c),  G(h) = h - h^3/3
    float vol = (2.f / 3.f) - lo;               // G(1) - G(c)  (x pi = cap volume)
    float p = vol * 0.75f;                      // / (4/3): acceptance probability per try
    if(p < 0.03f && u01(u.w) < __expf(1000.f * log1pf(-p)))
        return false;
    float y = lo + u01(u.x) * vol;              // G(h) = y
    y = fminf(fmaxf(y, -2.f / 3.f), 2.f / 3.f);
    float h = 2.f * cosf((acosf(-1.5f * y) + 12.566370614359172f) * (1.f / 3.f));
    h = fminf(fmaxf(h, c), 1.f);
    float rad = sqrtf(fmaxf(0.f, 1.f - h * h) * u01(u.y));
    float sn, cs;
    sincospif(2.f * u01(u.z), &sn, &cs);
    // orthonormal basis around the (unit) normal (branchless, Duff et al. 2017)
    float sg = copysignf(1.f, n.z);
    float a = -1.f / (sg + n.z);
    float bb = n.x * n.y * a;
    V3 t1 = mk(1.f + sg * n.x * n.x * a, sg * bb, -sg * n.x);
    V3 t2 = mk(bb, sg + n.y * n.y * a, -n.y);
    float ex = rad * cs, ey = rad * sn;
    V3 d = mk(h * n.x + ex * t1.x + ey * t2.x + k * r.x, h * n.y + ex * t1.y + ey * t2.y + k * r.y, h * n.z + ex * t1.z + ey * t2.z + k * r.z);
    out = Ar<false>::normalize(d);
    return true;
}


// Shades one hit.  Adds thr*emissive to `radiance`.  Returns true when the path continues
// (ps then holds the next ray), false when it ends here.
__device__ __forceinline__ bool shadeVertex(const DevScene &sc, PathState &ps, const Hit &hit, Philox &rng, uint32_t flags, V3 &radiance)
{
    const bool bilinear = flags & PTB_FLAG_BILINEAR;
    const DevMaterial &m = sc.materials[hit.mat];
    V3 hitPos = Ar<false>::point(ps.o, ps.d, hit.t);
    const int cm = m.constMask; // bare ColorTextures were folded into the material record at scene create
    const bool ends = (cm & PTB_MAT_TERMINAL) || ps.depth == 0u || ps.strength < PTB_EPS; // path-trace.h:105-108; a terminal material's child carries weight 0 (:161)
    // The material's five values at the hit point.  Slots that are not constants are evaluated here, in ONE loop
    // around the texture evaluator (one copy of its code in the kernel); a vertex that ends the path only emits.
    // Textures are pure functions of the position, so evaluating the reflectance before the refraction test
    // (the reference asks for it after, path-trace.h:110-160) changes nothing.
    V3 e = mk(m.emissive.x, m.emissive.y, m.emissive.z), refl = mk(m.reflect.x, m.reflect.y, m.reflect.z), tr = mk(m.transmit.x, m.transmit.y, m.transmit.z);
    float trc = m.trcF, scat = m.scatterF;
#pragma unroll 1
    for(int need = ~cm & (ends ? (1 << PTB_MAT_EMISSIVE) : 31); need; need &= need - 1)
    {
        const int k = __ffs(need) - 1;
        const float4 v = texEval(sc.texBaked, sc.texops, sc.texdata, sc.images, m.tex[k], hitPos.x, hitPos.y, hitPos.z, bilinear);
        if(k == PTB_MAT_EMISSIVE)
            e = mk(v.x, v.y, v.z);
        else if(k == PTB_MAT_REFLECT)
            refl = mk(v.x, v.y, v.z);
        else if(k == PTB_MAT_TRANSMIT)
            tr = mk(v.x, v.y, v.z);
        else if(k == PTB_MAT_SCATTER)
            scat = v.w;
        else
            trc = v.w;
    }
    radiance.x += ps.thr.x * e.x, radiance.y += ps.thr.y * e.y, radiance.z += ps.thr.z * e.z;
    if(ends)
        return false;
    float addFactor = 1.f;
    trc = fmaxf(0.f, fminf(1.f, trc));
    float refractFactor = trc > 0.f ? trc * refractStrength(ps.d, hit.ior, hit.n) : 0.f;
    if(refractFactor > PTB_EPS) // path-trace.h:111-121
    {
        V3 rd = vrefract(ps.d, hit.ior, hit.n);
        if(!isZero(rd))
        {
            float rest = 1.f - refractFactor;
            bool take;
            float w;
            if(rest < PTB_EPS)
                take = true, w = refractFactor; // the reflected branch is dropped at :122-125
            else
            {
                uint4 u = rngNext4(rng);
                take = u01(u.x) < refractFactor;
                w = 1.f; // refractFactor / P(refract)
            }
            if(take)
            {
                ps.strength = ps.strength * refractFactor * addFactor * Ar<false>::length(tr);
                ps.thr = mk(ps.thr.x * tr.x * w, ps.thr.y * tr.y * w, ps.thr.z * tr.z * w);
                ps.o = hitPos, ps.d = rd;
                ps.depth--;
                return !isZero(ps.thr);
            }
            addFactor = rest; // weight addFactor / P(reflect) = 1 is folded in below
        }
    }
    if(addFactor < PTB_EPS)
        return false;
    // diffuse / specular reflection, path-trace.h:127-163
    scat = fmaxf(0.f, fminf(1.f, scat));
    int count = (int)(10000.f * ps.strength * addFactor * scat);
    if(scat <= PTB_EPS || count == 0)
        count = 1;
    if(isZero(refl)) // zero-weight child: contributes exactly 0, not traced
        return false;
    V3 nd = vreflect(ps.d, hit.n);
    if(scat > PTB_EPS)
    {
        bool ok = (flags & PTB_FLAG_REJECTION) ? scatterRejection(rng, hit.n, nd, scat, nd) : scatterCap(rng, hit.n, nd, scat, nd);
        if(!ok)
            return false;
    }
    float factor = 1.f - (1.f - Ar<false>::dot(nd, hit.n)) * scat;
    ps.strength = ps.strength / (float)count * addFactor * factor * Ar<false>::length(refl);
    ps.thr = mk(ps.thr.x * factor * refl.x, ps.thr.y * factor * refl.y, ps.thr.z * factor * refl.z);
    ps.o = hitPos, ps.d = nd;
    ps.depth--;
    return !isZero(ps.thr);
}

} // namespace ptb
