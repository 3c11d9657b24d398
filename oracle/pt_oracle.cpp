// pt_oracle.cpp — TEST INFRASTRUCTURE: a CPU restatement of the reference's render hot path
// (programmerjake/path-trace) over the flattened scene of include/ptb200.h.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` leg may
// load this library, and only as the checker.  The product (libptb200.so) never links or calls
// it; it has no CPU fallback.
//
// Parity status: PINNED against the reference itself.  tests/test_oracle_vs_ref.py runs this
// file and the unmodified reference (oracle/_ref/libptref.so, built by oracle/Makefile from
// /root/reference) on the same seeded inputs and requires bit-identical span lists, traceRay
// colours and rendered pixels (same DefaultRandomEngine stream); tests/golden/ holds vectors
// generated from the reference for machines where /root/reference is absent.  The reference
// has no tests or golden vectors of its own.
//
// Every function cites the reference file:line it restates.  Arithmetic is IEEE binary32 in
// the reference's operation order; build with -ffp-contract=off.
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include <algorithm>
#include <vector>

#include "ptb200.h"

namespace
{
const float EPS = 1e-3;        // include/misc.h:7
const float MAX_VALUE = 1e20;  // include/misc.h:8

// ---- include/vector3d.h ----------------------------------------------------------------------
struct V3
{
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float x, float y, float z) : x(x), y(y), z(z) {}
    explicit V3(float v) : x(v), y(v), z(v) {}
};
inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return a * s; }
inline V3 operator/(V3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
inline bool isZero(V3 a) { return a.x == 0 && a.y == 0 && a.z == 0; }
inline float dot(V3 a, V3 b) // vector3d.h:102-106: products first, then (x+y)+z
{
    V3 p = a * b;
    return p.x + p.y + p.z;
}
inline float vabs(V3 v) { return sqrtf(dot(v, v)); }  // vector3d.h:111-114
inline V3 normalize(V3 v)                            // vector3d.h:115-120
{
    float m = vabs(v);
    if(m == 0)
        m = 1;
    return v / m;
}
inline V3 reflect(V3 d, V3 n) // vector3d.h:186-190 — magnitude of d is preserved
{
    n = normalize(n);
    return d - 2 * dot(d, n) * n;
}
float refractStrength(V3 d, float ior, V3 n) // vector3d.h:191-202
{
    if(ior < EPS || ior > 1 / EPS || isZero(n) || isZero(d))
        return 0;
    n = normalize(n);
    V3 in = normalize(d);
    float c = dot(in, n);
    float k = 1 - ior * ior * (1 - c * c);
    if(k <= 0)
        return 0;
    return (float)sqrt(sqrt((double)k)); // unqualified sqrt -> ::sqrt(double) at that point of the header
}
V3 refract(V3 d, float ior, V3 n) // vector3d.h:203-214
{
    if(ior < EPS || ior > 1 / EPS || isZero(n) || isZero(d))
        return V3();
    n = normalize(n);
    V3 in = normalize(d);
    float c = dot(in, n);
    float k = 1 - ior * ior * (1 - c * c);
    if(k < 0)
        return V3();
    return normalize(ior * in - (ior * c + sqrtf(k)) * n);
}

// ---- include/path-trace.h:21-54 — 64-bit LCG, high word out ---------------------------------
struct Lcg
{
    uint64_t v;
    explicit Lcg(unsigned seed = 0) : v(seed ^ 0x12476242) {}
    unsigned next()
    {
        v = 214013 * v + 2531011;
        return (unsigned)(v >> 32);
    }
};
// vector3d.h:15-34 with RG::min()=0, RG::max()=0xFFFFFFFF: closed [lo,hi]
inline float uniform(Lcg &r, float lo, float hi)
{
    float u = (float)r.next();
    u -= (float)(int64_t)0;
    u /= (float)(int64_t)0xFFFFFFFFll;
    u *= hi - lo;
    u += lo;
    return u;
}
V3 randBall(Lcg &r, float max, float min) // vector3d.h:164-185
{
    if(max == 0)
        return V3();
    V3 p;
    float mag;
    do
    {
        p.x = uniform(r, -max, max);
        p.y = uniform(r, -max, max);
        p.z = uniform(r, -max, max);
        mag = vabs(p);
    } while(mag > max || (min > 0 && mag == 0));
    if(min > 0)
        p = p * ((max + mag * (max - min)) / mag);
    return p;
}

// ---- include/transform.h:408-421 over the 12-float layout x00,x10,x20,x30,x01,... -----------
inline V3 apply(const float *m, V3 v)
{
    return V3(v.x * m[0] + v.y * m[1] + v.z * m[2] + m[3], v.x * m[4] + v.y * m[5] + v.z * m[6] + m[7], v.x * m[8] + v.y * m[9] + v.z * m[10] + m[11]);
}
inline V3 applyNoTranslate(const float *m, V3 v)
{
    return V3(v.x * m[0] + v.y * m[1] + v.z * m[2], v.x * m[4] + v.y * m[5] + v.z * m[6], v.x * m[8] + v.y * m[9] + v.z * m[10]);
}

// ---- include/span.h:12-120 ------------------------------------------------------------------
struct Span
{
    float start, end;
    V3 startNormal, endNormal;
    int startPrim, endPrim;
    void copyStartFromStart(const Span &s) { start = s.start, startPrim = s.startPrim, startNormal = s.startNormal; }
    void copyEndFromStart(const Span &s) { end = s.start, endPrim = s.startPrim, endNormal = -s.startNormal; }
    void copyStartFromEnd(const Span &s) { start = s.end, startPrim = s.endPrim, startNormal = -s.endNormal; }
    void copyEndFromEnd(const Span &s) { end = s.end, endPrim = s.endPrim, endNormal = s.endNormal; }
};

struct Scene
{
    const PtbSceneDesc *d;
};

// ---- pull iterators: include/span.h:129-171 -------------------------------------------------
struct Iter
{
    virtual ~Iter() {}
    virtual void init(V3 o, V3 dir) = 0;
    virtual bool atEnd() const = 0;
    virtual void next() = 0;
    virtual const Span &cur() const = 0;
};

// A primitive under its chain of TransformedObject wrappers (include/object.h:31-76): the ray
// goes down through each wrapper's m (include/transform.h:429-432); the span's normals come
// back up through each wrapper's inv, renormalised at every level (include/span.h:122-127).
struct LeafIter : Iter
{
    const PtbSceneDesc *d;
    int prim;
    Span span;
    bool ended;
    LeafIter(const PtbSceneDesc *d, int prim) : d(d), prim(prim), ended(true) {}
    virtual bool atEnd() const { return ended; }
    virtual void next() { ended = true; }
    virtual const Span &cur() const { return span; }
    virtual void init(V3 o, V3 dir)
    {
        const PtbPrim &p = d->prims[prim];
        if(p.chain >= 0)
        {
            const PtbChain &c = d->chains[p.chain];
            for(int k = 0; k < c.count; k++)
            {
                const float *m = d->xforms[d->chain_xf[c.first + k]].m;
                o = apply(m, o);
                dir = applyNoTranslate(m, dir);
            }
        }
        ended = false;
        span.startPrim = span.endPrim = prim;
        if(p.type == PTB_PRIM_SPHERE)
        { // src/sphere.cpp:31-49
            V3 center(p.p[0], p.p[1], p.p[2]);
            V3 oc = o - center;
            float a = dot(dir, dir);
            float b = dot(oc, dir);
            float c = dot(oc, oc) - p.p[3];
            float disc = b * b - a * c;
            if(disc <= EPS)
            {
                ended = true;
                return;
            }
            float s = sqrtf(disc);
            span.start = (-b - s) / a;
            span.end = (-b + s) / a;
            span.startNormal = normalize((o + span.start * dir) - center);
            span.endNormal = normalize((o + span.end * dir) - center);
        }
        else
        { // src/plane.cpp:35-63; normals pre-normalised at :29-30
            V3 n(p.p[0], p.p[1], p.p[2]);
            span.startNormal = span.endNormal = normalize(n);
            float divisor = dot(dir, n);
            float numerator = -p.p[3] - dot(o, n);
            float t = 0;
            if(fabsf(divisor) < EPS * EPS || fabsf(t = numerator / divisor) >= MAX_VALUE)
            {
                if(fabsf(numerator) < EPS * EPS)
                {
                    span.start = -MAX_VALUE;
                    span.end = MAX_VALUE;
                }
                else
                {
                    ended = true;
                    return;
                }
            }
            else if(divisor < 0)
            {
                span.start = t;
                span.end = MAX_VALUE;
            }
            else
            {
                span.start = -MAX_VALUE;
                span.end = t;
            }
        }
        if(p.chain >= 0)
        {
            const PtbChain &c = d->chains[p.chain];
            for(int k = c.count - 1; k >= 0; k--)
            {
                const float *inv = d->xforms[d->chain_xf[c.first + k]].inv;
                span.startNormal = normalize(applyNoTranslate(inv, span.startNormal));
                span.endNormal = normalize(applyNoTranslate(inv, span.endNormal));
            }
        }
    }
};

// Shared plumbing of the three combinators (src/union.cpp:21-67 and twins).
struct PairIter : Iter
{
    Iter *a, *b;
    Span spanA, spanB, result;
    bool ended, aEnded, bEnded;
    PairIter(Iter *a, Iter *b) : a(a), b(b), ended(true), aEnded(true), bEnded(true) {}
    virtual ~PairIter()
    {
        delete a;
        delete b;
    }
    void nextA()
    {
        if(a->atEnd())
            aEnded = true;
        else
        {
            spanA = a->cur();
            a->next();
        }
    }
    void nextB()
    {
        if(b->atEnd())
            bEnded = true;
        else
        {
            spanB = b->cur();
            b->next();
        }
    }
    virtual void init(V3 o, V3 dir)
    {
        a->init(o, dir);
        b->init(o, dir);
        ended = aEnded = bEnded = false;
        nextA();
        nextB();
        next();
    }
    virtual bool atEnd() const { return ended; }
    virtual const Span &cur() const { return result; }
};

struct UnionIter : PairIter // src/union.cpp:84-134
{
    UnionIter(Iter *a, Iter *b) : PairIter(a, b) {}
    virtual void next()
    {
        for(;;)
        {
            if(aEnded)
            {
                if(bEnded)
                {
                    ended = true;
                    return;
                }
                result = spanB;
                nextB();
                return;
            }
            if(bEnded || spanA.end < spanB.start)
            {
                result = spanA;
                nextA();
                return;
            }
            if(spanB.end < spanA.start)
            {
                result = spanB;
                nextB();
                return;
            }
            if(spanA.start < spanB.start)
            {
                if(spanA.end < spanB.end)
                    spanA.copyEndFromEnd(spanB);
                nextB();
            }
            else
            {
                if(spanA.end > spanB.end)
                    spanB.copyEndFromEnd(spanA);
                nextA();
            }
        }
    }
};

struct DifferenceIter : PairIter // src/difference.cpp:84-135 (quirk Q4 at :124-131 kept)
{
    DifferenceIter(Iter *a, Iter *b) : PairIter(a, b) {}
    virtual void next()
    {
        for(;;)
        {
            if(aEnded)
            {
                ended = true;
                return;
            }
            if(bEnded || spanA.end < spanB.start)
            {
                result = spanA;
                nextA();
                return;
            }
            if(spanB.end < spanA.start)
                nextB();
            else if(spanA.start < spanB.start)
            {
                if(spanA.end < spanB.end)
                {
                    spanA.copyEndFromStart(spanB);
                    result = spanA;
                    nextA();
                    return;
                }
                result = spanA;
                result.copyEndFromStart(spanB);
                spanA.copyStartFromEnd(spanB);
                nextB();
                return;
            }
            else
            {
                if(spanA.end > spanB.end)
                {
                    spanA.copyEndFromStart(spanB);
                    nextB();
                    continue;
                }
                nextA();
            }
        }
    }
};

struct IntersectionIter : PairIter // src/intersection.cpp:84-130
{
    IntersectionIter(Iter *a, Iter *b) : PairIter(a, b) {}
    virtual void next()
    {
        for(;;)
        {
            if(aEnded || bEnded)
            {
                ended = true;
                return;
            }
            if(spanA.end < spanB.start)
            {
                nextA();
                continue;
            }
            if(spanB.end < spanA.start)
            {
                nextB();
                continue;
            }
            if(spanA.start < spanB.start)
            {
                if(spanA.end < spanB.end)
                {
                    spanA.copyStartFromStart(spanB);
                    result = spanA;
                    nextA();
                    return;
                }
                result = spanB;
                nextB();
                return;
            }
            else
            {
                if(spanB.end < spanA.end)
                {
                    spanB.copyStartFromStart(spanA);
                    result = spanB;
                    nextB();
                    return;
                }
                result = spanA;
                nextA();
                return;
            }
        }
    }
};

// Postfix program -> iterator tree (what Object::makeSpanIterator builds, src/union.cpp:150-153).
Iter *buildIterator(const PtbSceneDesc *d)
{
    std::vector<Iter *> stack;
    for(int i = 0; i < d->n_ops; i++)
    {
        uint32_t op = d->ops[i];
        unsigned kind = PTB_OP_KIND(op);
        if(kind == PTB_OP_PUSH_PRIM)
        {
            stack.push_back(new LeafIter(d, (int)PTB_OP_ARG(op)));
            continue;
        }
        if(stack.size() < 2)
        {
            for(size_t k = 0; k < stack.size(); k++)
                delete stack[k];
            return NULL;
        }
        Iter *b = stack.back();
        stack.pop_back();
        Iter *a = stack.back();
        stack.pop_back();
        if(kind == PTB_OP_UNION)
            stack.push_back(new UnionIter(a, b));
        else if(kind == PTB_OP_DIFFERENCE)
            stack.push_back(new DifferenceIter(a, b));
        else
            stack.push_back(new IntersectionIter(a, b));
    }
    if(stack.size() != 1)
    {
        for(size_t k = 0; k < stack.size(); k++)
            delete stack[k];
        return NULL;
    }
    return stack[0];
}

// ---- textures --------------------------------------------------------------------------------
// Image::getPixel / getPixelAlpha, src/image.cpp:366-396: black / 0 outside.
inline void imagePixel(const PtbImageDesc &im, int x, int y, float out[4])
{
    if(!im.rgba || y < 0 || y >= im.height || x < 0 || x >= im.width)
    {
        out[0] = out[1] = out[2] = out[3] = 0;
        return;
    }
    const float *p = im.rgba + 4 * ((size_t)x + (size_t)y * (size_t)im.width);
    out[0] = p[0], out[1] = p[1], out[2] = p[2], out[3] = p[3];
}
// ImageTexture::getColor, include/image_texture.h:18-28 (nearest texel, frac x, 1 - frac y)
inline void imageLookup(const PtbImageDesc &im, V3 v, float out[4])
{
    float x = v.x, y = v.y;
    x -= floorf(x);
    y -= floorf(y);
    y = 1 - y;
    x *= (float)(unsigned)im.width;
    y *= (float)(unsigned)im.height;
    imagePixel(im, (int)floorf(x), (int)floorf(y), out);
}
// ImageSkyboxTexture::getColor(float,float,Image), include/image_texture.h:77-85 (double math)
inline void faceLookup(const PtbImageDesc &im, float x, float y, float out[4])
{
    x = x * 0.5 + 0.5;
    y = 0.5 - y * 0.5;
    x *= (float)(unsigned)im.width;
    y *= (float)(unsigned)im.height;
    imagePixel(im, (int)floorf(x), (int)floorf(y), out);
}
// face selection, include/image_texture.h:91-110; images: top,bottom,left,right,front,back
void skyboxLookup(const PtbImageDesc *im, V3 v, float out[4])
{
    if(isZero(v))
    {
        out[0] = out[1] = out[2] = out[3] = 0;
        return;
    }
    V3 a(fabsf(v.x), fabsf(v.y), fabsf(v.z));
    if(a.x > a.y && a.x > a.z)
    {
        if(v.x < 0)
            return faceLookup(im[2], -v.z / a.x, v.y / a.x, out);
        return faceLookup(im[3], v.z / a.x, v.y / a.x, out);
    }
    if(a.y > a.z)
    {
        if(v.y < 0)
            return faceLookup(im[1], -v.x / a.y, v.z / a.y, out);
        return faceLookup(im[0], v.x / a.y, v.z / a.y, out);
    }
    if(v.z < 0)
        return faceLookup(im[5], v.x / a.z, v.y / a.z, out);
    return faceLookup(im[4], -v.x / a.z, v.y / a.z, out);
}
inline float myLog(float v) // include/filter_texture.h:61-66
{
    if(v <= 1e-30)
        return 0;
    return 0.5f + logf(v) / (float)log(2.0) / 256;
}
inline float meanRgb(V3 c) { return (c.x + c.y + c.z) * (1.0f / 3.0f); } // include/texture.h:14-18

// Evaluates a texture chain.  wantFloat selects Texture::getFloat: position wrappers pass
// getFloat down (texture.h:78-81, transform_texture.h:30-33); filter wrappers do not override
// it, so below a filter the chain runs in getColor mode and the mean is taken above it.
void evalTexture(const PtbSceneDesc *d, int root, V3 pos, bool wantFloat, V3 &color, float &f)
{
    int i = root;
    int firstFilter = -1;
    for(;; i++)
    {
        const PtbTexOp &op = d->texops[i];
        if(PTB_TEX_IS_LEAF(op.kind))
            break;
        switch(op.kind)
        {
        case PTB_TEX_AFFINE: // texture.h:74-81
            pos = apply(d->texdata + op.arg, pos);
            break;
        case PTB_TEX_MIRRORBALL: // transform_texture.h:46-59
            if(isZero(pos))
                pos = V3();
            else
            {
                V3 v = normalize(pos);
                if(v.z <= -1)
                    pos = V3(0, 0.5, 0);
                else
                {
                    float dd = sqrtf(2 + 2 * v.z);
                    if(dd == 0)
                        pos = V3(0, 0.5, 0);
                    else
                    {
                        float xt = v.x / dd, yt = v.y / dd;
                        pos = V3(xt * 0.5 + 0.5, yt * 0.5 + 0.5, 0);
                    }
                }
            }
            break;
        case PTB_TEX_SPHERICAL: // transform_texture.h:73-85
            if(isZero(pos))
                pos = V3();
            else
            {
                V3 v = normalize(pos);
                float theta = atan2f(v.y, v.x);
                if(theta < -M_PI)
                    theta += 2 * M_PI;
                if(theta > M_PI)
                    theta -= 2 * M_PI;
                float phi = asinf(v.z);
                pos = V3(theta * 0.5 / M_PI + 0.5, phi / (M_PI / 2) * 0.5 + 0.5, 0);
            }
            break;
        default: // MULTIPLY / LOG: applied on the way back up
            if(firstFilter < 0)
                firstFilter = i;
            break;
        }
    }
    const PtbTexOp &leaf = d->texops[i];
    bool floatAtLeaf = wantFloat && firstFilter < 0;
    float px[4];
    switch(leaf.kind)
    {
    case PTB_TEX_CONST:
        color = V3(d->texdata[leaf.arg], d->texdata[leaf.arg + 1], d->texdata[leaf.arg + 2]);
        f = meanRgb(color);
        break;
    case PTB_TEX_IMAGE:
        imageLookup(d->images[leaf.arg], pos, px);
        color = V3(px[0], px[1], px[2]);
        f = meanRgb(color);
        break;
    case PTB_TEX_IMAGE_ALPHA: // image_texture.h:45-66: getColor = Color(alpha), getFloat = alpha
        imageLookup(d->images[leaf.arg], pos, px);
        color = V3(px[3]);
        f = px[3];
        break;
    case PTB_TEX_SKYBOX:
        skyboxLookup(d->images + leaf.arg, pos, px);
        color = V3(px[0], px[1], px[2]);
        f = meanRgb(color);
        break;
    default: // PTB_TEX_SKYBOX_ALPHA, image_texture.h:135-176
        skyboxLookup(d->images + leaf.arg, pos, px);
        color = V3(px[3]);
        f = px[3];
        break;
    }
    if(floatAtLeaf)
        return;
    for(int k = i - 1; k >= root; k--)
    {
        const PtbTexOp &op = d->texops[k];
        if(op.kind == PTB_TEX_MULTIPLY) // filter_texture.h:43-46
            color = color * V3(d->texdata[op.arg], d->texdata[op.arg + 1], d->texdata[op.arg + 2]);
        else if(op.kind == PTB_TEX_LOG) // filter_texture.h:68-71
            color = V3(myLog(color.x), myLog(color.y), myLog(color.z));
    }
    f = meanRgb(color);
}
inline V3 texColor(const PtbSceneDesc *d, int root, V3 pos)
{
    V3 c;
    float f;
    evalTexture(d, root, pos, false, c, f);
    return c;
}
inline float texFloat(const PtbSceneDesc *d, int root, V3 pos)
{
    V3 c;
    float f;
    evalTexture(d, root, pos, true, c, f);
    return f;
}

// ---- first-hit scan, include/path-trace.h:66-101 --------------------------------------------
struct Hit
{
    float t;
    V3 normal;
    int kind, prim;
    float ior; // relative ior: 1/mat.ior entering, mat.ior leaving
};
Hit firstHit(const PtbSceneDesc *d, Iter &it, V3 o, V3 dir, unsigned long long *rays)
{
    Hit h;
    h.t = -1, h.kind = PTB_HIT_MISS, h.prim = -1, h.ior = 1;
    if(rays)
        ++*rays;
    it.init(o, dir);
    for(; !it.atEnd(); it.next())
    {
        const Span &s = it.cur();
        if(s.start >= MAX_VALUE)
            return h;
        if(s.start >= EPS)
        {
            h.t = s.start, h.normal = s.startNormal, h.prim = s.startPrim, h.kind = PTB_HIT_ENTER;
            h.ior = 1.0 / d->materials[d->prims[h.prim].material].ior;
            return h;
        }
        if(s.end >= MAX_VALUE)
            return h;
        if(s.end >= EPS)
        {
            h.t = s.end, h.normal = -s.endNormal, h.prim = s.endPrim, h.kind = PTB_HIT_EXIT;
            h.ior = d->materials[d->prims[h.prim].material].ior;
            return h;
        }
    }
    return h;
}

// ---- traceRay, include/path-trace.h:58-165 — the branching estimator, literally -------------
V3 traceRay(const PtbSceneDesc *d, Iter &it, V3 o, V3 dir, int depth, Lcg &rng, float strength, unsigned long long *rays)
{
    Hit h = firstHit(d, it, o, dir, rays);
    if(h.kind == PTB_HIT_MISS)
        return V3();
    const PtbMaterial &m = d->materials[d->prims[h.prim].material];
    float t = h.t;
    V3 normal = h.normal;
    float ior = h.ior;
    V3 hitPos = o + t * dir;
    V3 retval = texColor(d, m.tex[PTB_MAT_EMISSIVE], hitPos);
    float addFactor = 1;
    if(depth <= 0 || strength < EPS)
        return retval;
    float refractFactor = std::max(0.0f, std::min(1.0f, texFloat(d, m.tex[PTB_MAT_TRC], hitPos))) * refractStrength(dir, ior, normal);
    if(refractFactor > EPS)
    {
        V3 rd = refract(dir, ior, normal);
        if(!isZero(rd))
        {
            V3 transmit = texColor(d, m.tex[PTB_MAT_TRANSMIT], hitPos);
            V3 sub = traceRay(d, it, hitPos, rd, depth - 1, rng, strength * refractFactor * addFactor * vabs(transmit), rays);
            retval = retval + addFactor * refractFactor * transmit * sub;
            addFactor *= 1 - refractFactor;
        }
    }
    if(addFactor < EPS)
        return retval;
    float sc = texFloat(d, m.tex[PTB_MAT_SCATTER], hitPos);
    sc = std::max(0.0f, std::min(1.0f, sc));
    int count = (int)(10000 * strength * addFactor * sc);
    if(sc <= EPS)
        count = 1;
    if(count == 0)
        count = 1;
    V3 refl = texColor(d, m.tex[PTB_MAT_REFLECT], hitPos);
    for(int i = 0; i < count; i++)
    {
        V3 r = reflect(dir, normal);
        V3 nd = r;
        if(sc > EPS)
        {
            int tries = 0;
            do
            {
                tries++;
                if(tries > 1000)
                    return retval;
                nd = randBall(rng, 1, 0);
                nd = nd + (1 / sc - 1) * r;
            } while(dot(normal, nd) <= EPS);
            nd = normalize(nd);
        }
        float factor = 1 - (1 - dot(nd, normal)) * sc;
        V3 sub = traceRay(d, it, o + t * dir, nd, depth - 1, rng, strength / count * addFactor * factor * vabs(refl), rays);
        retval = retval + addFactor / count * factor * refl * sub;
    }
    return retval;
}

// tracePixel with per-sample jitter, include/path-trace.h:187-201
V3 tracePixel(const PtbSceneDesc *d, Iter &it, int px, int py, int W, int H, int sampleCount, int depth, float sw, float sh, float sd, Lcg &rng,
              unsigned long long *rays)
{
    V3 sum;
    for(int i = 0; i < sampleCount; i++)
    {
        float x = 2 * (px + uniform(rng, 0, 1)) / W - 1;
        float y = 1 - 2 * (py + uniform(rng, 0, 1)) / H;
        sum = sum + traceRay(d, it, V3(), V3(x * sw, y * sh, -sd), depth, rng, 1.0, rays);
    }
    return sum / (float)sampleCount;
}

double nowSeconds()
{
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}
} // namespace

// ---- C API (ctypes) --------------------------------------------------------------------------
extern "C" {

// Root span list + first hit for each ray; same record layout as ptb200_intersect.
int orc_intersect(const PtbSceneDesc *d, const float *origins, const float *dirs, size_t n, PtbHit *hits, PtbSpan *spans, int max_spans)
{
    Iter *it = buildIterator(d);
    if(!it)
        return -1;
    for(size_t i = 0; i < n; i++)
    {
        V3 o(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]), dir(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]);
        int count = 0;
        it->init(o, dir);
        for(; !it->atEnd(); it->next())
        {
            if(spans && count < max_spans)
            {
                const Span &s = it->cur();
                PtbSpan &out = spans[i * (size_t)max_spans + count];
                out.start = s.start, out.end = s.end;
                out.start_normal[0] = s.startNormal.x, out.start_normal[1] = s.startNormal.y, out.start_normal[2] = s.startNormal.z;
                out.end_normal[0] = s.endNormal.x, out.end_normal[1] = s.endNormal.y, out.end_normal[2] = s.endNormal.z;
                out.start_prim = s.startPrim, out.end_prim = s.endPrim;
                out.start_material = d->prims[s.startPrim].material, out.end_material = d->prims[s.endPrim].material;
            }
            count++;
        }
        if(hits)
        {
            Hit h = firstHit(d, *it, o, dir, NULL);
            PtbHit &out = hits[i];
            out.t = h.t;
            out.normal[0] = h.normal.x, out.normal[1] = h.normal.y, out.normal[2] = h.normal.z;
            out.kind = h.kind, out.prim = h.prim;
            out.material = h.prim >= 0 ? d->prims[h.prim].material : -1;
            out.n_spans = count;
            if(h.kind == PTB_HIT_MISS)
                out.normal[0] = out.normal[1] = out.normal[2] = 0;
        }
    }
    delete it;
    return 0;
}

int orc_trace_rays(const PtbSceneDesc *d, const float *origins, const float *dirs, size_t n, int depth, unsigned seed, float strength, float *rgb)
{
    Iter *it = buildIterator(d);
    if(!it)
        return -1;
    Lcg rng(seed);
    for(size_t i = 0; i < n; i++)
    {
        V3 c = traceRay(d, *it, V3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]), V3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]), depth, rng, strength, NULL);
        rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
    }
    delete it;
    return 0;
}

int orc_render(const PtbSceneDesc *d, int W, int H, int x0, int y0, int step, int nx, int ny, int sampleCount, int depth, float sw, float sh, float sd,
               unsigned seed, float *rgb, unsigned long long *rays)
{
    Iter *it = buildIterator(d);
    if(!it)
        return -1;
    Lcg rng(seed);
    unsigned long long count = 0;
    for(int j = 0; j < ny; j++)
        for(int i = 0; i < nx; i++)
        {
            V3 c = tracePixel(d, *it, x0 + i * step, y0 + j * step, W, H, sampleCount, depth, sw, sh, sd, rng, &count);
            float *o = rgb + 3 * ((size_t)j * nx + i);
            o[0] = c.x, o[1] = c.y, o[2] = c.z;
        }
    delete it;
    if(rays)
        *rays = count;
    return 0;
}

int orc_tex_eval(const PtbSceneDesc *d, int tex_root, const float *pos, size_t n, float *rgb, float *f)
{
    if(tex_root < 0 || tex_root >= d->n_texops)
        return -1;
    for(size_t i = 0; i < n; i++)
    {
        V3 p(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
        if(rgb)
        {
            V3 c = texColor(d, tex_root, p);
            rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
        }
        if(f)
            f[i] = texFloat(d, tex_root, p);
    }
    return 0;
}

void orc_rng_u32(unsigned seed, size_t n, unsigned *out)
{
    Lcg r(seed);
    for(size_t i = 0; i < n; i++)
        out[i] = r.next();
}
void orc_rng_uniform(unsigned seed, size_t n, float lo, float hi, float *out)
{
    Lcg r(seed);
    for(size_t i = 0; i < n; i++)
        out[i] = uniform(r, lo, hi);
}
void orc_rng_ball(unsigned seed, size_t n, float max, float min, float *out)
{
    Lcg r(seed);
    for(size_t i = 0; i < n; i++)
    {
        V3 v = randBall(r, max, min);
        out[3 * i] = v.x, out[3 * i + 1] = v.y, out[3 * i + 2] = v.z;
    }
}
void orc_vec_reflect(const float *v, const float *n, float *out)
{
    V3 r = reflect(V3(v[0], v[1], v[2]), V3(n[0], n[1], n[2]));
    out[0] = r.x, out[1] = r.y, out[2] = r.z;
}
void orc_vec_refract(const float *v, float ior, const float *n, float *out)
{
    V3 r = refract(V3(v[0], v[1], v[2]), ior, V3(n[0], n[1], n[2]));
    out[0] = r.x, out[1] = r.y, out[2] = r.z;
}
float orc_vec_refract_strength(const float *v, float ior, const float *n) { return refractStrength(V3(v[0], v[1], v[2]), ior, V3(n[0], n[1], n[2])); }

} // extern "C"

// ---- multithreaded driver ("port" CPU baseline when oracle/_ref is unavailable) -------------
namespace
{
struct MtJob
{
    const PtbSceneDesc *d;
    int W, H, x0, y0, step, nx, ny, sampleCount, depth;
    float sw, sh, sd;
    unsigned seed;
    float *rgb;
    double deadline;
    pthread_mutex_t lock;
    int nextRow;
    unsigned long long rays, pixels;
};
void *mtWorker(void *arg)
{
    MtJob &job = *(MtJob *)arg;
    Iter *it = buildIterator(job.d);
    unsigned long long rays = 0, pixels = 0;
    for(;;)
    {
        pthread_mutex_lock(&job.lock);
        int row = job.nextRow < job.ny && nowSeconds() < job.deadline ? job.nextRow++ : -1;
        pthread_mutex_unlock(&job.lock);
        if(row < 0)
            break;
        Lcg rng(job.seed + 7919u * (unsigned)row);
        for(int i = 0; i < job.nx; i++)
        {
            V3 c = tracePixel(job.d, *it, job.x0 + i * job.step, job.y0 + row * job.step, job.W, job.H, job.sampleCount, job.depth, job.sw, job.sh, job.sd, rng, &rays);
            float *o = job.rgb + 3 * ((size_t)row * job.nx + i);
            o[0] = c.x, o[1] = c.y, o[2] = c.z;
            pixels++;
        }
    }
    delete it;
    pthread_mutex_lock(&job.lock);
    job.rays += rays;
    job.pixels += pixels;
    pthread_mutex_unlock(&job.lock);
    return NULL;
}
} // namespace

extern "C" {
int orc_hardware_concurrency(void) { return (int)sysconf(_SC_NPROCESSORS_ONLN); }

int orc_render_mt(const PtbSceneDesc *d, int W, int H, int x0, int y0, int step, int nx, int ny, int sampleCount, int depth, float sw, float sh, float sd,
                  unsigned seed, int threads, double max_seconds, float *rgb, unsigned long long *rays, unsigned long long *pixels, double *seconds)
{
    Iter *probe = buildIterator(d);
    if(!probe)
        return -1;
    delete probe;
    MtJob job;
    job.d = d;
    job.W = W, job.H = H, job.x0 = x0, job.y0 = y0, job.step = step, job.nx = nx, job.ny = ny;
    job.sampleCount = sampleCount, job.depth = depth, job.sw = sw, job.sh = sh, job.sd = sd, job.seed = seed, job.rgb = rgb;
    job.nextRow = 0, job.rays = 0, job.pixels = 0;
    pthread_mutex_init(&job.lock, NULL);
    double t0 = nowSeconds();
    job.deadline = max_seconds > 0 ? t0 + max_seconds : 1e300;
    std::vector<pthread_t> th(threads);
    for(int i = 0; i < threads; i++)
        pthread_create(&th[i], NULL, mtWorker, &job);
    for(int i = 0; i < threads; i++)
        pthread_join(th[i], NULL);
    double t1 = nowSeconds();
    pthread_mutex_destroy(&job.lock);
    if(rays)
        *rays = job.rays;
    if(pixels)
        *pixels = job.pixels;
    if(seconds)
        *seconds = t1 - t0;
    return job.nextRow;
}
} // extern "C"
