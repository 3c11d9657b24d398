// ptb_trace.cuh — the TRACE side of the wavefront as device code in a header: camera samples, the
// staging of the fast evaluator, first-hit resolution and the bodies of the trace / intersect kernels.
// It is compiled twice: by nvcc into libptb200.so with the interpreting evaluator (evalFast,
// ptb_csg_fast.cuh), and at run time by NVRTC with PTB_EVAL bound to the evaluator GENERATED for one
// scene (ptb_specialize.cpp), whose program words are compile-time constants.
#pragma once
#include "ptb_launch.h"
#include "ptb_csg_fast.cuh"

#ifndef PTB_TRACE_RAYS
#define PTB_TRACE_RAYS 1
#endif
#ifndef PTB_TRACE_MIN_CTAS
#define PTB_TRACE_MIN_CTAS 8
#endif
// The evaluator the bodies below run: R rays per lane, root span lists left in the pools, tops[] = their lengths.
// Straight-line (specialised) evaluators keep the warps of a CTA in step, so that they fetch the same
// instruction-cache lines; the interpreter's loop needs no such help.
#ifndef PTB_TRACE_TRIP_SYNC
#define PTB_TRACE_TRIP_SYNC()
#endif
#ifndef PTB_EVAL
#if defined(PTB_SPEC_GEN) && PTB_SPEC_GEN == 2
#define PTB_EVAL(GP, O0, R, cx, gprog, o, d, tops) tops[0] = specEvalList<O0>(cx, gprog, o[0], d[0])
#else
#define PTB_EVAL(GP, O0, R, cx, gprog, o, d, tops) evalFast<GP, O0, R>(cx, gprog, o, d, tops)
#endif
#endif

namespace ptb
{

#if defined(PTB_SPEC_GEN) && PTB_SPEC_GEN == 2
// Register-resident evaluator of ONE scene (generated: specTrace / specList, ptb_csg_reg.cuh).  A ray whose lists
// outgrow the shared pool flags it; the warp (all 32 lanes are here: idle ones carry null rays) is then rerun
// through the interpreter, which has the global overflow area — same spans, more instructions.
template <bool O0>
__device__ __forceinline__ REval specBegin(const FastCtx &cx)
{
    REval ev;
    ev.poolBase = cx.pool.base, ev.cap = cx.pool.cap, ev.top = 0, ev.ovf = false, ev.tab = cx.tab;
    return ev;
}
// The interpreter as the rare way out, out of line and with plain arguments so that it does not weigh on the
// register allocation of the generated code.  Returns the length of the root list (pool slots [0, n)).
template <bool O0>
static __device__ __noinline__ int specFallback(const uint4 *__restrict__ gprog, uint32_t poolBase, int cap, float *ovf, size_t ovfStride, uint32_t stk, float ox, float oy, float oz,
                                                float dx, float dy, float dz)
{
    FastCtx cx;
    cx.prog = 0u, cx.stk = stk, cx.stkStride = 0u, cx.poolStride = 0u, cx.ovfRay = 0, cx.tab = 0u;
    cx.pool.base = poolBase, cx.pool.cap = cap, cx.pool.ovf = ovf, cx.pool.ovfStride = ovfStride;
    V3 oo[1] = {mk(ox, oy, oz)}, dd[1] = {mk(dx, dy, dz)};
    int tops[1];
    evalFast<true, O0, 1>(cx, gprog, oo, dd, tops);
    return tops[0];
}
template <bool O0>
__device__ __forceinline__ int specEvalList(const FastCtx &cx, const uint4 *__restrict__ gprog, V3 o, V3 d)
{
    REval ev = specBegin<O0>(cx);
    int n = specList<O0>(ev, o, d);
    if(__any_sync(0xFFFFFFFFu, ev.ovf))
        n = specFallback<O0>(gprog, cx.pool.base, cx.pool.cap, cx.pool.ovf, cx.pool.ovfStride, cx.stk, o.x, o.y, o.z, d.x, d.y, d.z);
    return n;
}
template <bool O0>
__device__ __forceinline__ uint2 specTraceHit(const DevScene &sc, const FastCtx &cx, V3 o, V3 d)
{
    REval ev = specBegin<O0>(cx);
    float t = -1.f;
    uint32_t tag = 0u;
    int kind = specTrace<O0>(ev, o, d, t, tag);
    if(__any_sync(0xFFFFFFFFu, ev.ovf))
    {
        const int n = specFallback<O0>(sc.fastProgram, cx.pool.base, cx.pool.cap, cx.pool.ovf, cx.pool.ovfStride, cx.stk, o.x, o.y, o.z, d.x, d.y, d.z);
        t = -1.f, tag = 0u;
        kind = firstHit(cx.pool, n, t, tag);
    }
    return make_uint2(__float_as_uint(t), ((uint32_t)kind << 16) | tag);
}
#endif

// Fast evaluator: stage the program (unless GP: it stays in global memory) and lay out the pools
// and start stacks of the R rays each thread carries.
template <bool GP, int R>
__device__ __forceinline__ FastCtx setupFast(const DevScene &sc, int cap, float *overflow, size_t ovfStride)
{
    FastCtx c;
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(ptbSmem);
    uint32_t words = 0;
    if(!GP)
    {
        words = (uint32_t)sc.fastWords;
        for(uint32_t i = threadIdx.x; i < words; i += PTB_FAST_THREADS)
            sts128(base + 16u * i, __ldg(sc.fastProgram + i));
    }
    c.prog = base;
    c.poolStride = (uint32_t)cap * (12u * PTB_FAST_THREADS);
    c.stkStride = (uint32_t)(sc.maxStack + 1) * (4u * PTB_FAST_THREADS);
    c.pool.base = base + 16u * words + 4u * threadIdx.x;
    c.pool.cap = cap;
    c.pool.ovf = overflow ? overflow + (size_t)blockIdx.x * PTB_FAST_THREADS + threadIdx.x : nullptr;
    c.pool.ovfStride = ovfStride;
    c.ovfRay = (size_t)(sc.poolBound > cap ? sc.poolBound - cap : 0) * 3 * ovfStride;
    c.stk = c.pool.base + (uint32_t)R * c.poolStride;
    c.tab = base + 16u * words + (uint32_t)R * (c.poolStride + c.stkStride);
#if defined(PTB_SPEC_GEN) && PTB_SPEC_GEN == 2
    specStageTable(c.tab); // the generated module's leaf constants: constant memory -> shared memory, once per CTA
#endif
    __syncthreads();
    return c;
}

// Resolves the first boundary with t >= eps into normal, material and relative ior.
template <bool EX, class P>
__device__ __forceinline__ int resolveFirstHit(const DevScene &sc, const P &pool, int nSpans, V3 o, V3 d, Hit &hit, uint32_t &tag)
{
    float t = -1.f;
    tag = 0u;
    int kind = firstHit(pool, nSpans, t, tag);
    hit.t = t;
    if(kind == PTB_HIT_MISS)
        return kind;
    V3 n = boundaryNormal<EX>(sc, tag, t, o, d);
    int mat = __ldg(&sc.primMeta[tag >> 1]).y;
    hit.mat = mat;
    if(kind == PTB_HIT_ENTER)
    { // path-trace.h:72-81
        hit.n = n;
        hit.ior = sc.materials[mat].iorIn;
    }
    else
    { // path-trace.h:86-95
        hit.n = neg(n);
        hit.ior = sc.materials[mat].ior;
    }
    return kind;
}

template <bool EX, class P>
__device__ __forceinline__ void writeHitAndSpans(const DevScene &sc, const P &pool, size_t i, int kind, int nSpans, const Hit &hit, uint32_t tag, V3 o, V3 d,
                                                 PtbHit *__restrict__ hits, PtbSpan *__restrict__ spans, int maxSpans)
{
    float4 a, b;
    if(kind == PTB_HIT_MISS)
    {
        a = make_float4(hit.t, 0.f, 0.f, 0.f);
        b = make_float4(__int_as_float(kind), __int_as_float(-1), __int_as_float(-1), __int_as_float(nSpans));
    }
    else
    {
        a = make_float4(hit.t, hit.n.x, hit.n.y, hit.n.z);
        b = make_float4(__int_as_float(kind), __int_as_float((int)(tag >> 1)), __int_as_float(hit.mat), __int_as_float(nSpans));
    }
    // 32-byte record as two float4 stores
    float4 *dst = reinterpret_cast<float4 *>(hits + i);
    dst[0] = a;
    dst[1] = b;
    if(spans)
    {
        int m = nSpans < maxSpans ? nSpans : maxSpans;
        for(int s = 0; s < m; s++)
        {
            SpanRec r = pool.load(s);
            PtbSpan &ps = spans[i * (size_t)maxSpans + s];
            V3 sn = boundaryNormal<EX>(sc, startTag(r), r.t0, o, d);
            V3 en = boundaryNormal<EX>(sc, endTag(r), r.t1, o, d);
            ps.start = r.t0, ps.end = r.t1;
            ps.start_normal[0] = sn.x, ps.start_normal[1] = sn.y, ps.start_normal[2] = sn.z;
            ps.end_normal[0] = en.x, ps.end_normal[1] = en.y, ps.end_normal[2] = en.z;
            ps.start_prim = (int)(startTag(r) >> 1), ps.end_prim = (int)(endTag(r) >> 1);
            ps.start_material = __ldg(&sc.primMeta[ps.start_prim]).y;
            ps.end_material = __ldg(&sc.primMeta[ps.end_prim]).y;
        }
    }
}

// Fast policy: the evaluator the renderer runs.
template <bool GP, bool O0>
__device__ __forceinline__ void intersectFastBody(const DevScene &sc, int cap, float *overflow, size_t ovfStride, const float4 *__restrict__ origins, const float4 *__restrict__ dirs,
                                                  size_t n, PtbHit *__restrict__ hits, PtbSpan *__restrict__ spans, int maxSpans)
{
    FastCtx cx = setupFast<GP, 1>(sc, cap, overflow, ovfStride);
    size_t stride = (size_t)gridDim.x * PTB_FAST_THREADS;
    // warp-uniform trip count: the evaluator needs all 32 lanes (idle lanes carry a null ray)
    for(size_t base = (size_t)blockIdx.x * PTB_FAST_THREADS + (threadIdx.x & ~31u); base < n; base += stride)
    {
        size_t i = base + (threadIdx.x & 31u);
        bool valid = i < n;
        V3 o = mk(0.f, 0.f, 0.f), d = o;
        if(valid)
        {
            float4 o4 = __ldg(&origins[i]), d4 = __ldg(&dirs[i]);
            o = mk(o4.x, o4.y, o4.z), d = mk(d4.x, d4.y, d4.z);
        }
        V3 oo[1] = {o}, dd[1] = {d};
        int tops[1];
        PTB_EVAL(GP, O0, 1, cx, sc.fastProgram, oo, dd, tops);
        int nSpans = tops[0];
        if(valid)
        {
            Hit hit;
            uint32_t tag;
            int kind = resolveFirstHit<false>(sc, cx.pool, nSpans, o, d, hit, tag);
            writeHitAndSpans<false>(sc, cx.pool, i, kind, nSpans, hit, tag, o, d, hits, spans, maxSpans);
        }
    }
}

__device__ __forceinline__ void queueStore(const RayQueue &q, unsigned idx, const PathState &ps)
{
    q.q0[idx] = make_float4(ps.o.x, ps.o.y, ps.o.z, __uint_as_float(ps.pixel));
    q.q1[idx] = make_float4(ps.d.x, ps.d.y, ps.d.z, ps.strength);
    q.q2[idx] = make_float4(ps.thr.x, ps.thr.y, ps.thr.z, __uint_as_float(ps.sample));
    q.q3[idx] = make_float4(__uint_as_float(ps.depth), __uint_as_float(ps.rngBlock), 0.f, 0.f);
}
__device__ __forceinline__ void queueLoad(const RayQueue &q, unsigned idx, PathState &ps)
{
    float4 a = __ldcs(&q.q0[idx]), b = __ldcs(&q.q1[idx]), c = __ldcs(&q.q2[idx]), e = __ldcs(&q.q3[idx]);
    ps.o = mk(a.x, a.y, a.z), ps.pixel = __float_as_uint(a.w);
    ps.d = mk(b.x, b.y, b.z), ps.strength = b.w;
    ps.thr = mk(c.x, c.y, c.z), ps.sample = __float_as_uint(c.w);
    ps.depth = __float_as_uint(e.x), ps.rngBlock = __float_as_uint(e.y);
}

// Camera sample i of this pass: pixel, global sample index, jittered pinhole ray
// (tracePixel, reference include/path-trace.h:191-196).  Called by the primary trace kernel and
// again by the primary shade kernel (regenerating is cheaper than a 64-byte round trip to HBM).
__device__ __forceinline__ void cameraSample(const RenderPass &pass, unsigned nPix, unsigned long long i, PathState &ps)
{
    const unsigned i32 = (unsigned)i; // the queue capacity check keeps pixels x samples per pass below 2^32
    // PIXEL-major: work item i is sample i % S of pixel i / S (S = samples per pixel in this pass), so the 32 lanes of
    // a warp hold consecutive samples of one or two pixels.  The shade kernels sum a pixel's samples across the warp
    // and issue ONE float4 atomic per pixel and warp instead of one per path; sample-major order swept the 133 MB
    // accumulation buffer (> L2) once per sample: 32 B of DRAM traffic per path, the bulk of the primary shade
    // kernel's 49 B (profiles/r01n_spec_c5.txt).  Camera rays of a warp are also as coherent as they can be.
    (void)nPix;
    unsigned p = pass.bySamples.div(i32), s = i32 - p * (unsigned)pass.samplesThisPass;
    int px, py;
    if(pass.pixelList)
    { // adaptive rendering: an explicit list of lattice points
        const uint32_t packed = __ldg(pass.pixelList + p);
        px = (int)(packed & 0xFFFFu), py = (int)(packed >> 16);
    }
    else
    {
        unsigned row = pass.byRowWidth.div(p);
        px = pass.x0 + (int)(p - row * (unsigned)pass.rw), py = pass.y0 + (int)row;
    }
    ps.pixel = (unsigned)py * (unsigned)pass.width + (unsigned)px;
    ps.sample = (unsigned)(pass.sampleBase + (int)s * pass.sampleStride);
    const uint2 u = cameraJitter(pass.seed, ps.pixel, ps.sample);
    ps.rngBlock = 0u;
    float x = fmaf((float)px + u01(u.x), pass.twoOverWidth, -1.f);
    float y = fmaf((float)py + u01(u.y), -pass.twoOverHeight, 1.f);
    ps.o = mk(0.f, 0.f, 0.f);
    ps.d = mk(x * pass.sw, y * pass.sh, -pass.sd);
    ps.thr = mk(1.f, 1.f, 1.f);
    ps.strength = 1.f;
    ps.depth = (unsigned)pass.maxDepth;
}

// Evaluates the CSG program for PTB_TRACE_RAYS rays per lane (all 32 lanes take part: SKIP words
// vote) and packs each ray's first boundary with t >= eps.
template <bool GP, bool O0>
__device__ __forceinline__ void traceRays(const DevScene &sc, const FastCtx &cx, const V3 (&o)[PTB_TRACE_RAYS], const V3 (&d)[PTB_TRACE_RAYS], uint2 (&h)[PTB_TRACE_RAYS])
{
#if defined(PTB_SPEC_GEN) && PTB_SPEC_GEN == 2
    h[0] = specTraceHit<O0>(sc, cx, o[0], d[0]); // (the generated modules carry one ray per lane)
#else
    int tops[PTB_TRACE_RAYS];
    PTB_EVAL(GP, O0, PTB_TRACE_RAYS, cx, sc.fastProgram, o, d, tops);
#pragma unroll
    for(int r = 0; r < PTB_TRACE_RAYS; r++)
    {
        float t = -1.f;
        uint32_t tag = 0u;
        int kind = firstHit(cx.poolOf(r), tops[r], t, tag);
        h[r] = make_uint2(__float_as_uint(t), ((uint32_t)kind << 16) | tag);
    }
#endif
}

// A warp takes 32 * PTB_TRACE_RAYS consecutive rays per trip: lane l carries rays l, l + 32, ...
template <bool GP>
__device__ __forceinline__ void tracePrimaryBody(const DevScene &sc, int cap, float *overflow, size_t ovfStride, const RenderPass &pass, uint2 *__restrict__ hits)
{
    FastCtx cx = setupFast<GP, PTB_TRACE_RAYS>(sc, cap, overflow, ovfStride);
    const unsigned nPix = (unsigned)pass.rw * (unsigned)pass.rh;
    // (32-bit work indices: the host keeps paths per pass below 2^32 - 2^24, a grid has < 2^24 threads)
    const unsigned total = nPix * (unsigned)pass.samplesThisPass;
    const unsigned stride = gridDim.x * PTB_FAST_THREADS * PTB_TRACE_RAYS;
    // warp-uniform trip count; idle lanes of the last warp carry a null ray
    for(unsigned ctaBase = blockIdx.x * PTB_FAST_THREADS * PTB_TRACE_RAYS; ctaBase < total; ctaBase += stride)
    {
        const unsigned base = ctaBase + (threadIdx.x & ~31u) * PTB_TRACE_RAYS; // (a warp past the end carries null rays)
        PTB_TRACE_TRIP_SYNC();
        V3 o[PTB_TRACE_RAYS], d[PTB_TRACE_RAYS];
#pragma unroll
        for(int r = 0; r < PTB_TRACE_RAYS; r++)
        {
            const unsigned i = base + 32u * r + (threadIdx.x & 31u);
            o[r] = d[r] = mk(0.f, 0.f, 0.f);
            if(i < total)
            {
                PathState ps;
                cameraSample(pass, nPix, i, ps);
                d[r] = ps.d;
            }
        }
        uint2 h[PTB_TRACE_RAYS];
        traceRays<GP, true>(sc, cx, o, d, h);
#pragma unroll
        for(int r = 0; r < PTB_TRACE_RAYS; r++)
        {
            const unsigned i = base + 32u * r + (threadIdx.x & 31u);
            if(i < total)
                __stcs(&hits[i], h[r]);
        }
    }
}

template <bool GP>
__device__ __forceinline__ void traceQueueBody(const DevScene &sc, int cap, float *overflow, size_t ovfStride, const RayQueue &in, const unsigned *__restrict__ inCount,
                                               uint2 *__restrict__ hits)
{
    const unsigned total = *inCount;
    if(total == 0u)
        return;
    FastCtx cx = setupFast<GP, PTB_TRACE_RAYS>(sc, cap, overflow, ovfStride);
    const unsigned stride = gridDim.x * PTB_FAST_THREADS * PTB_TRACE_RAYS;
    for(unsigned ctaBase = blockIdx.x * PTB_FAST_THREADS * PTB_TRACE_RAYS; ctaBase < total; ctaBase += stride)
    {
        const unsigned base = ctaBase + (threadIdx.x & ~31u) * PTB_TRACE_RAYS; // (a warp past the end carries null rays)
        PTB_TRACE_TRIP_SYNC();
        V3 o[PTB_TRACE_RAYS], d[PTB_TRACE_RAYS];
#pragma unroll
        for(int r = 0; r < PTB_TRACE_RAYS; r++)
        {
            const unsigned i = base + 32u * r + (threadIdx.x & 31u);
            o[r] = d[r] = mk(0.f, 0.f, 0.f);
            if(i < total)
            {
                float4 a = __ldg(&in.q0[i]), b = __ldg(&in.q1[i]);
                o[r] = mk(a.x, a.y, a.z), d[r] = mk(b.x, b.y, b.z);
            }
        }
        uint2 h[PTB_TRACE_RAYS];
        traceRays<GP, false>(sc, cx, o, d, h);
#pragma unroll
        for(int r = 0; r < PTB_TRACE_RAYS; r++)
        {
            const unsigned i = base + 32u * r + (threadIdx.x & 31u);
            if(i < total)
                __stcs(&hits[i], h[r]);
        }
    }
}

} // namespace ptb
