// ptb_csg.cuh — span-list CSG on the device.
//
// The host compiles the postfix program of include/ptb200.h into a linear stream of 16-byte
// words (built in ptb_capi.cu, staged into shared memory by the kernels):
//     leaf:  header {kind, prim, nchain, flags}, geometry float4, then 3 float4 rows of `m`
//            for each TransformedObject wrapper above the leaf (outermost first) unless the
//            header says the previous leaf had the same chain (transformed ray is reused)
//     op:    header {kind}
// Every lane of a warp runs the same stream, so all program reads are warp-uniform
// (shared-memory broadcast) and only the span lists differ per lane.
//
// Span lists live in a per-thread stack ("pool") in shared memory, laid out
// [slot][thread] so that any mix of slot indices across a warp is bank-conflict free; slots
// beyond the shared capacity spill to a per-thread global overflow area (worst case = one span
// per primitive, reached only by rays that hit everything).  A span is 3 words:
// (start, end, tags) with tag = prim << 1 | flip for each boundary — normals and materials
// are recomputed from (prim, t) for the one boundary the tracer picks, instead of being copied
// through every merge as the reference does (include/span.h:93-120).
//
// The merges are the reference's streaming state machines, statement for statement, over
// cursors into the materialised child lists (src/union.cpp:84-134, src/difference.cpp:84-135,
// src/intersection.cpp:84-130).  Children are pure streams after init(), so materialising
// them first gives the identical output sequence.
#pragma once
#include "ptb_math.cuh"

namespace ptb
{

enum
{
    DK_SPHERE = 0,
    DK_PLANE = 1,
    DK_UNION = 2,
    DK_DIFFERENCE = 3,
    DK_INTERSECTION = 4,
    DK_END = 5
};
#define PTB_LEAF_REUSE_RAY 1u

struct SpanRec
{
    float t0, t1;
    uint32_t tags; // low 16: start tag, high 16: end tag; tag = prim << 1 | flip
};
__device__ __forceinline__ uint32_t startTag(const SpanRec &s) { return s.tags & 0xFFFFu; }
__device__ __forceinline__ uint32_t endTag(const SpanRec &s) { return s.tags >> 16; }
// include/span.h:93-120 — a start that becomes an end (or vice versa) flips its normal.
__device__ __forceinline__ void copyStartFromStart(SpanRec &d, const SpanRec &s) { d.t0 = s.t0, d.tags = (d.tags & 0xFFFF0000u) | startTag(s); }
__device__ __forceinline__ void copyEndFromStart(SpanRec &d, const SpanRec &s) { d.t1 = s.t0, d.tags = (d.tags & 0xFFFFu) | ((startTag(s) ^ 1u) << 16); }
__device__ __forceinline__ void copyStartFromEnd(SpanRec &d, const SpanRec &s) { d.t0 = s.t1, d.tags = (d.tags & 0xFFFF0000u) | (endTag(s) ^ 1u); }
__device__ __forceinline__ void copyEndFromEnd(SpanRec &d, const SpanRec &s) { d.t1 = s.t1, d.tags = (d.tags & 0xFFFFu) | (endTag(s) << 16); }

// Per-thread span stack.
struct Pool
{
    float *s0;        // &smem_t0[threadIdx.x]
    float *s1;        // &smem_t1[threadIdx.x]
    uint32_t *sg;     // &smem_tags[threadIdx.x]
    uint16_t *starts; // &smem_starts[threadIdx.x]: pool offset where each stack entry begins
    int stride;       // threads per block
    int cap;          // slots held in shared memory
    float *ovf;       // global overflow: [slot - cap][3][ovfStride] + thread, or NULL
    size_t ovfStride;

    __device__ __forceinline__ SpanRec load(int slot) const
    {
        SpanRec r;
        if(slot < cap)
        {
            r.t0 = s0[slot * stride];
            r.t1 = s1[slot * stride];
            r.tags = sg[slot * stride];
        }
        else
        {
            const float *p = ovf + (size_t)(slot - cap) * 3 * ovfStride;
            r.t0 = p[0];
            r.t1 = p[ovfStride];
            r.tags = __float_as_uint(p[2 * ovfStride]);
        }
        return r;
    }
    __device__ __forceinline__ void store(int slot, const SpanRec &r) const
    {
        if(slot < cap)
        {
            s0[slot * stride] = r.t0;
            s1[slot * stride] = r.t1;
            sg[slot * stride] = r.tags;
        }
        else
        {
            float *p = ovf + (size_t)(slot - cap) * 3 * ovfStride;
            p[0] = r.t0;
            p[ovfStride] = r.t1;
            p[2 * ovfStride] = __uint_as_float(r.tags);
        }
    }
};

// ---- primitives -----------------------------------------------------------------------------

// SphereSpanIterator::init, reference src/sphere.cpp:31-49 (normals deferred).
template <bool EX>
__device__ __forceinline__ bool sphereSpan(V3 o, V3 d, float a, float4 g, float &t0, float &t1)
{
    typedef Ar<EX> A;
    V3 oc = A::sub(o, mk(g.x, g.y, g.z));
    float b = A::dot(oc, d);
    float c = A::sub(A::dot(oc, oc), g.w);
    float disc = A::sub(A::mul(b, b), A::mul(a, c));
    if(disc <= PTB_EPS)
        return false;
    float s = A::sqrt(disc);
    if(EX)
    {
        t0 = __fdiv_rn(__fsub_rn(-b, s), a);
        t1 = __fdiv_rn(__fadd_rn(-b, s), a);
    }
    else
    {
        float ia = __fdividef(1.0f, a);
        t0 = (-b - s) * ia;
        t1 = (-b + s) * ia;
    }
    return true;
}

// PlaneSpanIterator::init, reference src/plane.cpp:35-63.
template <bool EX>
__device__ __forceinline__ bool planeSpan(V3 o, V3 d, float4 g, float &t0, float &t1)
{
    typedef Ar<EX> A;
    V3 n = mk(g.x, g.y, g.z);
    float divisor = A::dot(d, n);
    float numerator = A::sub(-g.w, A::dot(o, n));
    float t = 0.0f;
    bool parallel = fabsf(divisor) < PTB_EPS * PTB_EPS;
    if(!parallel)
    {
        t = A::div(numerator, divisor);
        parallel = fabsf(t) >= PTB_MAX_VALUE;
    }
    if(parallel)
    {
        if(!(fabsf(numerator) < PTB_EPS * PTB_EPS))
            return false;
        t0 = -PTB_MAX_VALUE;
        t1 = PTB_MAX_VALUE;
    }
    else if(divisor < 0.0f)
    {
        t0 = t;
        t1 = PTB_MAX_VALUE;
    }
    else
    {
        t0 = -PTB_MAX_VALUE;
        t1 = t;
    }
    return true;
}

// ---- merges ---------------------------------------------------------------------------------

template <class P>
struct MergeCursor
{
    const P &pool;
    int a, aEnd, b, bEnd, out;
    SpanRec spanA, spanB;
    bool aEnded, bEnded;
    // A = pool slots [a0, aEnd), B = [b0, bEnd); output from slot dst on
    __device__ __forceinline__ MergeCursor(const P &pool, int a0, int aEnd, int b0, int bEnd, int dst) : pool(pool), a(a0), aEnd(aEnd), b(b0), bEnd(bEnd), out(dst), aEnded(false), bEnded(false)
    {
        spanA.t0 = spanA.t1 = spanB.t0 = spanB.t1 = 0.0f;
        spanA.tags = spanB.tags = 0u;
    }
    // nextA()/nextB(): src/union.cpp:21-47 — take the child's current span, or mark it ended
    __device__ __forceinline__ void nextA()
    {
        if(a < aEnd)
            spanA = pool.load(a++);
        else
            aEnded = true;
    }
    __device__ __forceinline__ void nextB()
    {
        if(b < bEnd)
            spanB = pool.load(b++);
        else
            bEnded = true;
    }
    __device__ __forceinline__ void emit(const SpanRec &s) { pool.store(out++, s); }
};

// Runs one combinator to exhaustion, writing its output spans from `dst` on; returns the slot
// after the last output.
template <class P>
static __device__ __forceinline__ int mergeRangesT(const P &pool, int kind, int a0, int aEnd, int b0, int bEnd, int dst)
{
    MergeCursor<P> m(pool, a0, aEnd, b0, bEnd, dst);
    m.nextA();
    m.nextB();
    if(kind == DK_UNION)
    { // UnionSpanIterator::next, src/union.cpp:84-134
        for(;;)
        {
            if(m.aEnded)
            {
                if(m.bEnded)
                    break;
                m.emit(m.spanB);
                m.nextB();
                continue;
            }
            if(m.bEnded || m.spanA.t1 < m.spanB.t0)
            {
                m.emit(m.spanA);
                m.nextA();
                continue;
            }
            if(m.spanB.t1 < m.spanA.t0)
            {
                m.emit(m.spanB);
                m.nextB();
                continue;
            }
            if(m.spanA.t0 < m.spanB.t0)
            {
                if(m.spanA.t1 < m.spanB.t1)
                    copyEndFromEnd(m.spanA, m.spanB);
                m.nextB();
            }
            else
            {
                if(m.spanA.t1 > m.spanB.t1)
                    copyEndFromEnd(m.spanB, m.spanA);
                m.nextA();
            }
        }
    }
    else if(kind == DK_DIFFERENCE)
    { // DifferenceSpanIterator::next, src/difference.cpp:84-135 (Q4 at :124-131 reproduced)
        for(;;)
        {
            if(m.aEnded)
                break;
            if(m.bEnded || m.spanA.t1 < m.spanB.t0)
            {
                m.emit(m.spanA);
                m.nextA();
                continue;
            }
            if(m.spanB.t1 < m.spanA.t0)
                m.nextB();
            else if(m.spanA.t0 < m.spanB.t0)
            {
                if(m.spanA.t1 < m.spanB.t1)
                {
                    copyEndFromStart(m.spanA, m.spanB);
                    m.emit(m.spanA);
                    m.nextA();
                    continue;
                }
                SpanRec r = m.spanA;
                copyEndFromStart(r, m.spanB);
                m.emit(r);
                copyStartFromEnd(m.spanA, m.spanB);
                m.nextB();
            }
            else
            {
                if(m.spanA.t1 > m.spanB.t1)
                {
                    copyEndFromStart(m.spanA, m.spanB);
                    m.nextB();
                    continue;
                }
                m.nextA();
            }
        }
    }
    else
    { // IntersectionSpanIterator::next, src/intersection.cpp:84-130
        for(;;)
        {
            if(m.aEnded || m.bEnded)
                break;
            if(m.spanA.t1 < m.spanB.t0)
            {
                m.nextA();
                continue;
            }
            if(m.spanB.t1 < m.spanA.t0)
            {
                m.nextB();
                continue;
            }
            if(m.spanA.t0 < m.spanB.t0)
            {
                if(m.spanA.t1 < m.spanB.t1)
                {
                    copyStartFromStart(m.spanA, m.spanB);
                    m.emit(m.spanA);
                    m.nextA();
                    continue;
                }
                m.emit(m.spanB);
                m.nextB();
            }
            else
            {
                if(m.spanB.t1 < m.spanA.t1)
                {
                    copyStartFromStart(m.spanB, m.spanA);
                    m.emit(m.spanB);
                    m.nextB();
                    continue;
                }
                m.emit(m.spanA);
                m.nextA();
            }
        }
    }
    return m.out;
}

// The two lists on top of an operand stack: A = [a0, b0), B = [b0, top).
template <class P>
static __device__ __forceinline__ int mergeListsT(const P &pool, int kind, int a0, int b0, int top, int dst)
{
    return mergeRangesT(pool, kind, a0, b0, b0, top, dst);
}

static __device__ __noinline__ int mergeLists(const Pool &pool, int kind, int a0, int b0, int top, int dst) { return mergeListsT(pool, kind, a0, b0, top, dst); }

// ---- program evaluation ---------------------------------------------------------------------

// Runs the whole program for one ray; on return the root span list is pool slots [0, return).
// `prog` points at 16-byte words (shared or global memory).
template <bool EX>
__device__ __forceinline__ int evalProgram(const uint4 *__restrict__ prog, const Pool &pool, V3 o, V3 d)
{
    typedef Ar<EX> A;
    int top = 0, depth = 0;
    V3 lo = o, ld = d;             // ray in the current leaf's space
    float la = A::dot(d, d);       // abs_squared(dir) of that ray (src/sphere.cpp:34)
    float wa = la;
    bool atWorld = true;
    const uint4 *pc = prog;
    for(;;)
    {
        uint4 h = *pc++;
        uint32_t kind = h.x;
        if(kind == DK_END)
            break;
        if(kind <= DK_PLANE)
        {
            float4 g = *reinterpret_cast<const float4 *>(pc++);
            uint32_t nchain = h.z;
            if(!(h.w & PTB_LEAF_REUSE_RAY))
            {
                if(nchain == 0)
                {
                    if(!atWorld)
                        lo = o, ld = d, la = wa, atWorld = true;
                }
                else
                {
                    lo = o, ld = d;
                    for(uint32_t k = 0; k < nchain; k++)
                    { // transform(m, ray), include/transform.h:429-432, once per wrapper
                        float4 r0 = *reinterpret_cast<const float4 *>(pc);
                        float4 r1 = *reinterpret_cast<const float4 *>(pc + 1);
                        float4 r2 = *reinterpret_cast<const float4 *>(pc + 2);
                        pc += 3;
                        lo = A::apply(r0, r1, r2, lo);
                        ld = A::applyNoTranslate(r0, r1, r2, ld);
                    }
                    la = A::dot(ld, ld);
                    atWorld = false;
                }
            }
            pool.starts[depth * pool.stride] = (uint16_t)top;
            depth++;
            float t0, t1;
            bool hit = kind == DK_SPHERE ? sphereSpan<EX>(lo, ld, la, g, t0, t1) : planeSpan<EX>(lo, ld, g, t0, t1);
            if(hit)
            {
                SpanRec s;
                s.t0 = t0, s.t1 = t1;
                uint32_t tag = h.y << 1;
                s.tags = tag | (tag << 16);
                pool.store(top++, s);
            }
            continue;
        }
        // combinator: B is the top stack entry, A the one below (postfix order a, b, op)
        int b0 = pool.starts[(depth - 1) * pool.stride];
        int a0 = pool.starts[(depth - 2) * pool.stride];
        depth--;
        int la_ = b0 - a0, lb_ = top - b0;
        if(la_ == 0 || lb_ == 0)
        {
            // union with an empty side is the other side unchanged; A minus nothing is A; nothing
            // minus B, and any intersection with nothing, is empty (same as running the state
            // machines with aEnded/bEnded set from the start)
            if(kind != DK_UNION && (la_ == 0 || kind == DK_INTERSECTION))
                top = a0;
            continue;
        }
        if(la_ == 1 && lb_ == 1)
            top = mergeLists(pool, (int)kind, a0, b0, top, a0); // both inputs are in registers before the first output
        else
        {
            int end = mergeLists(pool, (int)kind, a0, b0, top, top);
            int n = end - top;
            for(int i = 0; i < n; i++)
                pool.store(a0 + i, pool.load(top + i));
            top = a0 + n;
        }
    }
    return top;
}

// First boundary with t >= eps: the scan of traceRay, reference include/path-trace.h:66-101.
// Returns PTB_HIT_* (0 miss, 1 entered at a span start, 2 left at a span end).
template <class P>
__device__ __forceinline__ int firstHit(const P &pool, int n, float &t, uint32_t &tag)
{
    for(int i = 0; i < n; i++)
    {
        SpanRec s = pool.load(i);
        if(s.t0 >= PTB_MAX_VALUE)
            return 0;
        if(s.t0 >= PTB_EPS)
        {
            t = s.t0, tag = startTag(s);
            return 1;
        }
        if(s.t1 >= PTB_MAX_VALUE)
            return 0;
        if(s.t1 >= PTB_EPS)
        {
            t = s.t1, tag = endTag(s);
            return 2;
        }
    }
    return 0;
}

} // namespace ptb
