// ptb_csg_fast.cuh — the span-list CSG evaluator the renderer runs (fast arithmetic policy).
//
// Same algorithm and the same merge state machines as ptb_csg.cuh (the bit-exact evaluator),
// re-laid-out for issue rate.  ncu on the generic evaluator (profiles/r01a_full_c4.txt): 12 100
// warp instructions per ray on the 70-primitive scene, only 20 % of them FP32 arithmetic — the
// rest was generic-address LD/ST, 64-bit address arithmetic, per-op dispatch and local memory.
//
// The host (ptb_capi.cu, compileFastProgram) turns the postfix program into a stream of
// 16-byte words in which
//   * every TransformedObject chain is pre-multiplied (in double) into ONE 3x4 matrix, and the
//     sphere centre is folded into its translation:  oc = M o + (t - c)   (3 FFMA per row);
//   * a transformed plane becomes a world-space plane: n' = M^T n, D' = D + n.t, so a plane
//     costs 6 FFMA whatever wraps it;
//   * the operator that follows a leaf rides in the leaf's header as its `action`, further
//     operators as 2-bit codes: no fetch/dispatch per operator;
//   * a SKIP word sits before the right operand of every Difference / Intersection: when no
//     lane of the warp has a non-empty left operand the whole right subtree is jumped over
//     ({} \ B = {} n B = {}: src/difference.cpp:88-89, src/intersection.cpp:88-90).
//
// EMPTY LISTS COST NOTHING.  Most leaves miss, so most lists on the operand stack are empty.
// The stack is therefore one bit per entry (`bits`, bit 0 = top of stack: list non-empty?) and
// only non-empty lists own pool slots and an entry on the start stack.  A u {} = A, A \ {} = A,
// {} \ B = {}, A n {} = {} are then pure bit arithmetic — exactly what the reference's state
// machines produce when aEnded / bEnded is set from the start (src/union.cpp:88-102,
// src/difference.cpp:88-99, src/intersection.cpp:88-90) — and the real streaming merges
// (mergeListsT, shared with the exact evaluator) only run for lanes where both lists have spans.
// The tree shape, the merge rules and hence span topology are untouched; only float rounding
// differs from the reference (|dt|/t ~ 1e-7, gate 1e-5).
//
// Shared memory is addressed with 32-bit shared-window addresses through ld.shared/st.shared
// (no generic pointers, no per-access window-base recomputation).  Layout per CTA:
//     [ program words ][ span pool: cap x {t0,t1,tags} x FT ][ start stack: depth x FT ]
// Pool and stack are [slot][thread]: conflict-free for any mix of slots across a warp.
#pragma once
#include "ptb_csg.cuh"

namespace ptb
{

#ifndef PTB_HOST_EMU
extern __shared__ uint4 ptbSmem[];
#endif

// header.x: bits 0-3 one-hot kind (cheap bit tests, no jump table), 4-5 leaf action / control
// sub-kind, 6 = shift the bit stack when the leaf misses (action == PUSH), 7 = the word has a tail
// (operators to run, or an Intersection action), 8-11 nPost, 12-29 up to 9 operators (2 bits each),
// 30 = union-only tail, 31 = leaf pair.  Sphere headers: z = -r^2, w = |t - c|^2 - r^2 (the constant
// term for rays from the origin, i.e. all camera rays: include/path-trace.h:196)
enum
{
    FK_SPHERE_XF = 1, // header {x, tags, -r^2, 0}, 3 rows of the folded matrix
    FK_PLANE = 2,     // header {x, tags, extra planes k, words to jump}, float4 (n.xyz, D) in world space, then k x {{tags}, plane}
    FK_SPHERE = 4,    // header {x, tags, -r^2, 0}, float4 (-c.xyz, 0)
    FK_CONTROL = 8,
    FK_END = 8,              // control sub-kind 0
    FK_OPS = 8 | (1 << 4),   // header only: operators that could not ride on the previous leaf
    FK_SKIP = 8 | (2 << 4),  // header only: y = words to jump when the top list is empty warp-wide
    FH_MISS_SHIFT = 1 << 6,
    FH_TAIL = 1 << 7,
    FH_UNION_TAIL = 1 << 30, // every operator of the tail is a Union (and the action is not Intersection)
    FH_RUN = 1 << 30,        // SPHERE_XF WITHOUT a tail: first of a run of (x >> 12) & 15 adjacent leaves forming a whole Union subtree;
                             // bits 16-29: word index of the run's bounding spheres {C.xyz, R^2}, one word a leaf, behind FK_END (0: none)
    FH_PAIR = 1u << 31       // SPHERE_XF without tail directly followed by another SPHERE_XF leaf
};
enum
{
    FA_PUSH = 0, // no operator follows the leaf directly
    FA_UNION = 1,
    FA_DIFFERENCE = 2,
    FA_INTERSECTION = 3
};
#define PTB_FAST_THREADS 128
// out-of-line or inline slow paths (measured on B200)
#ifndef PTB_SLOW_CALLS
#define PTB_SLOW_CALLS 0 /* bit 0: fastMerge out of line, bit 1: fastLeafHit / fastOpSlow out of line; 0 (all inline) measured fastest: C5 8.05 vs 6.87 Grays/s with 3 */
#endif
#if PTB_SLOW_CALLS & 1
#define PTB_MERGE_FN static __device__ __noinline__
#else
#define PTB_MERGE_FN __device__ __forceinline__
#endif
#if PTB_SLOW_CALLS & 2
#define PTB_SLOW_FN static __device__ __noinline__
#else
#define PTB_SLOW_FN __device__ __forceinline__
#endif
#define PTB_FAST_MAX_POST 9
#ifndef PTB_PLANE_SELECT
#define PTB_PLANE_SELECT 0
#endif
#ifndef PTB_INPLACE_A1
#define PTB_INPLACE_A1 1
#endif
#ifndef PTB_TAIL_CLOSED
#define PTB_TAIL_CLOSED 1
#endif
#define PTB_FAST_MAX_STACK 32

#ifdef PTB_HOST_EMU
static inline uint4 lds128(uint32_t a) { uint4 v; memcpy(&v, PTB_EMU_SMEM + a, 16); return v; }
static inline uint32_t lds32(uint32_t a) { uint32_t v; memcpy(&v, PTB_EMU_SMEM + a, 4); return v; }
static inline void sts32(uint32_t a, uint32_t v) { memcpy(PTB_EMU_SMEM + a, &v, 4); }
static inline void sts128(uint32_t a, uint4 v) { memcpy(PTB_EMU_SMEM + a, &v, 16); }
#else
__device__ __forceinline__ uint4 lds128(uint32_t a)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) { asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }

#endif

// Span pool in shared memory with a global overflow area for slots >= cap.
struct FastPool
{
    uint32_t base; // shared address of slot 0 for this thread
    int cap;
    float *ovf; // global overflow, already offset to this thread, or NULL
    size_t ovfStride;

    __device__ __forceinline__ SpanRec load(int slot) const
    {
        SpanRec r;
        if(slot < cap)
        {
            uint32_t a = base + (uint32_t)slot * (12u * PTB_FAST_THREADS);
            r.t0 = __uint_as_float(lds32(a));
            r.t1 = __uint_as_float(lds32(a + 4u * PTB_FAST_THREADS));
            r.tags = lds32(a + 8u * PTB_FAST_THREADS);
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
            uint32_t a = base + (uint32_t)slot * (12u * PTB_FAST_THREADS);
            sts32(a, __float_as_uint(r.t0));
            sts32(a + 4u * PTB_FAST_THREADS, __float_as_uint(r.t1));
            sts32(a + 8u * PTB_FAST_THREADS, r.tags);
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

struct FastCtx
{
    uint32_t prog;       // shared address of the first program word (unused with a global program)
    uint32_t stk;        // shared address of this thread's start stack (ray 0)
    uint32_t stkStride;  // bytes from ray r's start stack to ray r+1's
    uint32_t poolStride; // bytes from ray r's pool to ray r+1's
    size_t ovfRay;       // floats from ray r's overflow area to ray r+1's
    uint32_t tab;        // shared address of the constant table of a scene-specialised module (ptb_csg_reg.cuh), behind pools and stacks
    FastPool pool;       // ray 0's pool
    __device__ __forceinline__ FastPool poolOf(int r) const
    {
        FastPool p = pool;
        p.base += (uint32_t)r * poolStride;
        p.ovf += (size_t)r * ovfRay;
        return p;
    }
};

#ifdef PTB_HOST_EMU
static inline float rcpApprox(float x) { return 1.0f / x; }
static inline float sqrtApprox(float x) { return sqrtf(x); }
#else
__device__ __forceinline__ float rcpApprox(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float sqrtApprox(float x)
{
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
#endif

// Evaluator state of one ray.
struct FastState
{
    uint32_t bits; // operand stack, one bit per entry, bit 0 = top: does the list have spans?
    int top;       // pool slots in use
    uint32_t sp;   // shared address one past the top of the start stack (non-empty lists only)
};

// Real merge of the two top NON-EMPTY lists A = [a0, b0), B = [b0, top); returns the new top.
// One out-of-line copy of the three state machines, everything passed by value (registers).
PTB_MERGE_FN int fastMerge(uint32_t poolBase, int cap, float *ovf, size_t ovfStride, uint32_t op, int a0, int b0, int top)
{
    FastPool pool;
    pool.base = poolBase, pool.cap = cap, pool.ovf = ovf, pool.ovfStride = ovfStride;
    const int kind = (int)op - FA_UNION + DK_UNION;
    if(op == FA_UNION && b0 - a0 == 1 && top - b0 == 1)
    { // the commonest real merge (two single spans), UnionSpanIterator::next unrolled: src/union.cpp:104-133
        SpanRec A = pool.load(a0), B = pool.load(b0);
        if(A.t1 < B.t0)
            return top; // disjoint, already in order
        if(B.t1 < A.t0)
        {
            pool.store(a0, B);
            pool.store(b0, A);
            return top;
        }
        if(A.t0 < B.t0)
        {
            if(A.t1 < B.t1)
                copyEndFromEnd(A, B);
            pool.store(a0, A);
        }
        else
        {
            if(A.t1 > B.t1)
                copyEndFromEnd(B, A);
            pool.store(a0, B);
        }
        return a0 + 1;
    }
    // A single-span A is in a register before the first output is written, and of B only spans already
    // read can be overwritten (output k is written after k spans were read, and A's slot comes first):
    // such merges run in place.  (Difference may split A in two, so it needs a single-span B as well.)
    // Otherwise the output is appended behind B and moved down.  One copy of the state machines either way.
    const bool inPlace = b0 - a0 == 1 && (top - b0 == 1 || (PTB_INPLACE_A1 && (op == FA_UNION || op == FA_INTERSECTION)));
    const int dst = inPlace ? a0 : top;
    const int end = mergeListsT(pool, kind, a0, b0, top, dst);
    if(inPlace)
        return end;
    const int n = end - top;
    for(int i = 0; i < n; i++)
        pool.store(a0 + i, pool.load(top + i));
    return a0 + n;
}

// The slow paths are out of line and take / return everything by value (registers), so the
// hot loop stays a few hundred contiguous instructions (instruction-cache resident).
//
// Operator on the two top entries (ab: bit 1 = A, bit 0 = B) for lanes that need pool work.
// Returns {result bit, top, sp}.
PTB_SLOW_FN uint3 fastOpSlow(uint32_t poolBase, int cap, float *ovf, size_t ovfStride, uint32_t op, uint32_t ab, int top, uint32_t sp)
{
    const uint32_t W = 4u * PTB_FAST_THREADS;
    if(ab == 3u)
    {
        int b0 = (int)lds32(sp - W), a0 = (int)lds32(sp - 2u * W);
        top = fastMerge(poolBase, cap, ovf, ovfStride, op, a0, b0, top);
        if(top == a0) // the result has no spans: it gives up its start-stack entry too
            return make_uint3(0u, (uint32_t)top, sp - 2u * W);
        return make_uint3(1u, (uint32_t)top, sp - W);
    }
    // exactly one operand has spans and the operator throws them away
    sp -= W;
    return make_uint3(0u, lds32(sp), sp);
}

// One operator on the bit stack; pool work only where needed.
__device__ __forceinline__ void fastOp(const FastPool &pool, uint32_t op, FastState &st)
{
    const uint32_t ab = st.bits & 3u;
    bool work;
    uint32_t res;
    if(op == FA_UNION)
        work = ab == 3u, res = ab != 0u;
    else if(op == FA_DIFFERENCE)
        work = (ab & 1u) != 0u, res = ab >> 1; // B has spans: real merge (ab == 3) or B discarded (ab == 1)
    else
        work = ab != 0u, res = 0u; // both: real merge; one: discarded
    if(work)
    {
        uint3 r = fastOpSlow(pool.base, pool.cap, pool.ovf, pool.ovfStride, op, ab, st.top, st.sp);
        res = r.x, st.top = (int)r.y, st.sp = r.z;
    }
    st.bits = ((st.bits >> 2) << 1) | res;
}

// Discriminant of one folded transformed sphere: a = |d'|^2, b = oc.d', disc = b^2 - a c
// (src/sphere.cpp:33-38 on the ray transformed by the folded matrix).  O0: the ray starts at the
// world origin (every camera ray), so oc is the matrix translation and c is the header constant.
template <bool O0>
__device__ __forceinline__ void xfDisc(const uint4 &h, const uint4 &u0, const uint4 &u1, const uint4 &u2, V3 o, V3 d, float &a, float &b, float &disc)
{
    const float m00 = __uint_as_float(u0.x), m01 = __uint_as_float(u0.y), m02 = __uint_as_float(u0.z), m03 = __uint_as_float(u0.w);
    const float m10 = __uint_as_float(u1.x), m11 = __uint_as_float(u1.y), m12 = __uint_as_float(u1.z), m13 = __uint_as_float(u1.w);
    const float m20 = __uint_as_float(u2.x), m21 = __uint_as_float(u2.y), m22 = __uint_as_float(u2.z), m23 = __uint_as_float(u2.w);
    float dx = fmaf(d.x, m00, fmaf(d.y, m01, d.z * m02));
    float dy = fmaf(d.x, m10, fmaf(d.y, m11, d.z * m12));
    float dz = fmaf(d.x, m20, fmaf(d.y, m21, d.z * m22));
    a = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
    float cc;
    if(O0)
    {
        b = fmaf(m03, dx, fmaf(m13, dy, m23 * dz));
        cc = __uint_as_float(h.w);
    }
    else
    {
        float ox = fmaf(o.x, m00, fmaf(o.y, m01, fmaf(o.z, m02, m03)));
        float oy = fmaf(o.x, m10, fmaf(o.y, m11, fmaf(o.z, m12, m13)));
        float oz = fmaf(o.x, m20, fmaf(o.y, m21, fmaf(o.z, m22, m23)));
        b = fmaf(ox, dx, fmaf(oy, dy, oz * dz));
        cc = fmaf(ox, ox, fmaf(oy, oy, fmaf(oz, oz, __uint_as_float(h.z))));
    }
    disc = fmaf(b, b, -(a * cc));
}

// One plane of a plane record folded into the running {(-inf, loEnd], [upStart, +inf)} pair:
// PlaneSpanIterator::init, src/plane.cpp:35-63, on a world-space plane u0 = (n.xyz, D).
template <bool O0>
__device__ __forceinline__ void planeFold(const uint4 &u0, uint32_t tags, V3 o, V3 d, float &loEnd, float &upStart, uint32_t &loTags, uint32_t &upTags, bool &full)
{
    const float nx = __uint_as_float(u0.x), ny = __uint_as_float(u0.y), nz = __uint_as_float(u0.z);
    float divisor = fmaf(d.x, nx, fmaf(d.y, ny, d.z * nz));
    float numerator = O0 ? -__uint_as_float(u0.w) : -fmaf(o.x, nx, fmaf(o.y, ny, fmaf(o.z, nz, __uint_as_float(u0.w))));
    float t = numerator * rcpApprox(divisor);
    bool parallel = fabsf(divisor) < PTB_EPS * PTB_EPS || !(fabsf(t) < PTB_MAX_VALUE);
    if(parallel)
    {
        if(fabsf(numerator) < PTB_EPS * PTB_EPS)
            full = true, loTags = upTags = tags; // the ray lies in the plane: the whole line
    }
    else if(divisor < 0.0f)
    { // [t, +inf): the earliest start wins and brings both its tags (src/union.cpp:113-123)
        if(t <= upStart)
            upStart = t, upTags = tags;
    }
    else
    { // (-inf, t]: equal starts keep the later operand's start, the larger end its tag (:124-133)
        loTags = (loTags & 0xFFFF0000u) | (tags & 0xFFFFu);
        if(t >= loEnd)
            loEnd = t, loTags = (loTags & 0xFFFFu) | (tags & 0xFFFF0000u);
    }
}
// The folded pair as the record's leaf: nothing, one span, or two spans in order.
__device__ __forceinline__ void planeFinish(float loEnd, float upStart, uint32_t loTags, uint32_t upTags, bool full, bool &hit, bool &two, float &lt0, float &lt1, float &ut0,
                                            uint32_t &ltags, uint32_t &utags)
{
    const bool hasLo = loEnd > -3.0e38f, hasUp = upStart < 3.0e38f;
    hit = full || hasLo || hasUp;
    two = false;
    lt0 = -PTB_MAX_VALUE, lt1 = PTB_MAX_VALUE, ltags = upTags; // the whole line (a ray in the plane)
    ut0 = upStart, utags = upTags;
    if(full)
        ;
    else if(loEnd >= upStart) // two planes only: either order gives the whole line (src/union.cpp:113-133)
        ltags = (loTags & 0xFFFFu) | (upTags & 0xFFFF0000u);
    else if(hasLo)
        lt1 = loEnd, ltags = loTags, two = hasUp;
    else
        lt0 = upStart;
}
// Discriminant of an untransformed sphere: header h = {.., .., -r^2, |c|^2 - r^2}, u0 = (-c.xyz, 0); src/sphere.cpp:33-38.
template <bool O0>
__device__ __forceinline__ void sphereDisc(const uint4 &h, const uint4 &u0, V3 o, V3 d, float &a, float &b, float &e)
{
    a = fmaf(d.x, d.x, fmaf(d.y, d.y, d.z * d.z));
    float cc;
    if(O0)
    {
        b = fmaf(__uint_as_float(u0.x), d.x, fmaf(__uint_as_float(u0.y), d.y, __uint_as_float(u0.z) * d.z));
        cc = __uint_as_float(h.w);
    }
    else
    {
        float ox = o.x + __uint_as_float(u0.x), oy = o.y + __uint_as_float(u0.y), oz = o.z + __uint_as_float(u0.z);
        b = fmaf(ox, d.x, fmaf(oy, d.y, oz * d.z));
        cc = fmaf(ox, ox, fmaf(oy, oy, fmaf(oz, oz, __uint_as_float(h.z))));
    }
    e = fmaf(b, b, -(a * cc));
}

// What one program word does to one ray's operand stack once its leaf (if any) has been evaluated:
// push the leaf's list (one span; two for a plane cluster) or an empty one, then run the word's
// operators.  hx = the word's header, see the enum above.
__device__ __forceinline__ void fastPushOperate(uint32_t hx, bool hit, bool two, float lt0, float lt1, float ut0, uint32_t ltags, uint32_t utags, FastState &st,
                                                const FastPool &pool)
{
    const uint32_t W = 4u * PTB_FAST_THREADS;
    const bool isLeaf = !(hx & FK_CONTROL);
    const uint32_t action = (hx >> 4) & 3u, npost = (hx >> 8) & 15u;
    const bool tail = (hx & (FH_TAIL | FK_CONTROL)) != 0u, isPlane = (hx & FK_PLANE) != 0u;
    if(!tail && !hit)
    { // a missed leaf with nothing behind it is an empty list, or nothing at all
        st.bits <<= (hx >> 6) & 1u;
        return;
    }
    uint32_t first = 0u; // the leaf's action, if it has to run as an operator on two lists
    if(isLeaf)
    {
        if(!hit)
        {
            st.bits <<= (hx >> 6) & 1u; // a pushed leaf that missed is an empty list
            if(action == FA_INTERSECTION && (st.bits & 1u))
            { // A n {} = {}: a missed leaf drops the spans of the list below
                st.sp -= W;
                st.top = (int)lds32(st.sp);
                st.bits &= ~1u;
            }
        }
        else if(action <= FA_UNION || (st.bits & 1u))
        { // push the leaf's list ({} \ B and {} n B stay empty: B is never stored)
            SpanRec s;
            s.t0 = lt0, s.t1 = lt1, s.tags = ltags;
            sts32(st.sp, (uint32_t)st.top);
            pool.store(st.top, s);
            const bool second = isPlane && two;
            if(second)
            {
                s.t0 = ut0, s.t1 = PTB_MAX_VALUE, s.tags = utags;
                pool.store(st.top + 1, s);
            }
            const bool replaces = action == FA_UNION && !(st.bits & 1u); // {} u B = B: B takes the empty list's entry
            st.top += second ? 2 : 1, st.sp += W, st.bits = replaces ? st.bits | 1u : (st.bits << 1) | 1u;
            first = replaces ? 0u : action;
        }
    }
    if(first == 0u)
    {
        if(npost == 0u)
            return;
        const uint32_t involved = st.bits & ((2u << npost) - 1u);
        if(PTB_TAIL_CLOSED && (hx & FH_UNION_TAIL) && __popc(involved) <= 1)
        { // a run of Unions: the lists involved collapse into one; spans move only if two of them have any
            st.bits = ((st.bits >> npost) & ~1u) | (involved ? 1u : 0u);
            return;
        }
    }
    uint32_t ops = hx >> 12, left = npost, op = first;
    if(op == 0u)
        op = (ops & 3u) + FA_UNION, ops >>= 2, left--;
    for(;;)
    {
        fastOp(pool, op, st);
        if(left == 0u)
            break;
        op = (ops & 3u) + FA_UNION, ops >>= 2, left--;
    }
}

// Runs the whole program for R rays per lane, all 32 lanes of the warp together (SKIP words
// vote).  Ray r's root span list ends up in pool r, slots [0, topOut[r]).  (R = 2 was measured
// slower than R = 1 on B200 — occupancy beats per-thread ILP here — and is kept for experiments.)
template <bool GP, bool O0, int R>
__device__ __forceinline__ void evalFast(const FastCtx &c, const uint4 *__restrict__ gprog, const V3 (&o)[R], const V3 (&d)[R], int (&topOut)[R])
{
    FastState st[R];
#pragma unroll
    for(int r = 0; r < R; r++)
        st[r].bits = 0u, st[r].top = 0, st[r].sp = c.stk + (uint32_t)r * c.stkStride;
    uint32_t pc = GP ? 0u : c.prog; // GP: word index; else shared byte address
    uint32_t boundMask = 0u;        // warp-uniform: which of the coming leaves of a sphere run can be hit (> 1: inside a run)
#define PTB_PROG(i) (GP ? __ldg(gprog + pc + (i)) : lds128(pc + 16u * (i)))
#define PTB_ADVANCE(n) (pc += GP ? (uint32_t)(n) : 16u * (uint32_t)(n))
// The commonest trips by far, taken by the whole warp together (votes), so that the loop stays
// converged for the SKIP votes: no ray of the warp hits the leaf, and either no operator follows it in
// its word, or a run of Unions does that moves no spans (at most one of the lists involved has any).
#define PTB_MISS_EXIT()                                                                                      \
    {                                                                                                        \
        bool anyHit = false;                                                                                 \
        _Pragma("unroll") for(int r = 0; r < R; r++) anyHit = anyHit || hit[r];                              \
        if(!__any_sync(0xFFFFFFFFu, anyHit))                                                                 \
        {                                                                                                    \
            if(!(h.x & FH_TAIL))                                                                             \
            {                                                                                                \
                _Pragma("unroll") for(int r = 0; r < R; r++) st[r].bits <<= (h.x >> 6) & 1u;                 \
                continue;                                                                                    \
            }                                                                                                \
            if(PTB_TAIL_CLOSED && (h.x & FH_UNION_TAIL))                                                     \
            { /* the leaf's own action is Union or Difference: a miss changes nothing */                    \
                const uint32_t np = (h.x >> 8) & 15u, mask = (2u << np) - 1u;                                \
                bool simple = true;                                                                          \
                _Pragma("unroll") for(int r = 0; r < R; r++) simple = simple && __popc(st[r].bits & mask) <= 1; \
                if(__all_sync(0xFFFFFFFFu, simple))                                                          \
                {                                                                                            \
                    _Pragma("unroll") for(int r = 0; r < R; r++)                                             \
                        st[r].bits = ((st[r].bits >> np) & ~1u) | ((st[r].bits & mask) ? 1u : 0u);           \
                    continue;                                                                                \
                }                                                                                            \
            }                                                                                                \
        }                                                                                                    \
    }
    for(;;)
    {
        uint4 h = PTB_PROG(0);
        // what the leaf in this word produced: nothing, one span, or (plane clusters) two in order
        bool hit[R], two[R];
        float lt0[R], lt1[R], ut0[R];
        uint32_t ltags[R], utags[R];
#pragma unroll
        for(int r = 0; r < R; r++)
            ltags[r] = h.y; // (the rest is set by the branch that sets hit[r]; two[r], ut0[r], utags[r] by planes only)
        if(h.x & FK_SPHERE_XF)
        {
            bool skipLeaf = false; // warp-uniform: no ray's line meets this leaf's bounding sphere
            if(O0 && (h.x & (FH_RUN | FH_TAIL)) == FH_RUN && ((h.x >> 16) & 0x3FFFu) != 0u)
            { // Camera rays, first leaf of a run whose world-space bounding spheres follow the program
              // (compileFastProgram): one 16-byte word and 7 instructions per leaf instead of four words and ~32
              // decide which leaves some ray of the warp can hit at all.  A line through the origin along d meets the
              // sphere (C, R) iff (C.d)^2 >= |d|^2 (|C|^2 - R^2); the word holds C and sqrt(|C|^2 - R^2), rounded
              // down.  A line that misses a leaf's bounding sphere misses the leaf, at any t: skipping it is treating
              // it as the miss it is.  (ncu, C4: the 48 spheres cost 1 530 of a camera ray's 3 220 instructions to find
              // that the warp's two pixels see one or none of them, profiles/r02k_full_c4.txt.  For queued rays — 32
              // unrelated rays a warp, any origin — the vote seldom frees a leaf and the test costs what it saves.)
                const uint32_t n = (h.x >> 12) & 15u, bi = (h.x >> 16) & 0x3FFFu;
                float sd[R];
                uint32_t met[R];
#pragma unroll
                for(int r = 0; r < R; r++)
                    sd[r] = Ar<false>::sqrt(fmaf(d[r].x, d[r].x, fmaf(d[r].y, d[r].y, d[r].z * d[r].z))), met[r] = 0u;
                for(uint32_t k = 0; k < n; k++)
                {
                    const uint4 bw = GP ? __ldg(gprog + bi + k) : lds128(c.prog + 16u * (bi + k));
#pragma unroll
                    for(int r = 0; r < R; r++)
                    {
                        const float bb = fmaf(__uint_as_float(bw.x), d[r].x, fmaf(__uint_as_float(bw.y), d[r].y, __uint_as_float(bw.z) * d[r].z));
                        met[r] |= fabsf(bb) >= sd[r] * __uint_as_float(bw.w) ? 1u << k : 0u;
                    }
                }
                uint32_t mask = 0u;
#pragma unroll
                for(int r = 0; r < R; r++)
                    mask |= met[r];
                mask = __reduce_or_sync(0xFFFFFFFFu, mask);
                if(mask == 0u)
                { // no leaf can be hit: the subtree is one empty list
#pragma unroll
                    for(int r = 0; r < R; r++)
                        st[r].bits <<= 1;
                    PTB_ADVANCE(4u * n);
                    continue;
                }
                boundMask = mask | (1u << n); // bit k: leaf k of the run may be hit; the top bit marks the end
            }
            else if(O0 && (h.x & (FH_RUN | FH_TAIL)) == FH_RUN)
            { // camera rays: the whole Union subtree of spheres in one loop and one vote (see compileFastProgram).
              // ncu, C4: a leaf costs ~49 instructions one by one, 17 of them arithmetic; ~26 here.
                const uint32_t n = (h.x >> 12) & 15u;
                bool any = false;
                for(uint32_t k = 0; k < n; k++)
                {
                    const uint4 g = PTB_PROG(4u * k), v0 = PTB_PROG(4u * k + 1u), v1 = PTB_PROG(4u * k + 2u), v2 = PTB_PROG(4u * k + 3u);
#pragma unroll
                    for(int r = 0; r < R; r++)
                    {
                        float ra, rb, re;
                        xfDisc<O0>(g, v0, v1, v2, o[r], d[r], ra, rb, re);
                        any = any || re > PTB_EPS;
                    }
                }
                if(!__any_sync(0xFFFFFFFFu, any))
                { // every leaf missed: the subtree is one empty list
#pragma unroll
                    for(int r = 0; r < R; r++)
                        st[r].bits <<= 1;
                    PTB_ADVANCE(4u * n);
                    continue;
                }
            }
            if(boundMask > 1u)
                skipLeaf = !(boundMask & 1u), boundMask >>= 1;
            float a[R], b[R], e[R];
            if(skipLeaf)
            {
#pragma unroll
                for(int r = 0; r < R; r++)
                    a[r] = 1.f, b[r] = 0.f, e[r] = -1.f;
                PTB_ADVANCE(4);
            }
            else if((h.x & FH_PAIR) && boundMask <= 1u)
            { // the next word is a transformed sphere too and this one has no tail: both discriminants
              // are computed together (ILP).  If no ray of the warp hits the first, it is done with and
              // the second becomes the leaf of this trip; otherwise the second waits for the next trip.
                const uint4 u0 = PTB_PROG(1), u1 = PTB_PROG(2), u2 = PTB_PROG(3);
                const uint4 g = PTB_PROG(4), v0 = PTB_PROG(5), v1 = PTB_PROG(6), v2 = PTB_PROG(7);
                float a1[R], b1[R], e1[R];
                bool any = false;
#pragma unroll
                for(int r = 0; r < R; r++)
                {
                    xfDisc<O0>(h, u0, u1, u2, o[r], d[r], a[r], b[r], e[r]);
                    xfDisc<O0>(g, v0, v1, v2, o[r], d[r], a1[r], b1[r], e1[r]);
                    any = any || e[r] > PTB_EPS;
                }
                PTB_ADVANCE(4);
                if(!__any_sync(0xFFFFFFFFu, any))
                {
#pragma unroll
                    for(int r = 0; r < R; r++)
                    {
                        st[r].bits <<= (h.x >> 6) & 1u; // (a leaf under Intersection has a tail: never first of a pair)
                        a[r] = a1[r], b[r] = b1[r], e[r] = e1[r];
                        ltags[r] = g.y;
                    }
                    h = g;
                    PTB_ADVANCE(4);
                }
            }
            else
            {
                const uint4 u0 = PTB_PROG(1), u1 = PTB_PROG(2), u2 = PTB_PROG(3);
#pragma unroll
                for(int r = 0; r < R; r++)
                    xfDisc<O0>(h, u0, u1, u2, o[r], d[r], a[r], b[r], e[r]);
                PTB_ADVANCE(4);
            }
#pragma unroll
            for(int r = 0; r < R; r++)
                hit[r] = e[r] > PTB_EPS;
            PTB_MISS_EXIT();
#pragma unroll
            for(int r = 0; r < R; r++)
                if(hit[r])
                { // src/sphere.cpp:44-46
                    float q = sqrtApprox(e[r]), ia = rcpApprox(a[r]);
                    lt0[r] = (-b[r] - q) * ia, lt1[r] = (-b[r] + q) * ia;
                }
        }
        else if(h.x & FK_PLANE)
        { // PlaneSpanIterator::init, src/plane.cpp:35-63, on world-space planes.  A record holds 1 + h.z
          // planes: one leaf (h.z == 0), or the planes of a subtree made of Unions of planes only,
          // whose own words follow the record.  Every plane yields a half-line (-inf, t] or [t, +inf);
          // as long as every lower end lies below every upper start their Union is, whatever the
          // tree's shape, {(-inf, largest lower end], [smallest upper start, +inf)} (see
          // compileFastProgram), folded here in registers and applied as ONE leaf, and the subtree's
          // words (h.w of them) are jumped over.  If any ray of the warp is not in that case the warp
          // runs the subtree's words instead.  Two planes alone need no such guard: two single spans
          // that overlap always fuse into one.
            float loEnd[R], upStart[R];
            uint32_t loTags[R], upTags[R];
            bool full[R];
#pragma unroll
            for(int r = 0; r < R; r++)
                loEnd[r] = -3.0e38f, upStart[r] = 3.0e38f, loTags[r] = upTags[r] = 0u, full[r] = false;
            uint32_t left = h.z, tags = h.y;
            PTB_ADVANCE(1);
            for(;;)
            {
                const uint4 u0 = PTB_PROG(0);
#pragma unroll
                for(int r = 0; r < R; r++)
                {
                    planeFold<O0>(u0, tags, o[r], d[r], loEnd[r], upStart[r], loTags[r], upTags[r], full[r]);
                }
                if(left == 0u)
                {
                    PTB_ADVANCE(1);
                    break;
                }
                left--;
                tags = PTB_PROG(1).x; // next plane: tags word, then geometry
                PTB_ADVANCE(2);
            }
            if(h.z)
            {
                bool separate = true;
#pragma unroll
                for(int r = 0; r < R; r++)
                    separate = separate && !full[r] && (h.z == 1u || loEnd[r] < upStart[r]);
                if(!__all_sync(0xFFFFFFFFu, separate))
                    continue; // the subtree's own words are next
                PTB_ADVANCE(h.w);
            }
#pragma unroll
            for(int r = 0; r < R; r++)
                planeFinish(loEnd[r], upStart[r], loTags[r], upTags[r], full[r], hit[r], two[r], lt0[r], lt1[r], ut0[r], ltags[r], utags[r]);
            PTB_MISS_EXIT();
        }
        else if(h.x & FK_SPHERE)
        {
            const uint4 u0 = PTB_PROG(1);
            PTB_ADVANCE(2);
#pragma unroll
            for(int r = 0; r < R; r++)
            {
                float a, b, e;
                sphereDisc<O0>(h, u0, o[r], d[r], a, b, e);
                hit[r] = e > PTB_EPS;
                if(hit[r])
                { // src/sphere.cpp:44-46
                    float q = sqrtApprox(e), ia = rcpApprox(a);
                    lt0[r] = (-b - q) * ia, lt1[r] = (-b + q) * ia;
                }
            }
            PTB_MISS_EXIT();
        }
        else
        { // control word
            const uint32_t sub = (h.x >> 4) & 3u;
            if(sub == 0u)
                break;
            PTB_ADVANCE(1);
            if(sub == 2u)
            {
                uint32_t live = 0u;
#pragma unroll
                for(int r = 0; r < R; r++)
                    live |= st[r].bits;
                if(!__any_sync(0xFFFFFFFFu, live & 1u))
                    PTB_ADVANCE(h.y);
                continue;
            }
        }
        // ---- the ONE place where lists are pushed and operators run (the code of the three merge
        // state machines exists once in the kernel: the hot loop stays instruction-cache resident) ----
#pragma unroll
        for(int r = 0; r < R; r++)
            fastPushOperate(h.x, hit[r], two[r], lt0[r], lt1[r], ut0[r], ltags[r], utags[r], st[r], c.poolOf(r));
    }
#undef PTB_PROG
#undef PTB_ADVANCE
#undef PTB_MISS_EXIT
#pragma unroll
    for(int r = 0; r < R; r++)
        topOut[r] = st[r].top;
}

} // namespace ptb
