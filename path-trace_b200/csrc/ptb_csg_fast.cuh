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

extern __shared__ uint4 ptbSmem[];

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
#ifndef PTB_TAIL_CLOSED
#define PTB_TAIL_CLOSED 1
#endif
#define PTB_FAST_MAX_STACK 32

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
    FastPool pool;       // ray 0's pool
    __device__ __forceinline__ FastPool poolOf(int r) const
    {
        FastPool p = pool;
        p.base += (uint32_t)r * poolStride;
        p.ovf += (size_t)r * ovfRay;
        return p;
    }
};

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
    if(b0 - a0 == 1 && top - b0 == 1)
        return mergeListsT(pool, kind, a0, b0, top, a0); // both inputs are in registers before the first output
    int end = mergeListsT(pool, kind, a0, b0, top, top);
    int n = end - top;
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

// A leaf that produced one span (or a plane cluster that produced two, in order): push the list, or
// combine it with the list below per `action`.  Returns {bits, top, sp}.
PTB_SLOW_FN uint3 fastLeafHit2(uint32_t poolBase, int cap, float *ovf, size_t ovfStride, uint32_t action, bool two, const SpanRec &s, const SpanRec &s2, uint32_t bits,
                               int top, uint32_t sp)
{
    const uint32_t W = 4u * PTB_FAST_THREADS;
    FastPool pool;
    pool.base = poolBase, pool.cap = cap, pool.ovf = ovf, pool.ovfStride = ovfStride;
    const int n = two ? 2 : 1;
    const bool aHas = bits & 1u;
    if(action == FA_PUSH || (action == FA_UNION && !aHas))
    {
        sts32(sp, (uint32_t)top);
        pool.store(top, s);
        if(two)
            pool.store(top + 1, s2);
        return make_uint3(action == FA_PUSH ? (bits << 1) | 1u : bits | 1u, (uint32_t)(top + n), sp + W);
    }
    if(!aHas)
        return make_uint3(bits, (uint32_t)top, sp); // {} \ B and {} n B stay empty; B is never stored
    int a0 = (int)lds32(sp - W);
    pool.store(top, s);
    if(two)
        pool.store(top + 1, s2);
    int ntop = fastMerge(poolBase, cap, ovf, ovfStride, action, a0, top, top + n);
    if(ntop == a0)
        return make_uint3(bits & ~1u, (uint32_t)ntop, sp - W);
    return make_uint3(bits, (uint32_t)ntop, sp);
}
PTB_SLOW_FN uint3 fastLeafHit(uint32_t poolBase, int cap, float *ovf, size_t ovfStride, uint32_t action, float t0, float t1, uint32_t tags, uint32_t bits, int top,
                              uint32_t sp)
{
    SpanRec s;
    s.t0 = t0, s.t1 = t1, s.tags = tags;
    return fastLeafHit2(poolBase, cap, ovf, ovfStride, action, false, s, s, bits, top, sp);
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

// Leaf epilogue for one ray: the common case (miss) is one shift.
__device__ __forceinline__ void fastLeafDone(const FastPool &pool, uint32_t hx, uint32_t tags, bool hit, float a, float b, float disc, FastState &st)
{
    if(hit)
    { // src/sphere.cpp:44-46
        float s = sqrtApprox(disc), ia = rcpApprox(a);
        uint3 q = fastLeafHit(pool.base, pool.cap, pool.ovf, pool.ovfStride, (hx >> 4) & 3u, (-b - s) * ia, (-b + s) * ia, tags, st.bits, st.top, st.sp);
        st.bits = q.x, st.top = (int)q.y, st.sp = q.z;
    }
    else
        st.bits <<= (hx >> 6) & 1u;
}

// Runs the whole program for R rays per lane, all 32 lanes of the warp together (SKIP words
// vote).  Ray r's root span list ends up in pool r, slots [0, topOut[r]).  (R = 2 was measured
// slower than R = 1 on B200 — occupancy beats per-thread ILP here — and is kept for experiments.)
template <bool GP, bool O0, int R>
__device__ __forceinline__ void evalFast(const FastCtx &c, const uint4 *__restrict__ gprog, const V3 (&o)[R], const V3 (&d)[R], int (&topOut)[R])
{
    const uint32_t W = 4u * PTB_FAST_THREADS;
    FastState st[R];
#pragma unroll
    for(int r = 0; r < R; r++)
        st[r].bits = 0u, st[r].top = 0, st[r].sp = c.stk + (uint32_t)r * c.stkStride;
    uint32_t pc = GP ? 0u : c.prog; // GP: word index; else shared byte address
#define PTB_PROG(i) (GP ? __ldg(gprog + pc + (i)) : lds128(pc + 16u * (i)))
#define PTB_ADVANCE(n) (pc += GP ? (uint32_t)(n) : 16u * (uint32_t)(n))
    for(;;)
    {
        uint4 h = PTB_PROG(0);
        bool hit[R];
        if(h.x & FK_SPHERE_XF)
        {
            if(h.x & FH_PAIR)
            { // two tail-less... the first has no tail; both discriminants are computed together (ILP)
                const uint4 u0 = PTB_PROG(1), u1 = PTB_PROG(2), u2 = PTB_PROG(3);
                const uint4 g = PTB_PROG(4), v0 = PTB_PROG(5), v1 = PTB_PROG(6), v2 = PTB_PROG(7);
                PTB_ADVANCE(8);
                float a0[R], b0[R], e0[R], a1[R], b1[R], e1[R];
#pragma unroll
                for(int r = 0; r < R; r++)
                {
                    xfDisc<O0>(h, u0, u1, u2, o[r], d[r], a0[r], b0[r], e0[r]);
                    xfDisc<O0>(g, v0, v1, v2, o[r], d[r], a1[r], b1[r], e1[r]);
                }
#pragma unroll
                for(int r = 0; r < R; r++)
                {
                    fastLeafDone(c.poolOf(r), h.x, h.y, e0[r] > PTB_EPS, a0[r], b0[r], e0[r], st[r]);
                    hit[r] = e1[r] > PTB_EPS;
                    fastLeafDone(c.poolOf(r), g.x, g.y, hit[r], a1[r], b1[r], e1[r], st[r]);
                }
                h = g; // the second leaf's tail, if any, runs below
            }
            else
            {
                const uint4 u0 = PTB_PROG(1), u1 = PTB_PROG(2), u2 = PTB_PROG(3);
                PTB_ADVANCE(4);
#pragma unroll
                for(int r = 0; r < R; r++)
                {
                    float a, b, e;
                    xfDisc<O0>(h, u0, u1, u2, o[r], d[r], a, b, e);
                    hit[r] = e > PTB_EPS;
                    fastLeafDone(c.poolOf(r), h.x, h.y, hit[r], a, b, e, st[r]);
                }
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
                const float nx = __uint_as_float(u0.x), ny = __uint_as_float(u0.y), nz = __uint_as_float(u0.z);
#pragma unroll
                for(int r = 0; r < R; r++)
                {
                    float divisor = fmaf(d[r].x, nx, fmaf(d[r].y, ny, d[r].z * nz));
                    float numerator = O0 ? -__uint_as_float(u0.w) : -fmaf(o[r].x, nx, fmaf(o[r].y, ny, fmaf(o[r].z, nz, __uint_as_float(u0.w))));
                    float t = numerator * rcpApprox(divisor);
                    bool parallel = fabsf(divisor) < PTB_EPS * PTB_EPS || !(fabsf(t) < PTB_MAX_VALUE);
#if PTB_PLANE_SELECT
                    // selects, not branches: the cases differ from lane to lane
                    const bool inPlane = parallel && fabsf(numerator) < PTB_EPS * PTB_EPS; // the ray lies in the plane: the whole line
                    const bool upper = !parallel && divisor < 0.0f, lower = !parallel && !(divisor < 0.0f);
                    // [t, +inf): the earliest start wins and brings both its tags (src/union.cpp:113-123)
                    const bool newStart = (upper && t <= upStart[r]) || inPlane;
                    // (-inf, t]: equal starts keep the later operand's start, the larger end its tag (:124-133)
                    const bool newEnd = (lower && t >= loEnd[r]) || inPlane;
                    full[r] = full[r] || inPlane;
                    upStart[r] = newStart && !inPlane ? t : upStart[r];
                    upTags[r] = newStart ? tags : upTags[r];
                    loTags[r] = lower || inPlane ? (loTags[r] & 0xFFFF0000u) | (tags & 0xFFFFu) : loTags[r];
                    loEnd[r] = newEnd && !inPlane ? t : loEnd[r];
                    loTags[r] = newEnd ? (loTags[r] & 0xFFFFu) | (tags & 0xFFFF0000u) : loTags[r];
#else
                    if(parallel)
                    {
                        if(fabsf(numerator) < PTB_EPS * PTB_EPS)
                            full[r] = true, loTags[r] = upTags[r] = tags; // the ray lies in the plane: the whole line
                    }
                    else if(divisor < 0.0f)
                    { // [t, +inf): the earliest start wins and brings both its tags (src/union.cpp:113-123)
                        if(t <= upStart[r])
                            upStart[r] = t, upTags[r] = tags;
                    }
                    else
                    { // (-inf, t]: equal starts keep the later operand's start, the larger end its tag (:124-133)
                        loTags[r] = (loTags[r] & 0xFFFF0000u) | (tags & 0xFFFFu);
                        if(t >= loEnd[r])
                            loEnd[r] = t, loTags[r] = (loTags[r] & 0xFFFFu) | (tags & 0xFFFF0000u);
                    }
#endif
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
            {
                const bool hasLo = loEnd[r] > -3.0e38f, hasUp = upStart[r] < 3.0e38f;
                SpanRec s, s2;
                s2.t0 = upStart[r], s2.t1 = PTB_MAX_VALUE, s2.tags = upTags[r];
                bool two = false;
                if(full[r])
                    s.t0 = -PTB_MAX_VALUE, s.t1 = PTB_MAX_VALUE, s.tags = upTags[r];
                else if(loEnd[r] >= upStart[r]) // two planes only: either order gives the whole line (src/union.cpp:113-133)
                    s.t0 = -PTB_MAX_VALUE, s.t1 = PTB_MAX_VALUE, s.tags = (loTags[r] & 0xFFFFu) | (upTags[r] & 0xFFFF0000u);
                else if(hasLo)
                    s.t0 = -PTB_MAX_VALUE, s.t1 = loEnd[r], s.tags = loTags[r], two = hasUp;
                else
                    s = s2;
                hit[r] = full[r] || hasLo || hasUp;
                if(hit[r])
                {
                    FastPool pool = c.poolOf(r);
                    uint3 q = fastLeafHit2(pool.base, pool.cap, pool.ovf, pool.ovfStride, (h.x >> 4) & 3u, two, s, s2, st[r].bits, st[r].top, st[r].sp);
                    st[r].bits = q.x, st[r].top = (int)q.y, st[r].sp = q.z;
                }
                else
                    st[r].bits <<= (h.x >> 6) & 1u;
            }
        }
        else if(h.x & FK_SPHERE)
        {
            const uint4 u0 = PTB_PROG(1);
            PTB_ADVANCE(2);
#pragma unroll
            for(int r = 0; r < R; r++)
            {
                float a = fmaf(d[r].x, d[r].x, fmaf(d[r].y, d[r].y, d[r].z * d[r].z));
                float b, cc;
                if(O0)
                {
                    b = fmaf(__uint_as_float(u0.x), d[r].x, fmaf(__uint_as_float(u0.y), d[r].y, __uint_as_float(u0.z) * d[r].z));
                    cc = __uint_as_float(h.w);
                }
                else
                {
                    float ox = o[r].x + __uint_as_float(u0.x), oy = o[r].y + __uint_as_float(u0.y), oz = o[r].z + __uint_as_float(u0.z);
                    b = fmaf(ox, d[r].x, fmaf(oy, d[r].y, oz * d[r].z));
                    cc = fmaf(ox, ox, fmaf(oy, oy, fmaf(oz, oz, __uint_as_float(h.z))));
                }
                float e = fmaf(b, b, -(a * cc));
                hit[r] = e > PTB_EPS;
                fastLeafDone(c.poolOf(r), h.x, h.y, hit[r], a, b, e, st[r]);
            }
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
#pragma unroll
            for(int r = 0; r < R; r++)
                hit[r] = true; // an OPS word has no leaf: nothing to drop below
        }
        if(!(h.x & FH_TAIL))
            continue;
        uint32_t npost = (h.x >> 8) & 15u;
        if(h.x & FH_UNION_TAIL)
        { // a run of Unions: the lists involved collapse into one; spans move only if two of them have any
#pragma unroll
            for(int r = 0; r < R; r++)
            {
                const uint32_t involved = st[r].bits & ((2u << npost) - 1u);
                if(PTB_TAIL_CLOSED && __popc(involved) <= 1)
                {
                    st[r].bits = ((st[r].bits >> npost) & ~1u) | (involved ? 1u : 0u);
                    continue;
                }
                for(uint32_t k = npost; k; k--)
                {
                    if((st[r].bits & 3u) == 3u)
                    {
                        FastPool pool = c.poolOf(r);
                        uint3 q = fastOpSlow(pool.base, pool.cap, pool.ovf, pool.ovfStride, FA_UNION, 3u, st[r].top, st[r].sp);
                        st[r].top = (int)q.y, st[r].sp = q.z;
                    }
                    st[r].bits = (st[r].bits >> 1) | (st[r].bits & 1u);
                }
            }
            continue;
        }
        if(((h.x >> 4) & 3u) == FA_INTERSECTION && !(h.x & FK_CONTROL))
        { // A n {} = {}: a missed leaf drops the spans of the list below
#pragma unroll
            for(int r = 0; r < R; r++)
                if(!hit[r] && (st[r].bits & 1u))
                {
                    st[r].sp -= W;
                    st[r].top = (int)lds32(st[r].sp);
                    st[r].bits &= ~1u;
                }
        }
        uint32_t post = h.x >> 12;
        while(npost)
        {
            const uint32_t op = (post & 3u) + FA_UNION;
#pragma unroll
            for(int r = 0; r < R; r++)
                fastOp(c.poolOf(r), op, st[r]);
            post >>= 2, npost--;
        }
    }
#undef PTB_PROG
#undef PTB_ADVANCE
#pragma unroll
    for(int r = 0; r < R; r++)
        topOut[r] = st[r].top;
}

} // namespace ptb
