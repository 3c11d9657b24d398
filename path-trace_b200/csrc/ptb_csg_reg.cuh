// ptb_csg_reg.cuh — register-resident span lists for the scene-SPECIALISED evaluator (second generation).
//
// The interpreter (ptb_csg_fast.cuh) and the first generation of specialised kernels keep every span list in a
// shared-memory pool and drive it with a bit stack, a start stack and one out-of-line push-and-operate function.
// ncu on the specialised C5 camera-ray kernel (profiles/r01n_spec_c5.txt): 853 instructions per ray, 14 % of them
// FP32 — ~200 go into the 2.2 out-of-line calls per ray, ~150 into generic merge loops over the pool, ~50 into the
// first-hit scan; on the incoherent C3 batch 3 900 per ray (profiles/r02b_full_c3.txt).
//
// For a scene whose tree is known when the code is GENERATED none of that bookkeeping is needed:
//   * every node of the CSG tree has its own list variable: no operand stack, no bit stack, no start stack;
//   * a list of ONE span lives in registers (`RList::a`) — what a leaf yields when it is hit, and most merges of
//     two such lists; only lists of two or more spans go to the shared-memory pool, which is then a plain stack of
//     the pool-resident lists still pending (`REval::top`); the plane cluster that is the root Union's right
//     operand (the six sky planes of every scene built like the reference's, src/test.cpp:134-140) stays in
//     registers even with its two spans;
//   * an empty operand costs a compare (A u {} = A, {} \ B = {}, ...: what the reference's state machines emit
//     when aEnded / bEnded is set from the start, src/union.cpp:88-102, src/difference.cpp:88-99,
//     src/intersection.cpp:88-90), two single spans merge in closed form in registers (the three next()
//     machines unrolled for |A| = |B| = 1), anything longer goes through ONE out-of-line copy of the general
//     streaming merges (mergeRangesT, shared with the interpreter);
//   * the trace kernels only need the first boundary with t >= eps (include/path-trace.h:66-101): when the root is
//     a Union its stream is scanned while it is merged and never stored (rlUnionFirstHit).
// Lanes are independent (no warp votes): a miss is n = 0 and branches skip work per lane.
//
// Span values, tags and list lengths are exactly the interpreter's (same leaf arithmetic, same state machines),
// which the tests check bit for bit (tests/test_specialize.py on the GPU, tests/test_spec_emulation.py on the CPU).
// A ray whose lists outgrow the shared pool sets REval::ovf; the kernel then reruns the warp through the interpreter.
#pragma once
#include "ptb_csg_fast.cuh"

namespace ptb
{

struct RList
{
    int n;     // spans in the list
    SpanRec a; // the span when n == 1 (register-resident)
    int base;  // n >= 2: pool slot of span 0 (spans are pool slots [base, base + n))
};

// Per-ray evaluator state: the pool is a stack of the pool-resident lists that are still pending.
struct REval
{
    uint32_t poolBase; // shared address of this thread's slot 0 (FastPool::base)
    int cap;           // slots in shared memory
    int top;           // slots in use
    bool ovf;          // a list did not fit: the result of this ray is void, rerun it through the interpreter
    uint32_t tab;      // shared address of the module's constant table (shape functions read their leaves' constants there)
};

// One word of the constant table.  Not volatile: the table is read-only once staged, so the compiler may merge and
// move these loads.  (In constant memory the same reads were 13 indexed scalar LDCs per transformed sphere and the
// kernel stalled on the memory-input/output queue: profiles/r02f_full_c4_shapes.txt.)
#ifdef PTB_HOST_EMU
static inline uint4 ldsTab128(uint32_t a) { return lds128(a); }
#else
__device__ __forceinline__ uint4 ldsTab128(uint32_t a)
{
    uint4 v;
    asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
#endif
// A line through the origin along d against a bounding-sphere word {C, sqrt(|C|^2 - R^2)} (compileFastProgram,
// ptb_capi.cu): it meets the sphere iff |C.d| >= |d| sqrt(|C|^2 - R^2).  sd = |d|.
__device__ __forceinline__ bool rlBoundMet(const uint4 w, const V3 d, float sd)
{
    const float b = fmaf(__uint_as_float(w.x), d.x, fmaf(__uint_as_float(w.y), d.y, __uint_as_float(w.z) * d.z));
    return fabsf(b) >= sd * __uint_as_float(w.w);
}
// Defined by the generated module: copies its table (if it has one) to shared memory at `tab`; all threads call it.
__device__ __forceinline__ void specStageTable(uint32_t tab);

__device__ __forceinline__ void rpoolStore(uint32_t poolBase, int slot, const SpanRec &r)
{
    const uint32_t a = poolBase + (uint32_t)slot * (12u * PTB_FAST_THREADS);
    sts32(a, __float_as_uint(r.t0));
    sts32(a + 4u * PTB_FAST_THREADS, __float_as_uint(r.t1));
    sts32(a + 8u * PTB_FAST_THREADS, r.tags);
}
__device__ __forceinline__ SpanRec rpoolLoad(uint32_t poolBase, int slot)
{
    const uint32_t a = poolBase + (uint32_t)slot * (12u * PTB_FAST_THREADS);
    SpanRec r;
    r.t0 = __uint_as_float(lds32(a));
    r.t1 = __uint_as_float(lds32(a + 4u * PTB_FAST_THREADS));
    r.tags = lds32(a + 8u * PTB_FAST_THREADS);
    return r;
}
// The pool as the merge templates of ptb_csg.cuh see it (no overflow area: REval::ovf guards the capacity).
struct RPool
{
    uint32_t base;
    __device__ __forceinline__ SpanRec load(int slot) const { return rpoolLoad(base, slot); }
    __device__ __forceinline__ void store(int slot, const SpanRec &r) const { rpoolStore(base, slot, r); }
};

__device__ __forceinline__ SpanRec rlSpan(float t0, float t1, uint32_t tags)
{
    SpanRec s;
    s.t0 = t0, s.t1 = t1, s.tags = tags;
    return s;
}
__device__ __forceinline__ RList rlEmpty()
{
    RList r;
    r.n = 0, r.base = 0;
    r.a.t0 = r.a.t1 = 0.f;
    r.a.tags = 0u;
    return r;
}
// A leaf: nothing or one span.
__device__ __forceinline__ RList rlLeaf(bool hit, float t0, float t1, uint32_t tags)
{
    RList r;
    r.n = hit ? 1 : 0, r.base = 0;
    r.a.t0 = t0, r.a.t1 = t1, r.a.tags = tags;
    return r;
}
// Two spans (in order) as a new pool-resident list on top of the stack.
__device__ __forceinline__ RList rlPush2(REval &ev, const SpanRec &first, const SpanRec &second)
{
    RList r = rlEmpty();
    if(ev.top + 2 > ev.cap)
    {
        ev.ovf = true;
        return r;
    }
    rpoolStore(ev.poolBase, ev.top, first);
    rpoolStore(ev.poolBase, ev.top + 1, second);
    r.n = 2, r.base = ev.top;
    ev.top += 2;
    return r;
}
// A plane cluster (planeFinish): nothing, one span, or two in order, the second reaching to +max_value.
__device__ __forceinline__ RList rlCluster(REval &ev, bool hit, bool two, float lt0, float lt1, float ut0, uint32_t ltags, uint32_t utags)
{
    RList r = rlLeaf(hit, lt0, lt1, ltags);
    if(hit && two)
    {
        SpanRec up;
        up.t0 = ut0, up.t1 = PTB_MAX_VALUE, up.tags = utags;
        r = rlPush2(ev, r.a, up);
    }
    return r;
}

// ---- |A| = |B| = 1: the three next() state machines run to exhaustion by hand ---------------------------------
// UnionSpanIterator::next, src/union.cpp:84-134
__device__ __forceinline__ RList rlUnion11(REval &ev, SpanRec A, SpanRec B)
{
    if(A.t1 < B.t0)
        return rlPush2(ev, A, B); // disjoint, already in order
    if(B.t1 < A.t0)
        return rlPush2(ev, B, A);
    RList r = rlEmpty();
    r.n = 1;
    if(A.t0 < B.t0)
    { // A starts first and absorbs B (:113-123)
        if(A.t1 < B.t1)
            copyEndFromEnd(A, B);
        r.a = A;
    }
    else
    { // B absorbs A (:124-133)
        if(A.t1 > B.t1)
            copyEndFromEnd(B, A);
        r.a = B;
    }
    return r;
}
// DifferenceSpanIterator::next, src/difference.cpp:84-135 (quirk Q4 at :124-131: the inverted span)
__device__ __forceinline__ RList rlDifference11(REval &ev, SpanRec A, SpanRec B)
{
    RList r = rlEmpty();
    if(A.t1 < B.t0 || B.t1 < A.t0)
        r.n = 1, r.a = A; // B lies wholly before or behind A
    else if(A.t0 < B.t0)
    {
        if(A.t1 < B.t1)
        { // B cuts A's end off
            copyEndFromStart(A, B);
            r.n = 1, r.a = A;
        }
        else
        { // B lies inside A: two pieces
            SpanRec head = A;
            copyEndFromStart(head, B);
            copyStartFromEnd(A, B);
            r = rlPush2(ev, head, A);
        }
    }
    else if(A.t1 > B.t1)
    { // A starts inside B and outlives it: the reference sets A.end = B.start (Q4)
        copyEndFromStart(A, B);
        r.n = 1, r.a = A;
    }
    return r; // else: A inside B, nothing left
}
// IntersectionSpanIterator::next, src/intersection.cpp:84-130
__device__ __forceinline__ RList rlIntersection11(SpanRec A, SpanRec B)
{
    RList r = rlEmpty();
    if(A.t1 < B.t0 || B.t1 < A.t0)
        return r;
    r.n = 1;
    if(A.t0 < B.t0)
    {
        if(A.t1 < B.t1)
        {
            copyStartFromStart(A, B);
            r.a = A;
        }
        else
            r.a = B;
    }
    else if(B.t1 < A.t1)
    {
        copyStartFromStart(B, A);
        r.a = B;
    }
    else
        r.a = A;
    return r;
}

// ---- anything longer: ONE out-of-line copy of the general streaming merges ---------------------------------------
// counts = kind | |A| << 8 | |B| << 16.  A single-span operand arrives in registers and is pushed onto the pool
// first; the output is appended above everything (it never overwrites an operand) and then moved down to `low`, the
// first slot the operands occupied.  Returns the output's length; the caller has checked the capacity.
static __device__ __noinline__ int rlMergeGeneral(uint32_t poolBase, uint32_t counts, float a0t0, float a0t1, uint32_t a0tags, int aBase, float b0t0, float b0t1, uint32_t b0tags,
                                                  int bBase, int top)
{
    RPool pool;
    pool.base = poolBase;
    const int kind = (int)(counts & 0xFFu), nA = (int)((counts >> 8) & 0xFFu), nB = (int)(counts >> 16);
    const int low = nA >= 2 ? aBase : (nB >= 2 ? bBase : top);
    if(nA == 1)
    {
        SpanRec s;
        s.t0 = a0t0, s.t1 = a0t1, s.tags = a0tags;
        aBase = top;
        pool.store(top++, s);
    }
    if(nB == 1)
    {
        SpanRec s;
        s.t0 = b0t0, s.t1 = b0t1, s.tags = b0tags;
        bBase = top;
        pool.store(top++, s);
    }
    const int n = mergeRangesT(pool, kind, aBase, aBase + nA, bBase, bBase + nB, top) - top;
    if(low != top)
        for(int i = 0; i < n; i++)
            pool.store(low + i, pool.load(top + i));
    return n;
}
// Stack discipline: the operands are the topmost pending lists, so everything from `low` up is theirs and is
// released; a result of one span comes back into registers.
__device__ __forceinline__ RList rlGeneral(REval &ev, int kind, const RList &A, const RList &B)
{
    const int nA = A.n, nB = B.n;
    const int low = nA >= 2 ? A.base : (nB >= 2 ? B.base : ev.top);
    RList r = rlEmpty();
    if(ev.top + (nA == 1 ? 1 : 0) + (nB == 1 ? 1 : 0) + nA + nB > ev.cap)
    { // (a merge emits at most |A| + |B| spans)
        ev.ovf = true;
        return r;
    }
    const int n = rlMergeGeneral(ev.poolBase, (uint32_t)kind | ((uint32_t)nA << 8) | ((uint32_t)nB << 16), A.a.t0, A.a.t1, A.a.tags, A.base, B.a.t0, B.a.t1, B.a.tags, B.base, ev.top);
    r.n = n;
    if(n >= 2)
        r.base = low, ev.top = low + n;
    else
    {
        ev.top = low;
        if(n == 1)
            r.a = rpoolLoad(ev.poolBase, low);
    }
    return r;
}
// Releases the pool slots of operands whose spans an operator throws away.
__device__ __forceinline__ void rlRelease(REval &ev, const RList &A, const RList &B)
{
    if(A.n >= 2)
        ev.top = A.base;
    else if(B.n >= 2)
        ev.top = B.base;
}

// ---- the three operators ----------------------------------------------------------------------------------------------
__device__ __forceinline__ RList rlUnion(REval &ev, const RList &A, const RList &B)
{
    if(B.n == 0)
        return A;
    if(A.n == 0)
        return B;
    if(A.n == 1 && B.n == 1)
        return rlUnion11(ev, A.a, B.a);
    return rlGeneral(ev, DK_UNION, A, B);
}
__device__ __forceinline__ RList rlDifference(REval &ev, const RList &A, const RList &B)
{
    if(B.n == 0)
        return A;
    if(A.n == 0)
    {
        rlRelease(ev, A, B);
        return rlEmpty();
    }
    if(A.n == 1 && B.n == 1)
        return rlDifference11(ev, A.a, B.a);
    return rlGeneral(ev, DK_DIFFERENCE, A, B);
}
__device__ __forceinline__ RList rlIntersection(REval &ev, const RList &A, const RList &B)
{
    if(A.n == 0 || B.n == 0)
    {
        rlRelease(ev, A, B);
        return rlEmpty();
    }
    if(A.n == 1 && B.n == 1)
        return rlIntersection11(A.a, B.a);
    return rlGeneral(ev, DK_INTERSECTION, A, B);
}

// The same three, accumulating into the left operand (what the generated code calls: a node's list variable is
// its left operand's, so A u {} = A costs nothing at all).
__device__ __forceinline__ void rlUnionInto(REval &ev, RList &A, const RList &B)
{
    if(B.n == 0)
        return;
    if(A.n == 0)
        A = B;
    else if(A.n == 1 && B.n == 1)
        A = rlUnion11(ev, A.a, B.a);
    else
        A = rlGeneral(ev, DK_UNION, A, B);
}
__device__ __forceinline__ void rlDifferenceInto(REval &ev, RList &A, const RList &B)
{
    if(B.n == 0 || A.n == 0)
        return; // (the right operand of an empty left one is never evaluated: nothing to release)
    if(A.n == 1 && B.n == 1)
        A = rlDifference11(ev, A.a, B.a);
    else
        A = rlGeneral(ev, DK_DIFFERENCE, A, B);
}
__device__ __forceinline__ void rlIntersectionInto(REval &ev, RList &A, const RList &B)
{
    if(A.n == 0)
        return;
    if(B.n == 0)
    {
        rlRelease(ev, A, B);
        A.n = 0;
    }
    else if(A.n == 1 && B.n == 1)
        A = rlIntersection11(A.a, B.a);
    else
        A = rlGeneral(ev, DK_INTERSECTION, A, B);
}

// The same three with everything but the empty-operand tests in ONE out-of-line function, for large programs: inlined
// at every operator the closed forms and the call of the general merge are ~120 instructions each, most of them never
// executed by a warp of camera rays, and the straight-line code of a 70-leaf scene stops fitting the instruction
// cache (ncu, config 4: 5.4 stall cycles per issue waiting for instructions, profiles/r02w_spec_c4.txt).
struct RMerged
{
    RList L;
    int top;
    bool ovf;
};
static __device__ __noinline__ RMerged rlMergeSlow(int kind, uint32_t poolBase, int cap, int top, RList A, RList B)
{
    REval ev;
    ev.poolBase = poolBase, ev.cap = cap, ev.top = top, ev.ovf = false, ev.tab = 0u;
    RMerged r;
    if(A.n == 1 && B.n == 1)
        r.L = kind == DK_UNION ? rlUnion11(ev, A.a, B.a) : kind == DK_DIFFERENCE ? rlDifference11(ev, A.a, B.a) : rlIntersection11(A.a, B.a);
    else
        r.L = rlGeneral(ev, kind, A, B);
    r.top = ev.top, r.ovf = ev.ovf;
    return r;
}
__device__ __forceinline__ void rlMergeCold(REval &ev, int kind, RList &A, const RList &B)
{
    const RMerged r = rlMergeSlow(kind, ev.poolBase, ev.cap, ev.top, A, B);
    A = r.L, ev.top = r.top, ev.ovf = ev.ovf || r.ovf;
}
// A leaner form of the same: only the closed forms of two single spans go out of line, as ONE function with scalar
// parameters and a uint4 result (registers both ways): {n | overflow << 31, t0, t1, tags}; a result of two spans
// lies in pool slots [top, top + 2).  Lists of two or more spans take rlGeneral (whose merge is out of line already).
static __device__ __noinline__ uint4 rlMerge11Call(int kind, uint32_t poolBase, int cap, int top, float a0, float a1, uint32_t at, float b0, float b1, uint32_t bt)
{
    REval ev;
    ev.poolBase = poolBase, ev.cap = cap, ev.top = top, ev.ovf = false, ev.tab = 0u;
    const SpanRec A = rlSpan(a0, a1, at), B = rlSpan(b0, b1, bt);
    const RList r = kind == DK_UNION ? rlUnion11(ev, A, B) : rlDifference11(ev, A, B);
    return make_uint4((uint32_t)r.n | (ev.ovf ? 0x80000000u : 0u), __float_as_uint(r.a.t0), __float_as_uint(r.a.t1), r.a.tags);
}
__device__ __forceinline__ void rlMerge11(REval &ev, int kind, RList &A, const RList &B)
{
    const uint4 r = rlMerge11Call(kind, ev.poolBase, ev.cap, ev.top, A.a.t0, A.a.t1, A.a.tags, B.a.t0, B.a.t1, B.a.tags);
    A.n = (int)(r.x & 0x7FFFFFFFu), A.a = rlSpan(__uint_as_float(r.y), __uint_as_float(r.z), r.w), A.base = ev.top;
    ev.ovf = ev.ovf || (r.x >> 31) != 0u;
    if(A.n >= 2)
        ev.top += 2;
}
__device__ __forceinline__ void rlUnionIntoLean(REval &ev, RList &A, const RList &B)
{
    if(B.n == 0)
        return;
    if(A.n == 0)
        A = B;
    else if(A.n == 1 && B.n == 1)
        rlMerge11(ev, DK_UNION, A, B);
    else
        A = rlGeneral(ev, DK_UNION, A, B);
}
__device__ __forceinline__ void rlDifferenceIntoLean(REval &ev, RList &A, const RList &B)
{
    if(B.n == 0 || A.n == 0)
        return;
    if(A.n == 1 && B.n == 1)
        rlMerge11(ev, DK_DIFFERENCE, A, B);
    else
        A = rlGeneral(ev, DK_DIFFERENCE, A, B);
}
__device__ __forceinline__ void rlIntersectionIntoLean(REval &ev, RList &A, const RList &B) { rlIntersectionInto(ev, A, B); } // (its closed form pushes nothing and is short)

// (PTB_UNLIKELY: the merge proper goes to the end of the function, out of the way of the code every ray runs)
#ifdef PTB_HOST_EMU
#define PTB_UNLIKELY(x) (x)
#define PTB_LIKELY(x) (x)
#else
#define PTB_UNLIKELY(x) __builtin_expect(!!(x), 0)
#define PTB_LIKELY(x) __builtin_expect(!!(x), 1)
#endif
__device__ __forceinline__ void rlUnionIntoCold(REval &ev, RList &A, const RList &B)
{
    if(PTB_UNLIKELY(B.n != 0))
    {
        if(A.n == 0)
            A = B;
        else
            rlMergeCold(ev, DK_UNION, A, B);
    }
}
__device__ __forceinline__ void rlDifferenceIntoCold(REval &ev, RList &A, const RList &B)
{
    if(PTB_UNLIKELY(B.n != 0 && A.n != 0))
        rlMergeCold(ev, DK_DIFFERENCE, A, B);
}
__device__ __forceinline__ void rlIntersectionIntoCold(REval &ev, RList &A, const RList &B)
{
    if(PTB_UNLIKELY(A.n != 0))
    {
        if(B.n == 0)
        {
            rlRelease(ev, A, B);
            A.n = 0;
        }
        else
            rlMergeCold(ev, DK_INTERSECTION, A, B);
    }
}

// ---- the root -----------------------------------------------------------------------------------------------------------
__device__ __forceinline__ SpanRec rlGet(const REval &ev, const RList &L, int i)
{
    if(L.n == 1)
        return L.a;
    return rpoolLoad(ev.poolBase, L.base + i);
}
// One span of the root stream against the scan of traceRay (include/path-trace.h:66-96): -1 = go on to the next
// span, else PTB_HIT_* with t and tag set.  (Same tests, same order as firstHit(), ptb_csg.cuh.)
__device__ __forceinline__ int rlScanSpan(const SpanRec &s, float &t, uint32_t &tag)
{
    if(s.t0 >= PTB_MAX_VALUE)
        return PTB_HIT_MISS;
    if(s.t0 >= PTB_EPS)
    {
        t = s.t0, tag = startTag(s);
        return PTB_HIT_ENTER;
    }
    if(s.t1 >= PTB_MAX_VALUE)
        return PTB_HIT_MISS;
    if(s.t1 >= PTB_EPS)
    {
        t = s.t1, tag = endTag(s);
        return PTB_HIT_EXIT;
    }
    return -1;
}
// First boundary with t >= eps of a finished list.
__device__ __forceinline__ int rlFirstHit(const REval &ev, const RList &R, float &t, uint32_t &tag)
{
    for(int i = 0; i < R.n; i++)
    {
        const int k = rlScanSpan(rlGet(ev, R, i), t, tag);
        if(k >= 0)
            return k;
    }
    return PTB_HIT_MISS;
}
// First boundary with t >= eps of A u B, B being at most two spans held in registers (a leaf, or a plane cluster:
// the sky planes): UnionSpanIterator::next (src/union.cpp:84-134) with every span scanned the moment it is
// emitted, and nothing stored — the scan stops at the first boundary anyway.
__device__ __forceinline__ int rlUnionFirstHit(const REval &ev, const RList &A, int nB, SpanRec b0, const SpanRec &b1, float &t, uint32_t &tag)
{
    if(A.n == 0)
    { // aEnded from the start (src/union.cpp:88-95): the stream is B's spans as they are.  Two thirds of the camera rays
      // of a scene inside a sky box end here (ncu: the loop below ran 2.9 times per C5 camera ray at ~40 instructions
      // a turn, profiles/r02k_full_c5.txt).
        if(nB == 0)
            return PTB_HIT_MISS;
        int k = rlScanSpan(b0, t, tag);
        if(k < 0 && nB >= 2)
            k = rlScanSpan(b1, t, tag);
        return k < 0 ? PTB_HIT_MISS : k;
    }
    int ia = 1, ib = 0;
    bool aEnded = false, bEnded = nB == 0;
    SpanRec sa = rlGet(ev, A, 0), sb = b0;
    if(!bEnded)
    {
        ib = 1;
        // B's first span ends behind the origin and before A begins — the lower span of a sky box seen from inside,
        // src/test.cpp:134-140 — so the machine below would emit it first (:104-107) and the scan would pass over
        // it: drop it here.
        if(b0.t0 < PTB_EPS && b0.t1 < PTB_EPS && !(sa.t1 < b0.t0) && b0.t1 < sa.t0)
        {
            if(nB >= 2)
                sb = b1, ib = 2;
            else
                bEnded = true;
        }
    }
    for(;;)
    {
        bool fromB; // which operand's current span goes out next, in the order of the reference's tests
        if(aEnded)
        { // src/union.cpp:88-95
            if(bEnded)
                return PTB_HIT_MISS;
            fromB = true;
        }
        else if(bEnded || sa.t1 < sb.t0)
            fromB = false; // :96-103
        else if(sb.t1 < sa.t0)
            fromB = true; // :104-107
        else
        {
            // They overlap.  The span that starts first absorbs the other and from then on only its END moves (B is a
            // leaf's or a cluster's list: in order; when B absorbs, no further span of A may be left to take over),
            // so when it starts at or beyond eps the scan's answer is known now: the merged span's start.
            const bool aFirst = sa.t0 < sb.t0;
            const float w0 = aFirst ? sa.t0 : sb.t0;
            if(w0 >= PTB_EPS && (aFirst || ia >= A.n))
            {
                if(w0 >= PTB_MAX_VALUE)
                    return PTB_HIT_MISS;
                t = w0, tag = (aFirst ? sa.tags : sb.tags) & 0xFFFFu; // startTag
                return PTB_HIT_ENTER;
            }
            if(aFirst)
            { // A absorbs B (:113-123)
                if(sa.t1 < sb.t1)
                    copyEndFromEnd(sa, sb);
                if(ib < nB)
                    sb = b1, ib++;
                else
                    bEnded = true;
            }
            else
            { // B absorbs A (:124-133)
                if(sa.t1 > sb.t1)
                    copyEndFromEnd(sb, sa);
                if(ia < A.n)
                    sa = rlGet(ev, A, ia++);
                else
                    aEnded = true;
            }
            continue;
        }
        SpanRec cur; // (field by field: a `fromB ? sb : sa` reference becomes an address select through local memory)
        cur.t0 = fromB ? sb.t0 : sa.t0, cur.t1 = fromB ? sb.t1 : sa.t1, cur.tags = fromB ? sb.tags : sa.tags;
        const int k = rlScanSpan(cur, t, tag);
        if(k >= 0)
            return k;
        if(fromB)
        {
            if(ib < nB)
                sb = b1, ib++;
            else
                bEnded = true;
        }
        else if(ia < A.n)
            sa = rlGet(ev, A, ia++);
        else
            aEnded = true;
    }
}
// The root list into pool slots [0, n) for the ray-batch kernels (they read spans back through a FastPool); returns n.
__device__ __forceinline__ int rlToPool(REval &ev, const RList &R)
{
    if(R.n == 1)
    {
        if(ev.cap < 1)
        {
            ev.ovf = true;
            return 0;
        }
        rpoolStore(ev.poolBase, 0, R.a);
    }
    return R.n; // n >= 2: the root is the only pending list, so it starts at slot 0
}

// What the generator (ptb_specialize.cu) defines for one scene: specTrace in the renderer's module (first boundary
// with t >= eps: PTB_HIT_*, t, tag), specList in the ray-batch module (root list into pool slots [0, n), returns n).
template <bool O0>
__device__ __forceinline__ int specTrace(REval &ev, const V3 o, const V3 d, float &t, uint32_t &tag);
template <bool O0>
__device__ __forceinline__ int specList(REval &ev, const V3 o, const V3 d);

} // namespace ptb
