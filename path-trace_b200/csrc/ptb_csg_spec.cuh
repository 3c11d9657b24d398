// ptb_csg_spec.cuh — helpers of the scene-SPECIALISED evaluator.
//
// The interpreter (evalFast, ptb_csg_fast.cuh) spends more instructions on fetching and decoding
// program words than on arithmetic (ncu, C4 camera rays: 17 FP32 instructions of ~49 per transformed
// sphere).  For a scene that is going to be traced billions of times the host therefore prints the
// compiled fast stream as straight-line CUDA (ptb_specialize.cpp: one labelled block per word, jumps
// as gotos, every header, matrix and plane a literal) and compiles it with NVRTC.  The blocks call
// the SAME device functions as the interpreter (xfDisc, sphereDisc, planeFold, planeFinish,
// fastPushOperate), so results are identical; the compiler folds the decoding away and turns the
// geometry into immediates.  List pushes and operators stay ONE out-of-line function, so the code
// grows by ~40 instructions per leaf and not by three merge state machines per leaf.
#pragma once
#include "ptb_csg_fast.cuh"

namespace ptb
{

// PTB_MISS_EXIT of the interpreter: true when no lane of the warp hit the leaf and the word needs no
// pool work from anybody (the stack bits are updated here).
__device__ __forceinline__ bool specMissExit(uint32_t hx, bool hit, FastState &st)
{
    if(__any_sync(0xFFFFFFFFu, hit))
        return false;
    if(!(hx & FH_TAIL))
    {
        st.bits <<= (hx >> 6) & 1u;
        return true;
    }
    if(PTB_TAIL_CLOSED && (hx & FH_UNION_TAIL))
    { // the leaf's own action is Union or Difference: a miss changes nothing
        const uint32_t np = (hx >> 8) & 15u, mask = (2u << np) - 1u;
        const bool simple = __popc(st.bits & mask) <= 1;
        if(__all_sync(0xFFFFFFFFu, simple))
        {
            st.bits = ((st.bits >> np) & ~1u) | ((st.bits & mask) ? 1u : 0u);
            return true;
        }
    }
    return false;
}

// What the out-of-line functions need of the launch, written once per CTA by the kernel wrappers (every
// argument not passed is a register move saved at each of the ~100 call sites).
struct SpecShared
{
    float *ovf;        // span-pool overflow area of this launch (NULL when the pools fit in shared memory)
    size_t ovfStride;
    int cap;
};
__shared__ SpecShared ptbSpecShared;
__device__ __forceinline__ void specSharedInit(int cap, float *overflow, size_t ovfStride)
{
    if(threadIdx.x == 0)
        ptbSpecShared.ovf = overflow, ptbSpecShared.ovfStride = ovfStride, ptbSpecShared.cap = cap;
    // (setupFast ends with __syncthreads())
}
__device__ __forceinline__ FastPool specPool()
{ // = FastCtx::pool of setupFast<GP = true>: no program words in front of the pools
    FastPool pool;
    pool.base = (uint32_t)__cvta_generic_to_shared(ptbSmem) + 4u * threadIdx.x;
    pool.cap = ptbSpecShared.cap;
    pool.ovf = ptbSpecShared.ovf ? ptbSpecShared.ovf + (size_t)blockIdx.x * PTB_FAST_THREADS + threadIdx.x : nullptr;
    pool.ovfStride = ptbSpecShared.ovfStride;
    return pool;
}

// The out-of-line halves of a leaf block, entered by the whole warp when some lane hit the leaf (or its
// word has operators to run): everything but the discriminant.  State in and out by value (registers).
static __device__ __noinline__ uint3 specPushOperateSlow(uint32_t hx, bool hit, bool two, float lt0, float lt1, float ut0, uint32_t ltags, uint32_t utags, uint32_t bits, int top,
                                                         uint32_t sp)
{
    FastState st;
    st.bits = bits, st.top = top, st.sp = sp;
    fastPushOperate(hx, hit, two, lt0, lt1, ut0, ltags, utags, st, specPool());
    return make_uint3(st.bits, (uint32_t)st.top, st.sp);
}
static __device__ __noinline__ uint3 specSphereSlow(uint32_t hx, uint32_t tags, float a, float b, float e, uint32_t bits, int top, uint32_t sp)
{
    const bool hit = e > PTB_EPS;
    float lt0 = 0.f, lt1 = 0.f;
    if(hit)
    { // src/sphere.cpp:44-46
        const float q = sqrtApprox(e), ia = rcpApprox(a);
        lt0 = (-b - q) * ia, lt1 = (-b + q) * ia;
    }
    return specPushOperateSlow(hx, hit, false, lt0, lt1, 0.f, tags, 0u, bits, top, sp); // the ONE copy of the merge state machines
}
__device__ __forceinline__ void specSphere(uint32_t hx, uint32_t tags, float a, float b, float e, FastState &st)
{
    const uint3 r = specSphereSlow(hx, tags, a, b, e, st.bits, st.top, st.sp);
    st.bits = r.x, st.top = (int)r.y, st.sp = r.z;
}
__device__ __forceinline__ void specPushOperate(uint32_t hx, bool hit, bool two, float lt0, float lt1, float ut0, uint32_t ltags, uint32_t utags, FastState &st)
{
    const uint3 r = specPushOperateSlow(hx, hit, two, lt0, lt1, ut0, ltags, utags, st.bits, st.top, st.sp);
    st.bits = r.x, st.top = (int)r.y, st.sp = r.z;
}

} // namespace ptb
