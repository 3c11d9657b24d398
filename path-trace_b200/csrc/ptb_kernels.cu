// ptb_kernels.cu — the sm_100a kernels of the render path.
//
//   intersectKernel   ray batch -> CSG span lists -> first hit          (BASELINE config 3)
//   primaryKernel     raygen + intersect + shade, survivors -> HBM queue (wavefront stage 0)
//   bounceKernel      HBM queue -> intersect + shade -> HBM queue        (wavefront stage k)
//   resolveKernel     accumulation buffer -> mean radiance
//
// Work is branchy scalar FP32 (no tensor cores, no RT cores on B200).  Grids are persistent:
// a multiple of the SM count, each warp striding over the work; the CSG program is staged
// into shared memory once per CTA and read warp-uniformly; span lists sit in a bank-conflict
// free [slot][thread] shared-memory pool; queue records are four float4 SoA streams written
// through __ballot_sync/__popc warp-aggregated compaction so every store is coalesced.
#include <cub/device/device_radix_sort.cuh>

#include <stdlib.h>

#include <map>
#include <mutex>

// cudaFuncAttributeMaxDynamicSharedMemorySize is per kernel and device and only ever has to GROW: set it when a
// launch needs more than any before it, not at every launch (512 trace launches per C4 frame).
template <class K>
static void ensureDynamicSmem(K kernel, size_t bytes)
{
    static std::mutex lock;
    static std::map<std::pair<const void *, int>, size_t> have;
    int device = 0;
    cudaGetDevice(&device);
    std::lock_guard<std::mutex> g(lock);
    size_t &h = have[std::make_pair((const void *)kernel, device)];
    if(bytes > h)
    {
        cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        h = bytes;
    }
}

#include "ptb_launch.h"
#include "ptb_trace.cuh"
#include "ptb_shade.cuh"

namespace ptb
{

#define PTB_SHADE_THREADS 128
// CTAs per SM the shade kernels are compiled for: camera hits are coherent and want registers (128),
// bounce hits are scattered and want warps to hide their gathers (80 registers)
#ifndef PTB_SHADE_MIN_CTAS
#define PTB_SHADE_MIN_CTAS 6 /* measured on B200, C5 primary shade per frame: 4 CTAs (128 registers) 245 ms, 5: 227, 6 (80 registers): 214 */
#endif
#ifndef PTB_SHADE_MIN_CTAS_BOUNCE
#define PTB_SHADE_MIN_CTAS_BOUNCE 6
#endif
#ifndef PTB_TRACE_RAYS
#define PTB_TRACE_RAYS 1
#endif
#ifndef PTB_TRACE_MIN_CTAS
#define PTB_TRACE_MIN_CTAS 8
#endif

// ---- shared memory carve-up -----------------------------------------------------------------
size_t csgSharedBytes(const DevScene &sc, int threads, int cap, bool stageProgram)
{
    size_t bytes = stageProgram ? (size_t)sc.programWords * 16 : 0;
    bytes += (size_t)cap * threads * 12;                 // t0, t1, tags
    bytes += (size_t)sc.maxStack * threads * 2;          // stack-entry start offsets (u16)
    return (bytes + 15) & ~(size_t)15;
}
size_t fastSharedBytes(const DevScene &sc, int cap, bool stageProgram, int rays)
{
    size_t bytes = stageProgram ? (size_t)sc.fastWords * 16 : 0;
    bytes += (size_t)rays * cap * PTB_FAST_THREADS * 12;               // t0, t1, tags
    bytes += (size_t)rays * (sc.maxStack + 1) * PTB_FAST_THREADS * 4;  // list-start stack (u32)
    return (bytes + 15) & ~(size_t)15;
}
int traceRaysPerThread() { return PTB_TRACE_RAYS; }

struct CsgShared
{
    const uint4 *program;
    Pool pool;
};

__device__ __forceinline__ CsgShared setupCsg(const DevScene &sc, int cap, int stageProgram, float *overflow, size_t ovfStride)
{
    extern __shared__ uint4 smem[];
    CsgShared s;
    uint4 *p = smem;
    if(stageProgram)
    {
        for(int i = threadIdx.x; i < sc.programWords; i += blockDim.x)
            p[i] = sc.program[i];
        s.program = p;
        p += sc.programWords;
    }
    else
        s.program = sc.program;
    float *f = reinterpret_cast<float *>(p);
    int T = blockDim.x;
    s.pool.s0 = f + threadIdx.x;
    s.pool.s1 = f + (size_t)cap * T + threadIdx.x;
    s.pool.sg = reinterpret_cast<uint32_t *>(f + 2 * (size_t)cap * T) + threadIdx.x;
    s.pool.starts = reinterpret_cast<uint16_t *>(f + 3 * (size_t)cap * T) + threadIdx.x;
    s.pool.stride = T;
    s.pool.cap = cap;
    s.pool.ovf = overflow ? overflow + (size_t)blockIdx.x * T + threadIdx.x : nullptr;
    s.pool.ovfStride = ovfStride;
    __syncthreads();
    return s;
}

// ---- ray-batch intersection -------------------------------------------------------------------
// Exact policy: the generic evaluator, bit-identical to the reference.
__global__ void __launch_bounds__(128) intersectExactKernel(DevScene sc, int cap, int stageProgram, float *overflow, size_t ovfStride, const float4 *__restrict__ origins,
                                                            const float4 *__restrict__ dirs, size_t n, PtbHit *__restrict__ hits, PtbSpan *__restrict__ spans, int maxSpans)
{
    CsgShared cs = setupCsg(sc, cap, stageProgram, overflow, ovfStride);
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        float4 o4 = __ldg(&origins[i]), d4 = __ldg(&dirs[i]);
        V3 o = mk(o4.x, o4.y, o4.z), d = mk(d4.x, d4.y, d4.z);
        Hit hit;
        uint32_t tag;
        int nSpans = evalProgram<true>(cs.program, cs.pool, o, d);
        int kind = resolveFirstHit<true>(sc, cs.pool, nSpans, o, d, hit, tag);
        writeHitAndSpans<true>(sc, cs.pool, i, kind, nSpans, hit, tag, o, d, hits, spans, maxSpans);
    }
}

// Fast policy: the evaluator the renderer runs (body in ptb_trace.cuh).
template <bool GP, bool O0>
__global__ void __launch_bounds__(PTB_FAST_THREADS) intersectFastKernel(DevScene sc, int cap, float *overflow, size_t ovfStride, const float4 *__restrict__ origins,
                                                                        const float4 *__restrict__ dirs, size_t n, PtbHit *__restrict__ hits, PtbSpan *__restrict__ spans,
                                                                        int maxSpans)
{
    intersectFastBody<GP, O0>(sc, cap, overflow, ovfStride, origins, dirs, n, hits, spans, maxSpans);
}

void launchIntersect(const DevScene &sc, const CsgConfig &cfg, const float4 *origins, const float4 *dirs, size_t n, PtbHit *hits, PtbSpan *spans, int maxSpans,
                     bool exact, bool origin0, cudaStream_t stream)
{
    if(exact)
    {
        ensureDynamicSmem(intersectExactKernel, cfg.smemBytes);
        intersectExactKernel<<<cfg.blocks, cfg.threads, cfg.smemBytes, stream>>>(sc, cfg.cap, cfg.stageProgram, cfg.overflow, cfg.ovfStride, origins, dirs, n, hits, spans, maxSpans);
    }
    else
    {
#define PTB_LAUNCH_INTERSECT(GP, O0)                                                                                                                                      \
    ensureDynamicSmem(intersectFastKernel<GP, O0>, cfg.smemBytes);                                                   \
    intersectFastKernel<GP, O0><<<cfg.blocks, PTB_FAST_THREADS, cfg.smemBytes, stream>>>(sc, cfg.cap, cfg.overflow, cfg.ovfStride, origins, dirs, n, hits, spans, maxSpans)
        if(cfg.stageProgram && origin0)
        {
            PTB_LAUNCH_INTERSECT(false, true);
        }
        else if(cfg.stageProgram)
        {
            PTB_LAUNCH_INTERSECT(false, false);
        }
        else if(origin0)
        {
            PTB_LAUNCH_INTERSECT(true, true);
        }
        else
        {
            PTB_LAUNCH_INTERSECT(true, false);
        }
#undef PTB_LAUNCH_INTERSECT
    }
}

// ---- wavefront render stages --------------------------------------------------------------------
// One wavefront level = a TRACE kernel (rays -> first hit, 8 B per ray) and a SHADE kernel (hit ->
// radiance + next ray).  They are separate kernels on purpose: fused, the shading state (path
// throughput, RNG, texture interpreter) pushed the CSG loop to 94 registers / 5 CTAs per SM and
// ptxas spilled inside the loop when capped at 64 (profiles/r01c_full_c4.txt); alone, the trace
// kernel carries a ray and three words of evaluator state.
//
// Queue record (64 B, four coalesced float4 streams):
//   q0 = origin.xyz, pixel        q1 = dir.xyz, strength
//   q2 = throughput.rgb, sample   q3 = depth, rng block, -, -
// Hit record (8 B): t, kind << 16 | boundary tag.
// Trace stage kernels (bodies in ptb_trace.cuh).  A warp takes 32 consecutive rays per trip.
template <bool GP>
__global__ void __launch_bounds__(PTB_FAST_THREADS, PTB_TRACE_MIN_CTAS) tracePrimaryKernel(DevScene sc, int cap, float *overflow, size_t ovfStride, RenderPass pass,
                                                                                           uint2 *__restrict__ hits)
{
    tracePrimaryBody<GP>(sc, cap, overflow, ovfStride, pass, hits);
}
template <bool GP>
__global__ void __launch_bounds__(PTB_FAST_THREADS, PTB_TRACE_MIN_CTAS) traceQueueKernel(DevScene sc, int cap, float *overflow, size_t ovfStride, RayQueue in,
                                                                                         const unsigned *__restrict__ inCount, uint2 *__restrict__ hits)
{
    traceQueueBody<GP>(sc, cap, overflow, ovfStride, in, inCount, hits);
}

// Framebuffer accumulation: one 16-byte vector reduction per contribution.  rgb = radiance sum,
// a = completed camera samples (so a progressive reader can divide at any time).
__device__ __forceinline__ void accumulate(float4 *accum, uint32_t pixel, V3 rad, float done)
{
    if(rad.x != 0.f || rad.y != 0.f || rad.z != 0.f || done != 0.f)
        atomicAdd(&accum[pixel], make_float4(rad.x, rad.y, rad.z, done));
}

// The same for a whole warp (all 32 lanes must call it): lanes are summed over runs of equal `pixel` first — with the
// pixel-major work order a warp holds one or two pixels — and the first lane of each run issues the atomic.  Lanes
// with nothing to add pass pixel = 0xFFFFFFFF.
__device__ __forceinline__ void accumulateWarp(float4 *accum, uint32_t pixel, V3 rad, float done, unsigned lane)
{
    const uint32_t prev = __shfl_up_sync(0xFFFFFFFFu, pixel, 1);
    const bool head = lane == 0u || prev != pixel;
    const unsigned heads = __ballot_sync(0xFFFFFFFFu, head);
    float x = rad.x, y = rad.y, z = rad.z, w = done;
    if(heads == 0x00010001u || heads == 1u)
    { // The usual case of the camera-ray level: 16 (or a multiple of 32) samples per pixel and pass, so the warp holds
      // two aligned runs of 16 lanes (or one of 32).  Transposing butterfly: a lane gives away half of its values in
      // each of the first two rounds (x, y against z, w across lanes 8 apart; then one of the two left across lanes
      // 4 apart) and is left with ONE component summed over four lanes; two more rounds finish the sum over the 16
      // lanes.  5 shuffles (6 for a run of 32) + 3 to collect the float4 instead of 20 (ncu: accumulateWarp and its
      // shuffles were 78 of the 578 instructions of a camera-ray shade trip, profiles/r02k_full_c5.txt).
        const bool up8 = (lane & 8u) != 0u, up4 = (lane & 4u) != 0u;
        float k0 = up8 ? z : x, k1 = up8 ? w : y;
        k0 += __shfl_xor_sync(0xFFFFFFFFu, up8 ? x : z, 8);
        k1 += __shfl_xor_sync(0xFFFFFFFFu, up8 ? y : w, 8);
        float v = (up4 ? k1 : k0) + __shfl_xor_sync(0xFFFFFFFFu, up4 ? k0 : k1, 4);
        v += __shfl_xor_sync(0xFFFFFFFFu, v, 2);
        v += __shfl_xor_sync(0xFFFFFFFFu, v, 1);
        if(heads == 1u)
            v += __shfl_xor_sync(0xFFFFFFFFu, v, 16);
        // lanes 4c .. 4c+3 of each half now hold component c of that half's pixel
        const unsigned half = lane & 16u;
        const float sy = __shfl_sync(0xFFFFFFFFu, v, half + 4u), sz = __shfl_sync(0xFFFFFFFFu, v, half + 8u), sw = __shfl_sync(0xFFFFFFFFu, v, half + 12u);
        if(head && pixel != 0xFFFFFFFFu && (v != 0.f || sy != 0.f || sz != 0.f || sw != 0.f))
            atomicAdd(&accum[pixel], make_float4(v, sy, sz, sw));
        return;
    }
    // Runs of any length (bounce levels: the queue holds what survived, in order): segmented sum down the lanes.  (The
    // count of finished paths as a population count over the run instead of a fourth shuffled value executed as many
    // instructions and put ballot -> ffs -> popc in series: shadeKernel<bounce> 1.16 -> 1.27 ms, profiles/r02m.)
    const unsigned after = lane == 31u ? 0xFFFFFFFFu : (heads >> (lane + 1u)) | (0xFFFFFFFFu << (31u - lane)); // bit k-1: lane + k starts a new run (or is past the warp)
#pragma unroll
    for(unsigned d = 1u; d < 32u; d <<= 1)
    {
        const float tx = __shfl_down_sync(0xFFFFFFFFu, x, d), ty = __shfl_down_sync(0xFFFFFFFFu, y, d), tz = __shfl_down_sync(0xFFFFFFFFu, z, d),
                    tw = __shfl_down_sync(0xFFFFFFFFu, w, d);
        if((after & ((1u << d) - 1u)) == 0u) // lanes lane+1 .. lane+d all belong to this lane's run
            x += tx, y += ty, z += tz, w += tw;
    }
    if(head && pixel != 0xFFFFFFFFu && (x != 0.f || y != 0.f || z != 0.f || w != 0.f))
        atomicAdd(&accum[pixel], make_float4(x, y, z, w));
}

// One path vertex from a packed hit record (t, kind << 16 | boundary tag): normal through the un-folded
// transform chain, material, relative ior (path-trace.h:72-95), then shadeVertex.  Adds the vertex's radiance to
// `rad`; true while the path goes on (ps then holds the next ray).
__device__ __forceinline__ bool shadeHitRecord(const DevScene &sc, const RenderPass &pass, PathState &ps, uint2 h, V3 &rad)
{
    const int kind = (int)(h.y >> 16);
    if(kind == PTB_HIT_MISS)
        return false;
    const uint32_t tag = h.y & 0xFFFFu;
    Hit hit;
    hit.t = __uint_as_float(h.x);
    hit.mat = __ldg(&sc.primMeta[tag >> 1]).y;
    hit.n = mk(0.f, 0.f, 0.f), hit.ior = 1.f;
    if(!(sc.materials[hit.mat].constMask & PTB_MAT_TERMINAL))
    { // (a terminal material — every sky plane — never looks at the normal: two thirds of the camera hits of C5)
        V3 n = boundaryNormal<false>(sc, tag, hit.t, ps.o, ps.d);
        if(kind == PTB_HIT_ENTER)
            hit.n = n, hit.ior = sc.materials[hit.mat].iorIn; // path-trace.h:72-81
        else
            hit.n = neg(n), hit.ior = sc.materials[hit.mat].ior; // path-trace.h:86-95
    }
    Philox rng;
    rngInit(rng, pass.seed, ps.pixel, ps.sample, ps.rngBlock);
    const bool alive = shadeVertex(sc, ps, hit, rng, pass.flags, rad);
    ps.rngBlock = rng.block;
    return alive;
}

// Shade stage of one wavefront level: hit record -> normal, material, emitted radiance, next ray.
// PRIMARY: path state is regenerated from the sample index; otherwise it is read from `in`.
// Survivors are appended to `out` with __ballot_sync/__popc warp aggregation (one atomic per warp,
// coalesced stores).
template <bool PRIMARY, int MINCTAS = PRIMARY ? PTB_SHADE_MIN_CTAS : PTB_SHADE_MIN_CTAS_BOUNCE>
__global__ void __launch_bounds__(PTB_SHADE_THREADS, MINCTAS) shadeKernel(DevScene sc, RenderPass pass, float4 *accum, const uint2 *__restrict__ hits, RayQueue in,
                                                                 const unsigned *__restrict__ inCount, RayQueue out, unsigned *outCount)
{
    const unsigned nPix = (unsigned)pass.rw * (unsigned)pass.rh;
    // 32-bit work indices: the host keeps paths per pass below 2^32 - 2^24 (renderPasses) and a grid has < 2^24 threads
    const unsigned total = PRIMARY ? nPix * (unsigned)pass.samplesThisPass : *inCount;
    const unsigned stride = gridDim.x * PTB_SHADE_THREADS;
    __shared__ float4 stageAll[PTB_SHADE_THREADS / 32][4][64];
    float4(*stage)[64] = stageAll[threadIdx.x >> 5];
    const unsigned lane = threadIdx.x & 31u;
    unsigned staged = 0; // warp-uniform: records waiting in this warp's staging buffer (< 32 between trips)
    // software pipeline: the next trip's hit record (and queue record) is requested before this
    // trip is shaded, so its HBM latency hides behind the shading (ncu: 37-54 % of the stall samples
    // of this kernel were long-scoreboard waits on those loads)
    const unsigned first = blockIdx.x * PTB_SHADE_THREADS + (threadIdx.x & ~31u);
    uint2 hNext = make_uint2(0u, 0u);
    float4 n0 = make_float4(0.f, 0.f, 0.f, 0.f), n1 = n0, n2 = n0, n3 = n0;
    if(first + lane < total)
    {
        hNext = __ldcs(&hits[first + lane]);
        if(!PRIMARY)
            n0 = __ldcs(&in.q0[first + lane]), n1 = __ldcs(&in.q1[first + lane]), n2 = __ldcs(&in.q2[first + lane]), n3 = __ldcs(&in.q3[first + lane]);
    }
    for(unsigned base = first; base < total; base += stride)
    {
        const unsigned i = base + lane;
        bool alive = false;
        PathState ps;
        V3 rad = mk(0.f, 0.f, 0.f);
        uint32_t accPixel = 0xFFFFFFFFu;
        const uint2 h = hNext;
        const float4 c0 = n0, c1 = n1, c2 = n2, c3 = n3;
        if(i + stride < total)
        {
            hNext = __ldcs(&hits[i + stride]);
            if(!PRIMARY)
                n0 = __ldcs(&in.q0[i + stride]), n1 = __ldcs(&in.q1[i + stride]), n2 = __ldcs(&in.q2[i + stride]), n3 = __ldcs(&in.q3[i + stride]);
        }
        if(i < total)
        {
            if(PRIMARY)
                cameraSample(pass, nPix, i, ps);
            else
            {
                ps.o = mk(c0.x, c0.y, c0.z), ps.pixel = __float_as_uint(c0.w);
                ps.d = mk(c1.x, c1.y, c1.z), ps.strength = c1.w;
                ps.thr = mk(c2.x, c2.y, c2.z), ps.sample = __float_as_uint(c2.w);
                ps.depth = __float_as_uint(c3.x), ps.rngBlock = __float_as_uint(c3.y);
            }
            alive = shadeHitRecord(sc, pass, ps, h, rad);
            accPixel = ps.pixel;
        }
        accumulateWarp(accum, accPixel, rad, (accPixel != 0xFFFFFFFFu && !alive) ? 1.f : 0.f, lane);
        // survivors go to the warp's staging buffer in shared memory; full groups of 32 are flushed
        // with ONE atomic on the queue counter and four fully coalesced 512-byte stores.  (ncu: with
        // one atomic per warp iteration 20 % of the kernel's stall samples sat on its return value.)
        const unsigned mask = __ballot_sync(0xFFFFFFFFu, alive);
        if(alive)
        {
            unsigned at = staged + __popc(mask & ((1u << lane) - 1u));
            stage[0][at] = make_float4(ps.o.x, ps.o.y, ps.o.z, __uint_as_float(ps.pixel));
            stage[1][at] = make_float4(ps.d.x, ps.d.y, ps.d.z, ps.strength);
            stage[2][at] = make_float4(ps.thr.x, ps.thr.y, ps.thr.z, __uint_as_float(ps.sample));
            stage[3][at] = make_float4(__uint_as_float(ps.depth), __uint_as_float(ps.rngBlock), 0.f, 0.f);
        }
        staged += __popc(mask);
        __syncwarp();
        if(staged >= 32u)
        {
            unsigned slot = 0;
            if(lane == 0u)
                slot = atomicAdd(outCount, 32u);
            float4 a = stage[0][lane], b = stage[1][lane], c = stage[2][lane], e = stage[3][lane];
            slot = __shfl_sync(0xFFFFFFFFu, slot, 0) + lane;
            out.q0[slot] = a, out.q1[slot] = b, out.q2[slot] = c, out.q3[slot] = e;
            staged -= 32u;
            __syncwarp();
            if(lane < staged)
            { // at most 31 left: move them to the front
                a = stage[0][32u + lane], b = stage[1][32u + lane], c = stage[2][32u + lane], e = stage[3][32u + lane];
                stage[0][lane] = a, stage[1][lane] = b, stage[2][lane] = c, stage[3][lane] = e;
            }
            __syncwarp();
        }
    }
    if(staged)
    {
        unsigned slot = 0;
        if(lane == 0u)
            slot = atomicAdd(outCount, staged);
        slot = __shfl_sync(0xFFFFFFFFu, slot, 0) + lane;
        if(lane < staged)
            out.q0[slot] = stage[0][lane], out.q1[slot] = stage[1][lane], out.q2[slot] = stage[2][lane], out.q3[slot] = stage[3][lane];
    }
}

// Tail of a pass (bounce levels that hold a few thousand rays at most): every queued ray is followed to the end
// of its path inside ONE kernel — trace, shade, trace, ... — instead of a trace / shade launch pair per level on
// nearly empty queues.  Same device functions, same RNG blocks, hence the same radiance as the level-by-level
// wavefront (up to the order of the float atomics).  All 32 lanes run the evaluator every trip (its SKIP words and
// miss exits vote); lanes whose path has ended carry a null ray.  Register-heavy (trace + shade state together),
// which is why the big levels are NOT run this way.
template <bool GP>
__global__ void __launch_bounds__(PTB_FAST_THREADS) tailKernel(DevScene sc, int cap, float *overflow, size_t ovfStride, RenderPass pass, float4 *accum, RayQueue in,
                                                               const unsigned *__restrict__ inCount, unsigned *extraRays)
{
    const unsigned total = *inCount;
    if(total == 0u)
        return;
    FastCtx cx = setupFast<GP, 1>(sc, cap, overflow, ovfStride);
    const unsigned stride = gridDim.x * PTB_FAST_THREADS;
    const unsigned lane = threadIdx.x & 31u;
    unsigned extra = 0u;
    for(unsigned base = blockIdx.x * PTB_FAST_THREADS + (threadIdx.x & ~31u); base < total; base += stride)
    {
        const unsigned i = base + lane;
        bool alive = i < total;
        PathState ps;
        ps.o = ps.d = ps.thr = mk(0.f, 0.f, 0.f);
        ps.strength = 0.f, ps.pixel = ps.sample = ps.depth = ps.rngBlock = 0u;
        if(alive)
            queueLoad(in, (unsigned)i, ps);
        bool first = true;
        while(__any_sync(0xFFFFFFFFu, alive))
        {
            V3 o[1] = {alive ? ps.o : mk(0.f, 0.f, 0.f)}, d[1] = {alive ? ps.d : mk(0.f, 0.f, 0.f)};
            int tops[1];
            evalFast<GP, false, 1>(cx, sc.fastProgram, o, d, tops);
            if(alive)
            {
                float t = -1.f;
                uint32_t tag = 0u;
                const int kind = firstHit(cx.pool, tops[0], t, tag);
                extra += first ? 0u : 1u;
                V3 rad = mk(0.f, 0.f, 0.f);
                alive = shadeHitRecord(sc, pass, ps, make_uint2(__float_as_uint(t), ((uint32_t)kind << 16) | tag), rad);
                accumulate(accum, ps.pixel, rad, alive ? 0.f : 1.f);
            }
            first = false;
        }
    }
#pragma unroll
    for(int k = 16; k; k >>= 1)
        extra += __shfl_xor_sync(0xFFFFFFFFu, extra, k);
    if(lane == 0u && extra)
        atomicAdd(extraRays, extra);
}
void launchTail(const DevScene &sc, const CsgConfig &cfg, int blocks, const RenderPass &pass, float4 *accum, RayQueue in, const unsigned *inCount, unsigned *extraRays, cudaStream_t stream)
{
    if(cfg.stageProgram)
        tailKernel<false><<<blocks, PTB_FAST_THREADS, cfg.smemBytes, stream>>>(sc, cfg.cap, cfg.overflow, cfg.ovfStride, pass, accum, in, inCount, extraRays);
    else
        tailKernel<true><<<blocks, PTB_FAST_THREADS, cfg.smemBytes, stream>>>(sc, cfg.cap, cfg.overflow, cfg.ovfStride, pass, accum, in, inCount, extraRays);
}
int tailBlocksPerSm(bool globalProgram, size_t smemBytes)
{
    int n = 0;
    cudaError_t e;
    if(globalProgram)
    {
        ensureDynamicSmem(tailKernel<true>, smemBytes);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, tailKernel<true>, PTB_FAST_THREADS, smemBytes);
    }
    else
    {
        ensureDynamicSmem(tailKernel<false>, smemBytes);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, tailKernel<false>, PTB_FAST_THREADS, smemBytes);
    }
    if(e != cudaSuccess)
    {
        cudaGetLastError();
        return 0;
    }
    return n;
}

void launchTracePrimary(const DevScene &sc, const CsgConfig &cfg, const RenderPass &pass, uint2 *hits, cudaStream_t stream)
{
    if(cfg.stageProgram)
    {
        ensureDynamicSmem(tracePrimaryKernel<false>, cfg.smemBytes);
        tracePrimaryKernel<false><<<cfg.blocks, PTB_FAST_THREADS, cfg.smemBytes, stream>>>(sc, cfg.cap, cfg.overflow, cfg.ovfStride, pass, hits);
    }
    else
    {
        ensureDynamicSmem(tracePrimaryKernel<true>, cfg.smemBytes);
        tracePrimaryKernel<true><<<cfg.blocks, PTB_FAST_THREADS, cfg.smemBytes, stream>>>(sc, cfg.cap, cfg.overflow, cfg.ovfStride, pass, hits);
    }
}
void launchTraceQueue(const DevScene &sc, const CsgConfig &cfg, RayQueue in, const unsigned *inCount, uint2 *hits, cudaStream_t stream)
{
    if(cfg.stageProgram)
    {
        ensureDynamicSmem(traceQueueKernel<false>, cfg.smemBytes);
        traceQueueKernel<false><<<cfg.blocks, PTB_FAST_THREADS, cfg.smemBytes, stream>>>(sc, cfg.cap, cfg.overflow, cfg.ovfStride, in, inCount, hits);
    }
    else
    {
        ensureDynamicSmem(traceQueueKernel<true>, cfg.smemBytes);
        traceQueueKernel<true><<<cfg.blocks, PTB_FAST_THREADS, cfg.smemBytes, stream>>>(sc, cfg.cap, cfg.overflow, cfg.ovfStride, in, inCount, hits);
    }
}
// Resident CTAs per SM the shade kernels are compiled for, selectable for measurements: PTB_SHADE_CTAS (camera hits:
// 4 = 128 registers, 6 = 80 (default), 7, 8 = 64) and PTB_SHADE_CTAS_BOUNCE (4, 6, 8).
static int shadeVariant(bool primary)
{
    static const int p = getenv("PTB_SHADE_CTAS") ? atoi(getenv("PTB_SHADE_CTAS")) : PTB_SHADE_MIN_CTAS;
    static const int b = getenv("PTB_SHADE_CTAS_BOUNCE") ? atoi(getenv("PTB_SHADE_CTAS_BOUNCE")) : PTB_SHADE_MIN_CTAS_BOUNCE;
    return primary ? p : b;
}
void launchShade(const DevScene &sc, int blocks, bool primary, const RenderPass &pass, float4 *accum, const uint2 *hits, RayQueue in, const unsigned *inCount, RayQueue out,
                 unsigned *outCount, cudaStream_t stream)
{
    const int variant = shadeVariant(primary);
#define PTB_SHADE_LAUNCH(P, C) shadeKernel<P, C><<<blocks, PTB_SHADE_THREADS, 0, stream>>>(sc, pass, accum, hits, in, inCount, out, outCount)
    if(primary)
    {
        if(variant == 4)
            PTB_SHADE_LAUNCH(true, 4);
        else if(variant == 7)
            PTB_SHADE_LAUNCH(true, 7);
        else if(variant == 8)
            PTB_SHADE_LAUNCH(true, 8);
        else
            PTB_SHADE_LAUNCH(true, PTB_SHADE_MIN_CTAS);
    }
    else
    {
        if(variant == 8)
            PTB_SHADE_LAUNCH(false, 8);
        else if(variant == 4)
            PTB_SHADE_LAUNCH(false, 4);
        else
            PTB_SHADE_LAUNCH(false, PTB_SHADE_MIN_CTAS_BOUNCE);
    }
#undef PTB_SHADE_LAUNCH
}
int shadeBlocksPerSm(bool primary)
{
    int n = 0;
    const int variant = shadeVariant(primary);
    cudaError_t e;
    if(primary)
        e = variant == 4 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, shadeKernel<true, 4>, PTB_SHADE_THREADS, 0)
            : variant == 7 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, shadeKernel<true, 7>, PTB_SHADE_THREADS, 0)
            : variant == 8 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, shadeKernel<true, 8>, PTB_SHADE_THREADS, 0)
                           : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, shadeKernel<true, PTB_SHADE_MIN_CTAS>, PTB_SHADE_THREADS, 0);
    else
        e = variant == 8 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, shadeKernel<false, 8>, PTB_SHADE_THREADS, 0)
            : variant == 4 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, shadeKernel<false, 4>, PTB_SHADE_THREADS, 0)
                           : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, shadeKernel<false, PTB_SHADE_MIN_CTAS_BOUNCE>, PTB_SHADE_THREADS, 0);
    if(e != cudaSuccess)
    {
        cudaGetLastError();
        return 0;
    }
    return n;
}

// ---- resolve ------------------------------------------------------------------------------------
__global__ void resolveKernel(const float4 *__restrict__ accum, float4 *__restrict__ rgba, size_t n)
{
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        float4 a = accum[i];
        float inv = a.w > 0.f ? 1.f / a.w : 0.f;
        rgba[i] = make_float4(a.x * inv, a.y * inv, a.z * inv, a.w > 0.f ? 1.f : 0.f);
    }
}
void launchResolve(const float4 *accum, float4 *rgba, size_t nPixels, cudaStream_t stream)
{
    int blocks = (int)((nPixels + 255) / 256);
    if(blocks > 148 * 8)
        blocks = 148 * 8;
    if(blocks < 1)
        blocks = 1;
    resolveKernel<<<blocks, 256, 0, stream>>>(accum, rgba, nPixels);
}

// ---- 8-bit display conversion (reference src/test.cpp:992-995) ------------------------------------
// byte = max(0, min(255, (int)floor(256 * c / count))) per channel, float arithmetic in that order;
// count = the pixel's finished samples (alpha of an accumulation buffer), or 1 for a resolved image.
// 16 B read + 4 B written per pixel: HBM-bound, coalesced, one pixel per thread and trip.
__global__ void display8Kernel(const float4 *__restrict__ rgba, uchar4 *__restrict__ out, size_t n, int divideByAlpha)
{
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        float4 a = __ldcs(&rgba[i]);
        float count = divideByAlpha ? a.w : 1.f;
        uchar4 o = make_uchar4(0, 0, 0, 255);
        if(count > 0.f)
        {
            float r = floorf(__fdiv_rn(__fmul_rn(256.f, a.x), count)), g = floorf(__fdiv_rn(__fmul_rn(256.f, a.y), count)),
                  b = floorf(__fdiv_rn(__fmul_rn(256.f, a.z), count));
            // (int) of a NaN or of a value outside the int range is INT_MIN on the x86-64 reference build
            // (cvttss2si's "integer indefinite"), which the clamp then turns into 0
            o.x = (unsigned char)((r >= 2147483648.f || r < -2147483648.f || r != r) ? 0 : max(0, min(255, (int)r)));
            o.y = (unsigned char)((g >= 2147483648.f || g < -2147483648.f || g != g) ? 0 : max(0, min(255, (int)g)));
            o.z = (unsigned char)((b >= 2147483648.f || b < -2147483648.f || b != b) ? 0 : max(0, min(255, (int)b)));
        }
        out[i] = o;
    }
}
void launchDisplay8(const float4 *rgba, uchar4 *out, size_t nPixels, bool divideByAlpha, cudaStream_t stream)
{
    int blocks = (int)((nPixels + 255) / 256);
    if(blocks > 148 * 8)
        blocks = 148 * 8;
    if(blocks < 1)
        blocks = 1;
    display8Kernel<<<blocks, 256, 0, stream>>>(rgba, out, nPixels, divideByAlpha ? 1 : 0);
}

// ---- adaptive quadtree (reference src/test.cpp:423-499, RenderBlock::renderSquare) ---------------
// The reference walks each 32x32 tile depth-first; the walk only ever reads TRACED lattice points
// (an interpolated value is never a corner of a later square: interpolation writes the half-open
// square [x, x+s) x [y, y+s), corners of other squares lie on its closed far edges or were traced
// earlier), so one level of the tree can be decided for all tiles at once.  Arithmetic is the
// reference's, un-fused: Color = Vector3D, dot(v, v) = (x*x + y*y) + z*z (include/vector3d.h:102-110).
__device__ __forceinline__ float absSquaredRn(float x, float y, float z) { return __fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), __fmul_rn(z, z)); }
// colorCloseEnough (src/test.cpp:437-440): |a - b|^2 <= delta^2 * |a|^2
__device__ __forceinline__ bool colorCloseEnough(float4 a, float4 b, float delta2)
{
    return absSquaredRn(__fsub_rn(a.x, b.x), __fsub_rn(a.y, b.y), __fsub_rn(a.z, b.z)) <= __fmul_rn(delta2, absSquaredRn(a.x, a.y, a.z));
}
// a + f * (b - a) per channel (src/test.cpp:428-434)
__device__ __forceinline__ float4 lerpRn(float4 a, float4 b, float f)
{
    return make_float4(__fadd_rn(a.x, __fmul_rn(f, __fsub_rn(b.x, a.x))), __fadd_rn(a.y, __fmul_rn(f, __fsub_rn(b.y, a.y))), __fadd_rn(a.z, __fmul_rn(f, __fsub_rn(b.z, a.z))), 2.f);
}

__global__ void adaptiveDecideKernel(AdaptiveLevel lv)
{
    const unsigned stride = gridDim.x * blockDim.x;
    for(unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < lv.nSquares; i += stride)
    {
        const uint32_t packed = lv.squares[i];
        const int x = (int)(packed & 0xFFFFu), y = (int)(packed >> 16), s = lv.size;
        if(x > lv.width || y > lv.height) // src/test.cpp:468-471
            continue;
        const float4 tl = lv.value[(size_t)y * lv.pitch + x], tr = lv.value[(size_t)y * lv.pitch + x + s];
        const float4 bl = lv.value[(size_t)(y + s) * lv.pitch + x], br = lv.value[(size_t)(y + s) * lv.pitch + x + s];
        const float d2 = lv.delta2;
        if(colorCloseEnough(tl, tr, d2) && colorCloseEnough(tl, bl, d2) && colorCloseEnough(tl, br, d2) && colorCloseEnough(tr, bl, d2) && colorCloseEnough(tr, br, d2) &&
           colorCloseEnough(bl, br, d2) && s <= lv.maxSampleSize)
        { // interpolateSquare, src/test.cpp:423-436
            atomicAdd(&lv.counters[2], 1u);
            const int visW = min(s, lv.width - x), visH = min(s, lv.height - y);
            if(visW > 0 && visH > 0)
                atomicAdd(&lv.counters[3], (unsigned)(visW * visH));
            for(int yy = 0; yy < s && y + yy < lv.height; yy++)
            {
                const float fy = __fdiv_rn((float)yy, (float)s);
                const float4 l = lerpRn(tl, bl, fy), r = lerpRn(tr, br, fy);
                for(int xx = 0; xx < s && x + xx < lv.width; xx++)
                    lv.interp[(size_t)(y + yy) * lv.pitch + x + xx] = lerpRn(l, r, __fdiv_rn((float)xx, (float)s));
            }
            continue;
        }
        // subdivide (src/test.cpp:486-498): five more lattice points, four squares of half the size
        const int h = s / 2, cx = x + h, cy = y + h;
        const int mx[5] = {cx, x, cx, x + s, cx}, my[5] = {y, cy, cy, cy, y + s};
#pragma unroll
        for(int k = 0; k < 5; k++)
            if(atomicExch(&lv.requested[(size_t)my[k] * lv.pitch + mx[k]], 1u) == 0u)
                lv.pixelList[atomicAdd(&lv.counters[1], 1u)] = ((uint32_t)my[k] << 16) | (uint32_t)mx[k];
        if(h > 1)
        { // squares of size 1 only re-state their traced top-left pixel (src/test.cpp:472-476)
            const unsigned at = atomicAdd(&lv.counters[0], 4u);
            lv.childSquares[at] = ((uint32_t)y << 16) | (uint32_t)x, lv.childSquares[at + 1] = ((uint32_t)y << 16) | (uint32_t)cx;
            lv.childSquares[at + 2] = ((uint32_t)cy << 16) | (uint32_t)x, lv.childSquares[at + 3] = ((uint32_t)cy << 16) | (uint32_t)cx;
        }
    }
}
void launchAdaptiveDecide(const AdaptiveLevel &lv, cudaStream_t stream)
{
    if(lv.nSquares == 0u)
        return;
    int blocks = (int)((lv.nSquares + 127u) / 128u);
    adaptiveDecideKernel<<<blocks > 148 * 16 ? 148 * 16 : blocks, 128, 0, stream>>>(lv);
}

size_t adaptiveSortList(uint32_t *list, uint32_t *alt, unsigned n, void *temp, size_t tempBytes, cudaStream_t stream)
{
    size_t need = 0;
    cub::DeviceRadixSort::SortKeys(nullptr, need, list, alt, (int)n, 0, 32, stream);
    if(tempBytes == 0 || n < 2u)
        return need;
    if(tempBytes < need)
        return 0;
    cub::DeviceRadixSort::SortKeys(temp, need, list, alt, (int)n, 0, 32, stream);
    cudaMemcpyAsync(list, alt, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream);
    return need;
}

__global__ void adaptiveResolveListKernel(const float4 *__restrict__ accum, float4 *__restrict__ value, const uint32_t *__restrict__ pixelList, unsigned n, int pitch)
{
    const unsigned stride = gridDim.x * blockDim.x;
    for(unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        const uint32_t packed = pixelList[i];
        const size_t at = (size_t)(packed >> 16) * pitch + (packed & 0xFFFFu);
        const float4 a = accum[at];
        const float inv = a.w > 0.f ? 1.f / a.w : 0.f;
        value[at] = make_float4(a.x * inv, a.y * inv, a.z * inv, 1.f);
    }
}
void launchAdaptiveResolveList(const float4 *accum, float4 *value, const uint32_t *pixelList, unsigned n, int pitch, cudaStream_t stream)
{
    if(n == 0u)
        return;
    int blocks = (int)((n + 255u) / 256u);
    adaptiveResolveListKernel<<<blocks > 148 * 8 ? 148 * 8 : blocks, 256, 0, stream>>>(accum, value, pixelList, n, pitch);
}

// out[i] = mean radiance of listed pixel i (y << 16 | x) of an accumulation frame
__global__ void gatherPixelsKernel(const float4 *__restrict__ accum, const uint32_t *__restrict__ pixelList, unsigned n, int pitch, float4 *__restrict__ out)
{
    const unsigned stride = gridDim.x * blockDim.x;
    for(unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        const uint32_t packed = pixelList[i];
        const float4 a = accum[(size_t)(packed >> 16) * pitch + (packed & 0xFFFFu)];
        const float inv = a.w > 0.f ? 1.f / a.w : 0.f;
        out[i] = make_float4(a.x * inv, a.y * inv, a.z * inv, a.w > 0.f ? 1.f : 0.f);
    }
}
void launchGatherPixels(const float4 *accum, const uint32_t *pixelList, unsigned n, int pitch, float4 *out, cudaStream_t stream)
{
    if(n == 0u)
        return;
    int blocks = (int)((n + 255u) / 256u);
    gatherPixelsKernel<<<blocks > 148 * 8 ? 148 * 8 : blocks, 256, 0, stream>>>(accum, pixelList, n, pitch, out);
}

__global__ void adaptiveAssembleKernel(const float4 *__restrict__ value, const float4 *__restrict__ interp, int pitch, int width, int height, float4 *__restrict__ rgba,
                                       uint8_t *__restrict__ source)
{
    const size_t n = (size_t)width * height, stride = (size_t)gridDim.x * blockDim.x;
    for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        const size_t y = i / (size_t)width, x = i - y * (size_t)width, at = y * pitch + x;
        const float4 t = interp[at];
        const float4 v = t.w == 2.f ? t : value[at];
        rgba[i] = make_float4(v.x, v.y, v.z, v.w > 0.f ? 1.f : 0.f);
        if(source)
            source[i] = (uint8_t)v.w;
    }
}
void launchAdaptiveAssemble(const float4 *value, const float4 *interp, int pitch, int width, int height, float4 *rgba, uint8_t *source, cudaStream_t stream)
{
    size_t n = (size_t)width * height;
    int blocks = (int)((n + 255) / 256);
    adaptiveAssembleKernel<<<blocks > 148 * 8 ? 148 * 8 : (blocks < 1 ? 1 : blocks), 256, 0, stream>>>(value, interp, pitch, width, height, rgba, source);
}

// ---- inspection kernels (known-answer tests through the C ABI) ----------------------------------
__global__ void textureEvalKernel(DevScene sc, int root, const float *__restrict__ pos, size_t n, float *__restrict__ rgb, float *__restrict__ f, uint32_t flags)
{
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        V3 p = mk(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
        V3 c;
        float v;
        evalTexture(sc, root, p, false, flags & PTB_FLAG_BILINEAR, c, v);
        if(rgb)
            rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
        if(f)
            f[i] = v;
    }
}
void launchTextureEval(const DevScene &sc, int root, const float *pos, size_t n, float *rgb, float *f, uint32_t flags, cudaStream_t stream)
{
    int blocks = (int)((n + 127) / 128);
    blocks = blocks < 1 ? 1 : (blocks > 1184 ? 1184 : blocks);
    textureEvalKernel<<<blocks, 128, 0, stream>>>(sc, root, pos, n, rgb, f, flags);
}

// Draws n scatter directions for one (normal, mirror vector, scatter coefficient) so the lobe
// sampler can be checked statistically against the reference's rejection loops.  A direction
// of (0,0,0) marks a dropped vertex (the 1000-try give-up).
__global__ void scatterSampleKernel(V3 normal, V3 refl, float scatter, size_t n, uint64_t seed, uint32_t flags, float *__restrict__ dirs)
{
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        Philox rng;
        rngInit(rng, seed, (uint32_t)i, (uint32_t)(i >> 32), 0u);
        V3 d = mk(0.f, 0.f, 0.f);
        bool ok = (flags & PTB_FLAG_REJECTION) ? scatterRejection(rng, normal, refl, scatter, d) : scatterCap(rng, normal, refl, scatter, d);
        if(!ok)
            d = mk(0.f, 0.f, 0.f);
        dirs[3 * i] = d.x, dirs[3 * i + 1] = d.y, dirs[3 * i + 2] = d.z;
    }
}
void launchScatterSample(const float *normal, const float *refl, float scatter, size_t n, uint64_t seed, uint32_t flags, float *dirs, cudaStream_t stream)
{
    int blocks = (int)((n + 127) / 128);
    blocks = blocks < 1 ? 1 : (blocks > 1184 ? 1184 : blocks);
    scatterSampleKernel<<<blocks, 128, 0, stream>>>(mk(normal[0], normal[1], normal[2]), mk(refl[0], refl[1], refl[2]), scatter, n, seed, flags, dirs);
}

// Vector3D::reflect / refract / refractStrength as the shade kernel computes them (ptb_shade.cuh), for
// known-answer tests against the reference (include/vector3d.h:186-214).  op 0: reflect -> 3 floats,
// 1: refract -> 3 floats, 2: refractStrength -> 1 float.
__global__ void vectorEvalKernel(int op, const float *__restrict__ v, const float *__restrict__ nrm, const float *__restrict__ ior, size_t n, float *__restrict__ out)
{
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        const V3 d = mk(v[3 * i], v[3 * i + 1], v[3 * i + 2]), nn = mk(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]);
        if(op == 2)
            out[i] = refractStrength(d, ior[i], nn);
        else
        {
            const V3 r = op == 0 ? vreflect(d, nn) : vrefract(d, ior[i], nn);
            out[3 * i] = r.x, out[3 * i + 1] = r.y, out[3 * i + 2] = r.z;
        }
    }
}
void launchVectorEval(int op, const float *v, const float *nrm, const float *ior, size_t n, float *out, cudaStream_t stream)
{
    int blocks = (int)((n + 127) / 128);
    blocks = blocks < 1 ? 1 : (blocks > 1184 ? 1184 : blocks);
    vectorEvalKernel<<<blocks, 128, 0, stream>>>(op, v, nrm, ior, n, out);
}

// The random words the render kernels draw, for statistical tests (the reference's LCG,
// include/path-trace.h:21-54, is replaced by counter-based generators; parity is statistical).
//   which 0: cameraJitter(seed, pixel, sample)        -> 2 words (sub-pixel x, y)
//   which 1: Philox block `block` of (pixel, sample)  -> 4 words (shading decisions)
__global__ void rngWordsKernel(int which, uint64_t seed, const uint32_t *__restrict__ pixel, const uint32_t *__restrict__ sample, const uint32_t *__restrict__ block, size_t n,
                               uint32_t *__restrict__ out)
{
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        if(which == 0)
        {
            const uint2 u = cameraJitter(seed, pixel[i], sample[i]);
            out[2 * i] = u.x, out[2 * i + 1] = u.y;
        }
        else
        {
            Philox rng;
            rngInit(rng, seed, pixel[i], sample[i], block ? block[i] : 0u);
            const uint4 u = rngNext4(rng);
            out[4 * i] = u.x, out[4 * i + 1] = u.y, out[4 * i + 2] = u.z, out[4 * i + 3] = u.w;
        }
    }
}
void launchRngWords(int which, uint64_t seed, const uint32_t *pixel, const uint32_t *sample, const uint32_t *block, size_t n, uint32_t *out, cudaStream_t stream)
{
    int blocks = (int)((n + 255) / 256);
    blocks = blocks < 1 ? 1 : (blocks > 1184 ? 1184 : blocks);
    rngWordsKernel<<<blocks, 256, 0, stream>>>(which, seed, pixel, sample, block, n, out);
}

// ---- FP32 issue-peak microbenchmark (roofline denominator) -----------------------
// Eight independent FFMA chains per thread: enough ILP to saturate both FP32 datapaths.
__global__ void __launch_bounds__(256) fmaPeakKernel(float *out, int iters, float a, float b)
{
    float x0 = threadIdx.x, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
    for(int i = 0; i < iters; i++)
    {
#pragma unroll
        for(int k = 0; k < 16; k++)
        {
            x0 = fmaf(x0, a, b), x1 = fmaf(x1, a, b), x2 = fmaf(x2, a, b), x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b), x5 = fmaf(x5, a, b), x6 = fmaf(x6, a, b), x7 = fmaf(x7, a, b);
        }
    }
    float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if(s == 12345.678f)
        out[0] = s;
}
double launchFmaPeak(int blocks, int threads, int iters, cudaStream_t stream)
{
    static float *sink = nullptr;
    if(!sink)
        cudaMalloc(&sink, sizeof(float));
    fmaPeakKernel<<<blocks, threads, 0, stream>>>(sink, iters, 0.999f, 0.001f);
    return (double)blocks * threads * (double)iters * 16.0 * 8.0;
}

} // namespace ptb

namespace ptb
{
// Resident CTAs per SM for a CSG kernel at this shared-memory size (occupancy API).
template <class K>
static int occupancyOf(K kernel, int threads, size_t smemBytes)
{
    int n = 0;
    ensureDynamicSmem(kernel, smemBytes);
    if(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, threads, smemBytes) != cudaSuccess)
    {
        cudaGetLastError();
        return 0;
    }
    return n;
}
int csgBlocksPerSm(int which, int threads, size_t smemBytes, bool globalProgram)
{
    if(which == 1)
        return occupancyOf(intersectExactKernel, threads, smemBytes);
    if(which == 0)
        return globalProgram ? occupancyOf(intersectFastKernel<true, false>, threads, smemBytes) : occupancyOf(intersectFastKernel<false, false>, threads, smemBytes);
    return globalProgram ? occupancyOf(tracePrimaryKernel<true>, threads, smemBytes) : occupancyOf(tracePrimaryKernel<false>, threads, smemBytes);
}
} // namespace ptb
