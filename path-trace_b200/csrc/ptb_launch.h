// ptb_launch.h — internal interface between the C-ABI host code (ptb_capi.cu) and the kernels
// (ptb_kernels.cu).  Not part of the public boundary.
#pragma once
#ifndef __CUDACC_RTC__
#include <cuda_runtime.h>
#include <stdint.h>
#endif

#include "ptb_scene_dev.cuh"

namespace ptb
{

// How one kernel launch lays out its shared memory and span-pool overflow.
struct CsgConfig
{
    int threads;       // threads per block
    int blocks;        // grid size (persistent: a multiple of the SM count)
    int cap;           // span-pool slots per thread kept in shared memory
    int stageProgram;  // 1: copy the program stream into shared memory first
    int fast;          // 1: laid out for the fast evaluator (ptb_csg_fast.cuh), 128 threads
    size_t smemBytes;  // dynamic shared memory per block
    float *overflow;   // global spill area or NULL when cap >= scene.poolBound
    size_t ovfStride;  // = blocks * threads
};

// Ray queue: four float4 SoA streams (64 B per ray), see ptb_kernels.cu.
struct RayQueue
{
    float4 *q0, *q1, *q2, *q3;
};

// Division of a 32-bit index by a launch constant as multiply-high + shifts (three integer divisions
// per camera sample were 60 of the 175 instructions of cameraSample).  Round-up method with a 33-bit
// magic number: q = mulhi(m, n); q = (((n - q) >> 1) + q) >> (l - 1), l = ceil(log2 d), d >= 2.
struct FastDivisor
{
    uint32_t d, m, shift;
#ifndef __CUDACC_RTC__
    static FastDivisor make(uint32_t d)
    {
        FastDivisor f;
        f.d = d, f.m = 0u, f.shift = 0u;
        if(d >= 2u)
        {
            uint32_t l = 0;
            while((1ull << l) < d)
                l++;
            f.m = (uint32_t)(((1ull << 32) * ((1ull << l) - d)) / d + 1ull);
            f.shift = l - 1u;
        }
        return f;
    }
#endif
#ifdef __CUDACC__
    __device__ __forceinline__ uint32_t div(uint32_t n) const
    {
        if(d < 2u)
            return n;
        uint32_t q = __umulhi(m, n);
        return (((n - q) >> 1) + q) >> shift;
    }
#endif
};

struct RenderPass
{
    int width, height;          // full image
    int x0, y0, rw, rh;         // rectangle rendered
    int samplesThisPass;        // camera samples per pixel in this pass
    int sampleBase;             // global index of this pass's first sample
    int sampleStride;
    int maxDepth;
    float sw, sh, sd;
    float twoOverWidth, twoOverHeight; // 2 / width, 2 / height
    FastDivisor byPixels, byRowWidth;  // rw * rh and rw
    FastDivisor bySamples;             // samplesThisPass
    uint64_t seed;
    uint32_t flags;
    // adaptive rendering: when non-null, sample index p of a pass names lattice point pixelList[p]
    // (y << 16 | x) instead of a pixel of the rectangle; rw = list length, rh = 1, width = frame pitch
    const uint32_t *pixelList;
};

#ifndef __CUDACC_RTC__ /* host-side launchers: not seen by the run-time compiled device code */
void launchIntersect(const DevScene &sc, const CsgConfig &cfg, const float4 *origins, const float4 *dirs, size_t n, PtbHit *hits, PtbSpan *spans, int maxSpans,
                     bool exact, bool origin0, cudaStream_t stream);
// Trace stage of the first wavefront level: camera rays for rw*rh*samplesThisPass samples -> hits[i].
void launchTracePrimary(const DevScene &sc, const CsgConfig &cfg, const RenderPass &pass, uint2 *hits, cudaStream_t stream);
// Trace stage of a bounce level: *inCount rays of `in` -> hits[i].
void launchTraceQueue(const DevScene &sc, const CsgConfig &cfg, RayQueue in, const unsigned *inCount, uint2 *hits, cudaStream_t stream);
// Shade stage: hits -> accumulate + survivors appended to `out` (*outCount counts them).  primary:
// path state regenerated from the sample index; else read from `in` (*inCount entries).
void launchShade(const DevScene &sc, int blocks, bool primary, const RenderPass &pass, float4 *accum, const uint2 *hits, RayQueue in, const unsigned *inCount, RayQueue out,
                 unsigned *outCount, cudaStream_t stream);
int shadeBlocksPerSm(bool primary);
// Tail of a pass: every ray of `in` (*inCount of them, bounce level >= 2) is followed to the end of its path by
// ONE kernel — trace, shade, trace, ... in a loop, no queue in between.  *extraRays += rays traced beyond the first
// of each path.  Uses the trace stage's launch shape (cfg: shared-memory layout of the interpreting evaluator).
void launchTail(const DevScene &sc, const CsgConfig &cfg, int blocks, const RenderPass &pass, float4 *accum, RayQueue in, const unsigned *inCount, unsigned *extraRays,
                cudaStream_t stream);
int tailBlocksPerSm(bool globalProgram, size_t smemBytes);
void launchResolve(const float4 *accum, float4 *rgba, size_t nPixels, cudaStream_t stream);
void launchDisplay8(const float4 *rgba, uchar4 *out, size_t nPixels, bool divideByAlpha, cudaStream_t stream);

// Adaptive quadtree of RenderBlock::renderSquare (reference src/test.cpp:423-499), level-synchronous.
struct AdaptiveLevel
{
    int width, height;       // image
    int pitch;               // frame pitch (lattice points per row, > width)
    int size;                // edge of the squares of this level
    int maxSampleSize;       // maximumSampleSize (src/test.cpp:50)
    float delta2;            // minimumColorDelta * minimumColorDelta (float product, src/test.cpp:439)
    const uint32_t *squares; // y << 16 | x of every square of this level
    unsigned nSquares;
    const float4 *value;     // traced lattice points: mean radiance, alpha 1
    float4 *interp;          // interpolated pixels: colour, alpha 2
    unsigned *requested;     // per lattice point: 1 once it has been put on a pixel list
    uint32_t *childSquares;  // out: squares of the next level (size / 2)
    uint32_t *pixelList;     // out: lattice points to trace before the next level
    unsigned *counters;      // [0] = child squares, [1] = pixel list length, [2] = squares interpolated, [3] = pixels interpolated
};
void launchAdaptiveDecide(const AdaptiveLevel &lv, cudaStream_t stream);
// Sorts a pixel list by (y, x): the decide kernel appends in atomic order, and camera rays of scattered
// pixels are as incoherent as bounce rays (2x slower to trace).  tempBytes = 0: returns the scratch size.
size_t adaptiveSortList(uint32_t *list, uint32_t *alt, unsigned n, void *temp, size_t tempBytes, cudaStream_t stream);
// value[p] = (accum[p].rgb / accum[p].a, 1) for every listed lattice point
void launchAdaptiveResolveList(const float4 *accum, float4 *value, const uint32_t *pixelList, unsigned n, int pitch, cudaStream_t stream);
// rgba (width x height) = interpolated pixel where there is one, else the traced one; source: 1 traced, 2 interpolated
void launchAdaptiveAssemble(const float4 *value, const float4 *interp, int pitch, int width, int height, float4 *rgba, uint8_t *source, cudaStream_t stream);
void launchGatherPixels(const float4 *accum, const uint32_t *pixelList, unsigned n, int pitch, float4 *out, cudaStream_t stream);
void launchTextureEval(const DevScene &sc, int root, const float *pos, size_t n, float *rgb, float *f, uint32_t flags, cudaStream_t stream);
void launchScatterSample(const float *normal, const float *refl, float scatter, size_t n, uint64_t seed, uint32_t flags, float *dirs, cudaStream_t stream);
void launchVectorEval(int op, const float *v, const float *nrm, const float *ior, size_t n, float *out, cudaStream_t stream);
void launchRngWords(int which, uint64_t seed, const uint32_t *pixel, const uint32_t *sample, const uint32_t *block, size_t n, uint32_t *out, cudaStream_t stream);
// FMA-chain microbenchmark: returns lane-FMAs executed.
double launchFmaPeak(int blocks, int threads, int iters, cudaStream_t stream);
size_t csgSharedBytes(const DevScene &sc, int threads, int cap, bool stageProgram);
size_t fastSharedBytes(const DevScene &sc, int cap, bool stageProgram, int rays);
int traceRaysPerThread(); // rays each thread of the trace kernels carries
#endif

} // namespace ptb

namespace ptb
{
// which: 0 = intersect (fast), 1 = intersect (exact), 2 = render stages; globalProgram: the
// program stream is too large for shared memory and is read through the read-only cache
#ifndef __CUDACC_RTC__
int csgBlocksPerSm(int which, int threads, size_t smemBytes, bool globalProgram);
#endif
}
