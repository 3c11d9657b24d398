// ptb_noalloc.cu — ptb200_intersect_device_static: the ray-batch CSG kernels with their scratch (the span-pool
// OVERFLOW area: span lists of rays that outgrow the shared-memory pool spill there) in a file-scope device array
// instead of a cudaMalloc'ed block, for callers that cannot allocate: code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc.  No counterpart in the reference.
// The entry point calls no cudaMalloc / cudaFree / allocating library routine, compiles nothing (the
// interpreting kernels only) and leaves every existing code path as it is.
//
// The array is per GPU (every device that loads the library holds PTB_STATIC_SCRATCH_BYTES until exit) and shared
// by every call: calls that overlap on the GPU (other streams, other scenes, other threads) race on it.  Nothing in
// it has to start at zero — a ray's spans are written to a slot before that slot is read (stack discipline of the
// evaluators, csrc/ptb_csg.cuh) — so the contents left by the previous call are harmless and it is never cleared.
#include <mutex>
#include <string>

#include "ptb_internal.h"
#include "ptb_csg_fast.cuh"

using namespace ptb;

// 16 MiB: enough for scenes whose worst-case span lists exceed the shared pool by up to 73 slots per ray at one
// CTA per SM on a 148-SM B200 (12 B per slot x 128 threads x 148 CTAs x 73).
static __device__ __align__(256) unsigned char ptbStaticScratch[PTB_STATIC_SCRATCH_BYTES];

namespace
{
// In host code `ptbStaticScratch` is a host-side stand-in, not a device address: the address a kernel can use
// comes from cudaGetSymbolAddress, once per device, with that device current.
const int kMaxDevices = 64;
void *g_scratchPtr[kMaxDevices];
bool g_scratchKnown[kMaxDevices];
std::mutex g_scratchLock;

int staticScratch(int device, float **out)
{
    if(device < 0 || device >= kMaxDevices)
        return fail(PTB_EINVAL, "device index beyond the static scratch table");
    std::lock_guard<std::mutex> g(g_scratchLock);
    if(!g_scratchKnown[device])
    {
        void *p = nullptr;
        CU(cudaGetSymbolAddress(&p, ptbStaticScratch)); // (the caller made `device` current)
        cudaPointerAttributes attr;
        CU(cudaPointerGetAttributes(&attr, p));
        if(attr.type != cudaMemoryTypeDevice || attr.device != device)
            return fail(PTB_ECUDA, "cudaGetSymbolAddress did not return device memory of the current device");
        g_scratchPtr[device] = p, g_scratchKnown[device] = true;
    }
    *out = (float *)g_scratchPtr[device];
    return PTB_OK;
}
} // namespace

extern "C" size_t ptb200_static_scratch_bytes(void) { return (size_t)PTB_STATIC_SCRATCH_BYTES; }

extern "C" int ptb200_intersect_device_static(PtbScene *s, const void *d_origins4, const void *d_dirs4, size_t n, void *d_hits, void *d_spans, int max_spans, uint32_t flags,
                                              void *stream)
{
    if(!s || !d_origins4 || !d_dirs4 || !d_hits)
        return fail(PTB_EINVAL, "null argument");
    if(d_spans && max_spans <= 0)
        return fail(PTB_EINVAL, "max_spans must be positive when spans are requested");
    if(flags & PTB_FLAG_SPECIALIZE)
        return fail(PTB_EINVAL, "ptb200_intersect_device_static runs the interpreting kernels only (a run-time compile allocates)");
    CU(cudaSetDevice(s->device));
    if(n == 0)
        return PTB_OK;
    const DevScene &dev = s->dev;
    const bool exact = (flags & PTB_FLAG_EXACT) != 0;
    // launch shape as chooseConfig (ptb_capi.cu) makes it, minus the allocation
    CsgConfig cfg;
    cfg.threads = 128;
    cfg.fast = exact ? 0 : 1;
    cfg.cap = dev.poolBound < (exact ? 16 : 12) ? dev.poolBound : (exact ? 16 : 12);
    if(cfg.fast)
    {
        cfg.stageProgram = (size_t)dev.fastWords * 16 <= 96 * 1024 ? 1 : 0;
        cfg.smemBytes = fastSharedBytes(dev, cfg.cap, cfg.stageProgram, 1);
    }
    else
    {
        cfg.stageProgram = (size_t)dev.programWords * 16 <= 64 * 1024 ? 1 : 0;
        cfg.smemBytes = csgSharedBytes(dev, cfg.threads, cfg.cap, cfg.stageProgram);
    }
    if(cfg.smemBytes > s->smemOptin)
        return fail(PTB_EINVAL, "scene needs more shared memory per CTA than the device offers");
    const int perSm = csgBlocksPerSm(exact ? 1 : 0, cfg.threads, cfg.smemBytes, !cfg.stageProgram);
    if(perSm <= 0)
        return fail(PTB_ECUDA, "kernel does not fit on an SM with this shared-memory size");
    long long blocks = (long long)perSm * s->smCount;
    cfg.overflow = nullptr;
    if(dev.poolBound > cfg.cap)
    { // the scratch: (poolBound - cap) slots x {t0, t1, tags} per resident thread.  The grid shrinks until it fits,
      // but not below one CTA per SM: beyond that the scene is too large for this entry point.
        const size_t perBlock = (size_t)(dev.poolBound - cfg.cap) * 3 * sizeof(float) * cfg.threads;
        const long long fit = (long long)((size_t)PTB_STATIC_SCRATCH_BYTES / perBlock);
        if(fit < s->smCount)
            return fail(PTB_EINVAL, "scene too large for ptb200_intersect_device_static: its worst-case span lists need " + std::to_string(perBlock * (size_t)s->smCount) +
                                        " bytes of span-pool overflow at one CTA per SM, the static scratch holds " + std::to_string((size_t)PTB_STATIC_SCRATCH_BYTES) +
                                        " (PTB_STATIC_SCRATCH_BYTES); use ptb200_intersect_device");
        if(blocks > fit)
            blocks = fit;
        float *scratch = nullptr;
        if(int rc = staticScratch(s->device, &scratch))
            return rc;
        cfg.overflow = scratch;
    }
    cfg.blocks = (int)blocks;
    cfg.ovfStride = (size_t)cfg.blocks * cfg.threads;
    launchIntersect(dev, cfg, (const float4 *)d_origins4, (const float4 *)d_dirs4, n, (PtbHit *)d_hits, (PtbSpan *)d_spans, max_spans, exact, (flags & PTB_FLAG_ORIGIN0) != 0,
                    (cudaStream_t)stream);
    CU(cudaGetLastError());
    return PTB_OK;
}
