// ptb_internal.h — what the translation units behind the C ABI share (ptb_capi.cu: scenes, ray batches,
// the wavefront driver; ptb_multi.cu: the multi-GPU entry points).  Not part of the public boundary.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "ptb_launch.h"
#include "ptb_specialize.h"

namespace ptb
{

// thread-local message behind ptb200_last_error(); returns `code`
int fail(int code, const std::string &msg);
const std::string &lastError();
void setLastError(const std::string &msg);
#define CU(call)                                                                                           \
    do                                                                                                     \
    {                                                                                                      \
        cudaError_t e_ = (call);                                                                           \
        if(e_ != cudaSuccess)                                                                              \
            return ptb::fail(PTB_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_));               \
    } while(0)

// Device block that goes to a per-process cache instead of back to the driver (cudaMalloc / cudaFree of
// the multi-GB ray queues cost from milliseconds to seconds).  A block may only be released once no
// kernel uses it any more: PtbScene::quiesce() is called before a scene's scratch shrinks, grows or dies.
struct DeviceBuffer
{
    void *p = nullptr;
    size_t bytes = 0;
    int device = 0;
    ~DeviceBuffer() { release(); }
    void release();
    cudaError_t ensure(size_t n);
};
void cacheTrim(int device); // device < 0: every device

struct ImageSlot
{
    cudaArray_t array = nullptr;
    cudaTextureObject_t point = 0, linear = 0;
    int w = 0, h = 0;
};

// Per-scene scratch for rendering (queues, counters, events).
struct RenderScratch
{
    DeviceBuffer queues, hits, counters, overflow, accum;
    DeviceBuffer adaptive; // value + interp frames, request flags, square and pixel lists of ptb200_render_adaptive
    DeviceBuffer multi;    // ptb200_render_multi: the padded accumulation frame the reduce-scatter runs on
    unsigned *hostCounters = nullptr; // pinned
    size_t hostCountersLen = 0;
    std::vector<cudaEvent_t> events;
    ~RenderScratch();
};

} // namespace ptb

// One scene on one device.  The scene DATA is immutable after create and may be shared by threads; the
// per-scene SCRATCH (ray queues, hit records, counters, span-pool overflow) is one set, so calls on a scene
// are serialised: `lock` serialises submission on the host, and `lastUse` (an event recorded behind the last
// kernel of every call) orders the device side — a call on another stream first waits for it, and destroy /
// scratch reallocation synchronise on it.
struct PtbScene
{
    int device = 0;
    int smCount = 148;
    size_t smemOptin = 0;
    ptb::DevScene dev;
    std::vector<void *> allocs;
    std::vector<ptb::ImageSlot> images;
    ptb::DevImage *dImages = nullptr;
    int imageCapacity = 0;
    ptb::DeviceBuffer tables; // every scene table (programs, primitives, transforms, materials, textures, image table) in ONE block, taken from and returned to the library's block cache
    std::mutex lock; // one render / intersect call at a time per scene
    ptb::RenderScratch scratch;
    bool asyncBusy = false;
    cudaEvent_t lastUse = nullptr;      // recorded on the stream of the last call that launched kernels on this scene
    cudaStream_t lastUseStream = nullptr;
    bool lastUsePending = false;
    // scene-specialised trace kernels (ptb_specialize.cu): compiled on first use by a render or ray batch
    // large enough to pay for the compile, or at once with PTB_FLAG_SPECIALIZE
    std::vector<uint4> fastHost;
    ptb::FastTree fastTree;
    std::string specSource[2]; // SPEC_TRACE, SPEC_INTERSECT
    bool specNever[2] = {false, false}; // a large program whose generated code has no bounding-sphere tests: stays interpreted
    bool lastSpecCameraOnly = false; // the last render ran the specialised kernel for camera rays only (a large program)
    int lastSpecPasses = 0, lastPasses = 0; // of the last render call with stats
    int specSeen[2] = {0, 0};               // calls that wanted the specialised kernels of each kind so far
    ptb::DeviceBuffer specOverflow;

    // Device-side ordering of the shared scratch (see above).  begin(): called with `lock` held before the
    // first launch of a call on `stream`; end(): behind its last launch; quiesce(): host waits for the last call.
    int begin(cudaStream_t stream);
    int end(cudaStream_t stream);
    void quiesce();
};

struct PtbJob
{
    PtbScene *scene = nullptr;
    PtbRenderParams params;
    std::thread worker;
    volatile int done = 0;
    int status = 0;
    std::string error;
    PtbStats stats;
    cudaStream_t readStream = nullptr;
    float4 *accum = nullptr; // device, owned by the job
    std::vector<float> staging;
};

namespace ptb
{

// Adaptive rendering traces explicit lists of lattice points instead of a rectangle.
struct PixelListArg
{
    const uint32_t *list; // device: y << 16 | x
    unsigned count;
    int pitch;            // lattice points per row of the accumulation frame
};

// Where the pixels of scene image i come from at create time: the host pointer of the description, or (multi-GPU
// create: uploaded once by the root and broadcast over NVLink) a linear RGBA32F device buffer.
struct ImageSource
{
    const float4 *device; // NULL: take PtbImageDesc::rgba
};
int sceneCreateImpl(const PtbSceneDesc *d, PtbScene *s, std::vector<uint4> *fastOnly = nullptr, const std::vector<ImageSource> *sources = nullptr, FastTree *treeOnly = nullptr);
int checkParams(const PtbRenderParams *p, int &x0, int &y0, int &rw, int &rh);
// The wavefront driver.  Caller holds scene->lock and has made scene->device current.
int renderPasses(PtbScene *s, const PtbRenderParams *p, float4 *accum, PtbStats *stats, cudaStream_t stream, const PixelListArg *points = nullptr);
// Host -> linear device memory through the pinned staging halves (pageable source), on the default stream; synchronous.
bool hostIsPinned(const void *p);
int uploadLinear(void *dst, const void *src, size_t bytes);

} // namespace ptb
