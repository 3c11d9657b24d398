// ptb200_render.hpp — the integrator side of the host API: what the reference exposes in
// include/path-trace.h (DefaultRandomEngine, tracePixel) and what its one caller, the
// BlockRenderer in src/test.cpp:310-518, looks like — here backed by the CUDA wavefront
// integrator through the C ABI.  Non-zero C-ABI statuses become C++ exceptions, as the
// reference's API reports errors.
#ifndef PTB200_RENDER_HPP_INCLUDED
#define PTB200_RENDER_HPP_INCLUDED

#include "ptb200_scene.hpp"

namespace ptb200
{

class DeviceError : public std::runtime_error
{
public:
    int status;
    DeviceError(int status, const std::string &what) : std::runtime_error(what), status(status) {}
};
void check(int status); // throws DeviceError(ptb200_last_error()) when status != 0

// An Object graph flattened and resident on the current CUDA device.
class DeviceScene
{
public:
    explicit DeviceScene(const PathTrace::Object *world);
    ~DeviceScene();
    PtbScene *handle() const { return scene; }

private:
    PtbScene *scene;
    DeviceScene(const DeviceScene &);
    DeviceScene &operator=(const DeviceScene &);
};

// Camera of RenderBlock::calcPixelColor (reference src/test.cpp:450): screen = (W, H, 2*min(W,H)).
PtbRenderParams defaultParams(int width, int height, int spp, int depth = 16);

// Renders the whole frame (or params.x0..y1) synchronously; rgba = width*height*4 floats.
void renderFrame(const PathTrace::Object *world, const PtbRenderParams &params, float *rgba, PtbStats *stats = NULL);
void renderFrame(const DeviceScene &scene, const PtbRenderParams &params, float *rgba, PtbStats *stats = NULL);

// The reference's adaptive tile renderer (RenderBlock::renderSquare, src/test.cpp:423-499) for the
// whole frame: pixels inside squares whose corners agree are interpolated instead of traced.
// adaptiveDefaults() gives the reference's constants for a frame size (src/test.cpp:38-50).
PtbAdaptiveParams adaptiveDefaults(int width, int height);
void renderFrameAdaptive(const DeviceScene &scene, const PtbRenderParams &params, const PtbAdaptiveParams &adaptive, float *rgba, PtbStats *stats = NULL,
                         PtbAdaptiveStats *adaptiveStats = NULL);

// One process driving several GPUs (ptb200_comm_init_all): the scene replicated on every device (image
// pixels cross PCIe once, NCCL broadcast to the others), frames rendered sample-sharded with one NCCL
// reduce-scatter, every device delivering its own rows of `rgba`.  The counterpart of the reference's render
// farm client loop (src/test.cpp:520-630), without the sockets.
class MultiDeviceScene
{
public:
    MultiDeviceScene(const PathTrace::Object *world, const std::vector<int> &devices); // devices empty: all of them
    ~MultiDeviceScene();
    int deviceCount() const { return (int)scenes.size(); }
    void render(const PtbRenderParams &params, float *rgba, PtbStats *statsPerDevice = NULL);

private:
    PtbComm *comm;
    std::vector<PtbScene *> scenes;
    MultiDeviceScene(const MultiDeviceScene &);
    MultiDeviceScene &operator=(const MultiDeviceScene &);
};

// Drop-in for the reference's BlockRenderer (src/test.cpp:310-323): construction starts the
// work asynchronously, done() polls, copyToBuffer() may be called repeatedly while it runs and
// writes only pixels that already have a value.
class GpuBlockRenderer
{
public:
    GpuBlockRenderer(const DeviceScene &scene, int x, int y, int size, const PtbRenderParams &frame);
    ~GpuBlockRenderer();
    bool done();
    void copyToBuffer(PathTrace::Color *screenBuffer, int w, int h);
    PtbStats finish();

private:
    PtbJob *job;
    int x0, y0, x1, y1;
    std::vector<float> staging;
    GpuBlockRenderer(const GpuBlockRenderer &);
    GpuBlockRenderer &operator=(const GpuBlockRenderer &);
};

} // namespace ptb200

namespace PathTrace
{

// reference include/path-trace.h:21-54.  Kept for source compatibility; the GPU path draws from
// a counter-based per-(pixel,sample) generator and uses this engine only to derive its seed.
class DefaultRandomEngine
{
    uint64_t v;

public:
    DefaultRandomEngine() { seed(0); }
    static unsigned min() { return 0; }
    static unsigned max() { return 0xFFFFFFFF; }
    void seed(unsigned s) { v = s ^ 0x12476242; }
    unsigned operator()()
    {
        v = 214013 * v + 2531011;
        return (unsigned)(v >> 32);
    }
    void discard(unsigned long long n)
    {
        while(n--)
            operator()();
    }
};
extern DefaultRandomEngine defaultRandomEngine;
const int DefaultRayDepth = 16;
const int DefaultSampleCount = 200;
const float DefaultScreenWidth = 4.0 / 3.0;
const float DefaultScreenHeight = 1.0;
const float DefaultScreenDistance = 2.0;

// tracePixel (reference include/path-trace.h:187-207) for an iterator obtained from
// Object::makeSpanIterator(): the mean of sampleCount jittered camera samples of pixel
// (px, py), computed on the GPU.  One launch per call — use renderFrame / GpuBlockRenderer for
// throughput.  Throws std::invalid_argument for iterators that did not come from this library.
Color tracePixel(SpanIterator &spanIterator, int px, int py, int screenXResolution, int screenYResolution, int sampleCount = DefaultSampleCount,
                 int rayDepth = DefaultRayDepth, float screenWidth = DefaultScreenWidth, float screenHeight = DefaultScreenHeight,
                 float screenDistance = DefaultScreenDistance);
template <typename T>
inline Color tracePixel(SpanIterator &spanIterator, int px, int py, int screenXResolution, int screenYResolution, int sampleCount, int rayDepth, float screenWidth,
                        float screenHeight, float screenDistance, T &randomEngine)
{
    extern Color tracePixelSeeded(SpanIterator &, int, int, int, int, int, int, float, float, float, uint64_t);
    uint64_t seed = ((uint64_t)randomEngine() << 32) | randomEngine();
    return tracePixelSeeded(spanIterator, px, py, screenXResolution, screenYResolution, sampleCount, rayDepth, screenWidth, screenHeight, screenDistance, seed);
}
Color tracePixelSeeded(SpanIterator &spanIterator, int px, int py, int W, int H, int sampleCount, int rayDepth, float sw, float sh, float sd, uint64_t seed);

} // namespace PathTrace

#endif // PTB200_RENDER_HPP_INCLUDED
