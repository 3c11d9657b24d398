// ptb200_device.cpp — the parts of the host C++ API that talk to the GPU through the C ABI.
#include "ptb200_render.hpp"

#include <string.h>

namespace ptb200
{

void check(int status)
{
    if(status != 0)
        throw DeviceError(status, ptb200_last_error());
}

DeviceScene::DeviceScene(const PathTrace::Object *world) : scene(NULL)
{
    Flattener f;
    world->flatten(f); // std::domain_error on a singular TransformedObject matrix
    PtbSceneDesc d = f.desc();
    check(ptb200_scene_create(&d, &scene));
}
DeviceScene::~DeviceScene() { ptb200_scene_destroy(scene); }

PtbRenderParams defaultParams(int width, int height, int spp, int depth)
{
    PtbRenderParams p;
    memset(&p, 0, sizeof(p));
    p.width = width, p.height = height;
    p.spp = spp;
    p.sample_offset = 0, p.sample_stride = 1;
    p.max_depth = depth;
    p.screen_width = (float)width, p.screen_height = (float)height;
    p.screen_distance = (float)(std::min(width, height) * 2); // src/test.cpp:450
    p.seed = 0xB200;
    return p;
}

void renderFrame(const DeviceScene &scene, const PtbRenderParams &params, float *rgba, PtbStats *stats) { check(ptb200_render(scene.handle(), &params, rgba, stats)); }
void renderFrame(const PathTrace::Object *world, const PtbRenderParams &params, float *rgba, PtbStats *stats)
{
    DeviceScene scene(world);
    renderFrame(scene, params, rgba, stats);
}

PtbAdaptiveParams adaptiveDefaults(int width, int height)
{
    // getBlockSize(ScreenWidth / 8) and maximumSampleSize = ScreenHeight / (480 / 4), src/test.cpp:40-50
    int count = width / 8, size = 1024;
    while(size > count && size > 4)
        size /= 2;
    PtbAdaptiveParams a;
    a.block_size = std::max(2, size / 4);
    a.max_sample_size = height / (480 / 4);
    a.min_color_delta = 0.003f; // minimumColorDelta, src/test.cpp:38
    a.reserved = 0;
    return a;
}
void renderFrameAdaptive(const DeviceScene &scene, const PtbRenderParams &params, const PtbAdaptiveParams &adaptive, float *rgba, PtbStats *stats, PtbAdaptiveStats *adaptiveStats)
{
    check(ptb200_render_adaptive(scene.handle(), &params, &adaptive, rgba, NULL, NULL, stats, adaptiveStats));
}

MultiDeviceScene::MultiDeviceScene(const PathTrace::Object *world, const std::vector<int> &devices) : comm(NULL)
{
    std::vector<int> devs = devices;
    if(devs.empty())
        for(int i = 0; i < ptb200_device_count(); i++)
            devs.push_back(i);
    Flattener f;
    world->flatten(f);
    PtbSceneDesc d = f.desc();
    check(ptb200_comm_init_all(devs.data(), (int)devs.size(), &comm));
    scenes.assign(devs.size(), (PtbScene *)NULL);
    int rc = ptb200_scene_create_multi(comm, &d, scenes.data());
    if(rc != 0)
    {
        std::string msg = ptb200_last_error();
        ptb200_comm_destroy(comm);
        throw DeviceError(rc, msg);
    }
}
MultiDeviceScene::~MultiDeviceScene()
{
    for(size_t i = 0; i < scenes.size(); i++)
        ptb200_scene_destroy(scenes[i]);
    ptb200_comm_destroy(comm);
}
void MultiDeviceScene::render(const PtbRenderParams &params, float *rgba, PtbStats *statsPerDevice)
{
    check(ptb200_render_multi(comm, scenes.data(), &params, rgba, NULL, statsPerDevice, NULL));
}

GpuBlockRenderer::GpuBlockRenderer(const DeviceScene &scene, int x, int y, int size, const PtbRenderParams &frame) : job(NULL)
{
    PtbRenderParams p = frame;
    x0 = x, y0 = y;
    x1 = std::min(x + size, frame.width), y1 = std::min(y + size, frame.height);
    p.x0 = x0, p.y0 = y0, p.x1 = x1, p.y1 = y1;
    check(ptb200_render_async(scene.handle(), &p, &job));
    staging.assign((size_t)frame.width * frame.height * 4, 0.0f);
}
GpuBlockRenderer::~GpuBlockRenderer() { ptb200_job_destroy(job); }
bool GpuBlockRenderer::done()
{
    int r = ptb200_job_done(job);
    if(r < 0)
        check(r);
    return r == 1;
}
void GpuBlockRenderer::copyToBuffer(PathTrace::Color *screenBuffer, int w, int h)
{
    std::fill(staging.begin(), staging.end(), 0.0f);
    check(ptb200_job_read(job, staging.data(), w, h));
    for(int y = y0; y < y1 && y < h; y++)
        for(int x = x0; x < x1 && x < w; x++)
        {
            const float *p = &staging[4 * ((size_t)y * w + x)];
            if(p[3] > 0)
                screenBuffer[x + (size_t)w * y] = PathTrace::Color(p[0], p[1], p[2]);
        }
}
PtbStats GpuBlockRenderer::finish()
{
    PtbStats s;
    check(ptb200_job_wait(job, &s));
    return s;
}

// GPU-backed SpanIterator: init(ray) runs the CSG kernel for that ray and the iterator walks the
// root span list it returned.  Materials are mapped back to the caller's Material pointers.
class DeviceSpanIterator : public PathTrace::SpanIterator
{
public:
    DeviceScene scene;
    std::vector<const PathTrace::Material *> materials;
    std::vector<PtbSpan> spans;
    PathTrace::Span current;
    int count, index;
    DeviceSpanIterator(const PathTrace::Object *world, const std::vector<const PathTrace::Material *> &materials) : scene(world), materials(materials), spans(16), count(0), index(0) {}
    void load()
    {
        if(index >= count)
            return;
        const PtbSpan &s = spans[index];
        current.start = s.start, current.end = s.end;
        current.startNormal = PathTrace::Vector3D(s.start_normal[0], s.start_normal[1], s.start_normal[2]);
        current.endNormal = PathTrace::Vector3D(s.end_normal[0], s.end_normal[1], s.end_normal[2]);
        current.startMaterial = materials[s.start_material];
        current.endMaterial = materials[s.end_material];
    }
    virtual const PathTrace::Span &operator*() const { return current; }
    virtual const PathTrace::Span *operator->() const { return &current; }
    virtual bool isAtEnd() const { return index >= count; }
    virtual void next()
    {
        index++;
        load();
    }
    virtual void init(const PathTrace::Ray &ray)
    {
        float o[3] = {ray.origin.x, ray.origin.y, ray.origin.z}, d[3] = {ray.dir.x, ray.dir.y, ray.dir.z};
        PtbHit hit;
        for(;;)
        {
            check(ptb200_intersect(scene.handle(), o, d, 1, &hit, spans.data(), (int)spans.size(), PTB_FLAG_EXACT));
            if(hit.n_spans <= (int)spans.size())
                break;
            spans.resize(hit.n_spans);
        }
        count = hit.n_spans, index = 0;
        load();
    }
};

} // namespace ptb200

namespace PathTrace
{

DefaultRandomEngine defaultRandomEngine;

SpanIterator *Object::makeSpanIterator() const
{
    // flatten once just to learn the Material* order (materials are numbered in first-use order)
    ptb200::Flattener f;
    flatten(f);
    std::vector<const void *> keys = f.materialKeys();
    std::vector<const Material *> mats(keys.size());
    for(size_t i = 0; i < keys.size(); i++)
        mats[i] = (const Material *)keys[i];
    return new ptb200::DeviceSpanIterator(this, mats);
}

Color tracePixelSeeded(SpanIterator &spanIterator, int px, int py, int W, int H, int sampleCount, int rayDepth, float sw, float sh, float sd, uint64_t seed)
{
    ptb200::DeviceSpanIterator *it = dynamic_cast<ptb200::DeviceSpanIterator *>(&spanIterator);
    if(!it)
        throw std::invalid_argument("tracePixel needs an iterator from Object::makeSpanIterator()");
    PtbRenderParams p = ptb200::defaultParams(W, H, sampleCount, rayDepth);
    p.screen_width = sw, p.screen_height = sh, p.screen_distance = sd;
    p.x0 = px, p.y0 = py, p.x1 = px + 1, p.y1 = py + 1;
    p.seed = seed;
    if(px < 0 || py < 0 || px >= W || py >= H)
        throw std::invalid_argument("tracePixel: pixel outside the image");
    // a full-frame staging buffer per pixel would be wasteful: render into a 1-pixel-tall strip
    std::vector<float> frame((size_t)W * H * 4, 0.0f);
    ptb200::check(ptb200_render(it->scene.handle(), &p, frame.data(), NULL));
    const float *c = &frame[4 * ((size_t)py * W + px)];
    return Color(c[0], c[1], c[2]);
}

Color tracePixel(SpanIterator &spanIterator, int px, int py, int W, int H, int sampleCount, int rayDepth, float sw, float sh, float sd)
{
    return tracePixel(spanIterator, px, py, W, H, sampleCount, rayDepth, sw, sh, sd, defaultRandomEngine);
}

// Texture::getColor / getFloat: evaluate the flattened texture program at one point on the GPU.
static void evalOne(const Texture *t, Vector3D pos, float rgb[3], float *f)
{
    ptb200::Flattener fl;
    int m = fl.beginMaterial(t, 1.0f);
    for(int slot = 0; slot < 5; slot++)
        fl.beginTexture(m, slot);
    t->flattenTexture(fl);
    fl.addSphere(Vector3D(0, 0, 0), 1.0f, m);
    PtbSceneDesc d = fl.desc();
    PtbScene *scene = NULL;
    ptb200::check(ptb200_scene_create(&d, &scene));
    float p[3] = {pos.x, pos.y, pos.z};
    int rc = ptb200_texture_eval(scene, d.materials[0].tex[0], p, 1, rgb, f, 0);
    ptb200_scene_destroy(scene);
    ptb200::check(rc);
}
Color Texture::getColor(Vector3D pos) const
{
    float rgb[3], f;
    evalOne(this, pos, rgb, &f);
    return Color(rgb[0], rgb[1], rgb[2]);
}
float Texture::getFloat(Vector3D pos) const
{
    float rgb[3], f;
    evalOne(this, pos, rgb, &f);
    return f;
}

} // namespace PathTrace
