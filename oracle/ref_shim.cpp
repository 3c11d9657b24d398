// ref_shim.cpp — TEST INFRASTRUCTURE.  Drives the UNMODIFIED reference implementation
// (/root/reference/include/*.h + src/*.cpp, compiled where they lie by oracle/Makefile into
// oracle/_ref/libptref.so) through a flat C API so the tests and bench.py's cpu_baseline /
// `--impl reference` leg can call it with ctypes.  Nothing in the product path may load this.
//
// Everything that computes here is the reference's own code: Object::makeSpanIterator,
// SpanIterator::init/next, traceRay, tracePixel, Texture::getColor/getFloat, Image, writeHDR.
// This file only marshals arrays in and out.
#include "path-trace.h"
#include "image_texture.h"
#include "filter_texture.h"
#include "transform_texture.h"
#include "png_decoder.h"
#include "thread.h"

#include <pthread.h>
#include <time.h>
#include <string.h>
#include <unistd.h>

#include "ptb200.h" // PtbSpan / PtbHit record layouts shared with the CUDA path

// libpng 1.2 is not in this image: PNG decoding is stubbed out.
PngDecoder::PngDecoder(string fileName) : w(0), h(0), data(NULL) { throw PngLoadError("png decoding unavailable in oracle build: " + fileName); }

#define SC(name) ref_##name
#include "../path-trace_b200/host/scene_capi.inc"

using namespace PathTrace;

namespace
{
// Pass-through iterator that counts root init() calls = ray segments.
class CountingIterator : public SpanIterator
{
public:
    SpanIterator *inner;
    unsigned long long rays;
    CountingIterator(SpanIterator *inner) : inner(inner), rays(0) {}
    virtual ~CountingIterator() { delete inner; }
    virtual const Span &operator*() const { return **inner; }
    virtual const Span *operator->() const { return &**inner; }
    virtual bool isAtEnd() const { return inner->isAtEnd(); }
    virtual void next() { inner->next(); }
    virtual void init(const Ray &ray)
    {
        rays++;
        inner->init(ray);
    }
};

double nowSeconds()
{
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}
} // namespace

extern "C" {

// Root span list of every ray, in iterator order.  spans: n x max_spans records (prim ids are
// unknown to the reference -> -1; materials by creation index).  counts: n.
int ref_trace_spans(void *world, const float *origins, const float *dirs, size_t n, PtbSpan *spans, int max_spans, int *counts)
{
    SCAPI_TRY
    SpanIterator *it = ((Object *)world)->makeSpanIterator();
    for(size_t i = 0; i < n; i++)
    {
        Ray ray(Vector3D(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]), Vector3D(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
        it->init(ray);
        int c = 0;
        for(; *it; (*it)++)
        {
            if(c < max_spans)
            {
                PtbSpan &o = spans[i * (size_t)max_spans + c];
                const Span &s = **it;
                o.start = s.start, o.end = s.end;
                o.start_normal[0] = s.startNormal.x, o.start_normal[1] = s.startNormal.y, o.start_normal[2] = s.startNormal.z;
                o.end_normal[0] = s.endNormal.x, o.end_normal[1] = s.endNormal.y, o.end_normal[2] = s.endNormal.z;
                o.start_prim = o.end_prim = -1;
                o.start_material = scapi::materialIndex(s.startMaterial);
                o.end_material = scapi::materialIndex(s.endMaterial);
            }
            c++;
        }
        counts[i] = c;
    }
    delete it;
    return 0;
    SCAPI_CATCH(-1)
}

// One traceRay call per ray (reference include/path-trace.h:58-165), engine seeded per call.
int ref_trace_rays(void *world, const float *origins, const float *dirs, size_t n, int depth, unsigned seed, float strength, float *rgb)
{
    SCAPI_TRY
    SpanIterator *it = ((Object *)world)->makeSpanIterator();
    DefaultRandomEngine rng;
    rng.seed(seed);
    for(size_t i = 0; i < n; i++)
    {
        Ray ray(Vector3D(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]), Vector3D(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
        Color c = traceRay(ray, *it, depth, rng, strength);
        rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
    }
    delete it;
    return 0;
    SCAPI_CATCH(-1)
}

// Single-threaded seeded render of the pixel lattice x = x0 + i*step (i < nx), y = y0 + j*step
// through the reference's jittering tracePixel (include/path-trace.h:187-201), row-major.
int ref_render(void *world, int W, int H, int x0, int y0, int step, int nx, int ny, int sampleCount, int depth, float sw, float sh, float sd, unsigned seed,
               float *rgb, unsigned long long *rays)
{
    SCAPI_TRY
    CountingIterator it(((Object *)world)->makeSpanIterator());
    DefaultRandomEngine rng;
    rng.seed(seed);
    for(int j = 0; j < ny; j++)
        for(int i = 0; i < nx; i++)
        {
            Color c = tracePixel(it, x0 + i * step, y0 + j * step, W, H, sampleCount, depth, sw, sh, sd, rng);
            float *o = rgb + 3 * ((size_t)j * nx + i);
            o[0] = c.x, o[1] = c.y, o[2] = c.z;
        }
    if(rays)
        *rays = it.rays;
    return 0;
    SCAPI_CATCH(-1)
}
} // extern "C"

// ---- multithreaded baseline driver: pthread workers, one makeSpanIterator()
//      per worker as RenderBlock::run (reference src/test.cpp:501-507), private engine each ----
namespace
{
struct MtJob
{
    Object *world;
    int W, H, x0, y0, step, nx, ny, sampleCount, depth;
    float sw, sh, sd;
    unsigned seed;
    float *rgb;
    double deadline; // wall-clock budget: workers stop taking rows after it
    pthread_mutex_t lock;
    int nextRow;
    unsigned long long rays, pixels;
};
void *mtWorker(void *arg)
{
    MtJob &job = *(MtJob *)arg;
    CountingIterator it(job.world->makeSpanIterator());
    unsigned long long pixels = 0;
    for(;;)
    {
        pthread_mutex_lock(&job.lock);
        int row = job.nextRow < job.ny && nowSeconds() < job.deadline ? job.nextRow++ : -1;
        pthread_mutex_unlock(&job.lock);
        if(row < 0)
            break;
        DefaultRandomEngine rng;
        rng.seed(job.seed + 7919u * (unsigned)row);
        for(int i = 0; i < job.nx; i++)
        {
            Color c = tracePixel(it, job.x0 + i * job.step, job.y0 + row * job.step, job.W, job.H, job.sampleCount, job.depth, job.sw, job.sh, job.sd, rng);
            float *o = job.rgb + 3 * ((size_t)row * job.nx + i);
            o[0] = c.x, o[1] = c.y, o[2] = c.z;
            pixels++;
        }
    }
    pthread_mutex_lock(&job.lock);
    job.rays += it.rays;
    job.pixels += pixels;
    pthread_mutex_unlock(&job.lock);
    return NULL;
}
} // namespace

extern "C" {
int ref_hardware_concurrency(void) { return (int)sysconf(_SC_NPROCESSORS_ONLN); }

// Renders lattice rows on `threads` workers until done or `max_seconds` elapsed.  Returns rows
// completed (rows are handed out in order, so rows [0, return) are valid once all workers exit).
int ref_render_mt(void *world, int W, int H, int x0, int y0, int step, int nx, int ny, int sampleCount, int depth, float sw, float sh, float sd, unsigned seed,
                  int threads, double max_seconds, float *rgb, unsigned long long *rays, unsigned long long *pixels, double *seconds)
{
    SCAPI_TRY
    MtJob job;
    job.world = (Object *)world;
    job.W = W, job.H = H, job.x0 = x0, job.y0 = y0, job.step = step, job.nx = nx, job.ny = ny;
    job.sampleCount = sampleCount, job.depth = depth, job.sw = sw, job.sh = sh, job.sd = sd, job.seed = seed, job.rgb = rgb;
    job.nextRow = 0, job.rays = 0, job.pixels = 0;
    pthread_mutex_init(&job.lock, NULL);
    double t0 = nowSeconds();
    job.deadline = max_seconds > 0 ? t0 + max_seconds : 1e300;
    std::vector<pthread_t> th(threads);
    for(int i = 0; i < threads; i++)
        pthread_create(&th[i], NULL, mtWorker, &job);
    for(int i = 0; i < threads; i++)
        pthread_join(th[i], NULL);
    double t1 = nowSeconds();
    pthread_mutex_destroy(&job.lock);
    if(rays)
        *rays = job.rays;
    if(pixels)
        *pixels = job.pixels;
    if(seconds)
        *seconds = t1 - t0;
    return job.nextRow;
    SCAPI_CATCH(-1)
}

// Texture::getColor / getFloat at n positions.
int ref_tex_eval(void *tex, const float *pos, size_t n, float *rgb, float *f)
{
    SCAPI_TRY
    Texture *t = (Texture *)tex;
    for(size_t i = 0; i < n; i++)
    {
        Vector3D p(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
        if(rgb)
        {
            Color c = t->getColor(p);
            rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
        }
        if(f)
            f[i] = t->getFloat(p);
    }
    return 0;
    SCAPI_CATCH(-1)
}

// DefaultRandomEngine / uniform_real_distribution / Vector3D::rand known-answer streams.
void ref_rng_u32(unsigned seed, size_t n, unsigned *out)
{
    DefaultRandomEngine rng;
    rng.seed(seed);
    for(size_t i = 0; i < n; i++)
        out[i] = rng();
}
void ref_rng_uniform(unsigned seed, size_t n, float lo, float hi, float *out)
{
    DefaultRandomEngine rng;
    rng.seed(seed);
    uniform_real_distribution<float> d(lo, hi);
    for(size_t i = 0; i < n; i++)
        out[i] = d(rng);
}
void ref_rng_ball(unsigned seed, size_t n, float max, float min, float *out)
{
    DefaultRandomEngine rng;
    rng.seed(seed);
    for(size_t i = 0; i < n; i++)
    {
        Vector3D v = Vector3D::rand(rng, max, min);
        out[3 * i] = v.x, out[3 * i + 1] = v.y, out[3 * i + 2] = v.z;
    }
}
} // extern "C"
