// ptb200_host_capi.cpp — flat C wrappers (ptbh_*) over this repository's C++ scene API, from the
// same source as the reference-side wrappers (scene_capi.inc), plus flatten / render entry
// points so Python tests and bench.py can drive the C++ host layer with ctypes.
#include "ptb200_scene.hpp"
#include "ptb200_render.hpp"

#define SC(name) ptbh_##name
#include "scene_capi.inc"

extern "C" {

// Flattens an Object graph.  The returned handle owns the arrays the PtbSceneDesc points into.
void *ptbh_flatten(void *world)
{
    SCAPI_TRY
    ptb200::Flattener *f = new ptb200::Flattener;
    try
    {
        ((PathTrace::Object *)world)->flatten(*f);
    }
    catch(...)
    {
        delete f;
        throw;
    }
    return f;
    SCAPI_CATCH(NULL)
}
// Flattens a lone texture as material 0 / slot 0 of an otherwise empty description (for
// ptb200_texture_eval known-answer tests).  Root = desc.materials[0].tex[0].
void *ptbh_flatten_texture(void *tex)
{
    SCAPI_TRY
    ptb200::Flattener *f = new ptb200::Flattener;
    try
    {
        int m = f->beginMaterial(tex, 1.0f);
        for(int slot = 0; slot < 5; slot++)
            f->beginTexture(m, slot);
        ((PathTrace::Texture *)tex)->flattenTexture(*f);
        f->addSphere(PathTrace::Vector3D(0, 0, 0), 1.0f, m);
    }
    catch(...)
    {
        delete f;
        throw;
    }
    return f;
    SCAPI_CATCH(NULL)
}
void ptbh_flatten_desc(void *flat, PtbSceneDesc *out) { *out = ((ptb200::Flattener *)flat)->desc(); }
void ptbh_flatten_free(void *flat) { delete(ptb200::Flattener *)flat; }

// Whole-frame render through the C++ host layer (ptb200::renderFrame).
int ptbh_render(void *world, const PtbRenderParams *params, float *rgba, PtbStats *stats)
{
    SCAPI_TRY
    ptb200::renderFrame((const PathTrace::Object *)world, *params, rgba, stats);
    return 0;
    SCAPI_CATCH(-1)
}

// Walks a GPU-backed SpanIterator from Object::makeSpanIterator() for each ray, the way the
// reference-side ref_trace_spans walks the reference's iterators.
int ptbh_trace_spans(void *world, const float *origins, const float *dirs, size_t n, PtbSpan *spans, int max_spans, int *counts)
{
    SCAPI_TRY
    PathTrace::SpanIterator *it = ((PathTrace::Object *)world)->makeSpanIterator();
    PathTrace::AutoDestruct<PathTrace::SpanIterator> guard(it);
    for(size_t i = 0; i < n; i++)
    {
        PathTrace::Ray ray(PathTrace::Vector3D(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]), PathTrace::Vector3D(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
        it->init(ray);
        int c = 0;
        for(; *it; (*it)++)
        {
            if(c < max_spans)
            {
                PtbSpan &o = spans[i * (size_t)max_spans + c];
                const PathTrace::Span &s = **it;
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
    return 0;
    SCAPI_CATCH(-1)
}

int ptbh_tex_eval(void *tex, const float *pos, size_t n, float *rgb, float *f)
{
    SCAPI_TRY
    PathTrace::Texture *t = (PathTrace::Texture *)tex;
    for(size_t i = 0; i < n; i++)
    {
        PathTrace::Vector3D p(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
        if(rgb)
        {
            PathTrace::Color c = t->getColor(p);
            rgb[3 * i] = c.x, rgb[3 * i + 1] = c.y, rgb[3 * i + 2] = c.z;
        }
        if(f)
            f[i] = t->getFloat(p);
    }
    return 0;
    SCAPI_CATCH(-1)
}

} // extern "C"
