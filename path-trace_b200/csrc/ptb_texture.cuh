// ptb_texture.cuh — texture-program interpreter.
//
// A texture is a linear chain of wrapper ops ending in one leaf (include/ptb200.h).  Position
// wrappers act on the way down, colour filters on the way back up.  Image fetches go through
// texture objects on float4 CUDA arrays with POINT filtering (the reference is nearest-texel,
// SURVEY.md Q8) and border addressing (black outside, as Image::getPixel,
// reference src/image.cpp:366-380).
#pragma once
#include "ptb_scene_dev.cuh"

namespace ptb
{

// ImageTexture::getColor addressing, reference include/image_texture.h:18-28
__device__ __forceinline__ float4 imageLookup(const DevImage &im, V3 v, bool bilinear)
{
    float x = v.x - floorf(v.x);
    float y = v.y - floorf(v.y);
    y = 1.0f - y;
    x *= (float)im.w;
    y *= (float)im.h;
    if(bilinear)
        return tex2D<float4>(im.texLinear, x, y);
    return tex2D<float4>(im.tex, floorf(x) + 0.5f, floorf(y) + 0.5f);
}
// ImageSkyboxTexture::getColor(x, y, image), reference include/image_texture.h:77-85 (the
// reference evaluates x*0.5+0.5 in double and narrows; done the same way here)
__device__ __forceinline__ float4 faceLookup(const DevImage &im, float x, float y, bool bilinear)
{
    x = (float)((double)x * 0.5 + 0.5);
    y = (float)(0.5 - (double)y * 0.5);
    x *= (float)im.w;
    y *= (float)im.h;
    if(bilinear)
        return tex2D<float4>(im.texLinear, x, y);
    return tex2D<float4>(im.tex, floorf(x) + 0.5f, floorf(y) + 0.5f);
}
// cube face by dominant axis with the reference's strict '>' tie rules,
// include/image_texture.h:91-110.  Faces: top, bottom, left, right, front, back.
__device__ __forceinline__ float4 skyboxLookup(const DevImage *im, V3 v, bool bilinear)
{
    if(isZero(v))
        return make_float4(0.f, 0.f, 0.f, 0.f);
    float ax = fabsf(v.x), ay = fabsf(v.y), az = fabsf(v.z);
    if(ax > ay && ax > az)
    {
        if(v.x < 0.f)
            return faceLookup(im[2], __fdiv_rn(-v.z, ax), __fdiv_rn(v.y, ax), bilinear);
        return faceLookup(im[3], __fdiv_rn(v.z, ax), __fdiv_rn(v.y, ax), bilinear);
    }
    if(ay > az)
    {
        if(v.y < 0.f)
            return faceLookup(im[1], __fdiv_rn(-v.x, ay), __fdiv_rn(v.z, ay), bilinear);
        return faceLookup(im[0], __fdiv_rn(v.x, ay), __fdiv_rn(v.z, ay), bilinear);
    }
    if(v.z < 0.f)
        return faceLookup(im[5], __fdiv_rn(v.x, az), __fdiv_rn(v.y, az), bilinear);
    return faceLookup(im[4], __fdiv_rn(-v.x, az), __fdiv_rn(v.y, az), bilinear);
}
// LogTexture::myLog, reference include/filter_texture.h:61-66
__device__ __forceinline__ float texLog(float v)
{
    if(v <= 1e-30f)
        return 0.f;
    return 0.5f + logf(v) / 0.693147180559945f / 256.f;
}
// Texture::getFloat default, reference include/texture.h:14-18
__device__ __forceinline__ float meanRgb(V3 c) { return (c.x + c.y + c.z) * (1.0f / 3.0f); }

// Position wrappers: the affine one uses exact IEEE arithmetic in the reference's order (a few ops
// per shaded hit, decides which texel a point sample lands on).
__device__ __forceinline__ V3 texMirrorBall(V3 p) // include/transform_texture.h:46-59
{
    if(isZero(p))
        return p;
    V3 v = Ar<false>::normalize(p);
    if(v.z <= -1.f)
        return mk(0.f, 0.5f, 0.f);
    float d = 2.f + 2.f * v.z;
    if(d <= 0.f)
        return mk(0.f, 0.5f, 0.f);
    float h = 0.5f * rsqrtf(d);
    return mk(fmaf(v.x, h, 0.5f), fmaf(v.y, h, 0.5f), 0.f);
}
// The reference narrows atan2/asin to float and then maps to [0,1] in double; binary32 throughout
// here (FP64 issues at 1/64 rate on B200).  A lookup moves across a texel edge for < 0.1 % of
// positions (tests/test_gpu_render.py::test_texture_known_answers allows 1 %).
__device__ __forceinline__ V3 texSpherical(V3 p) // include/transform_texture.h:73-85
{
    if(isZero(p))
        return p;
    V3 v = Ar<false>::normalize(p);
    float theta = atan2f(v.y, v.x); // |theta| <= pi in float: the reference's +-2pi wrap never fires
    float phi = asinf(fminf(fmaxf(v.z, -1.f), 1.f));
    return mk(fmaf(theta, 0.15915494309189535f, 0.5f), fmaf(phi, 0.3183098861837907f, 0.5f), 0.f);
}

// One texture chain at one position: rgb in xyz, Texture::getFloat in w.  Inlined ONCE per kernel: shadeVertex
// evaluates the material slots that need it in a loop around this function.  (Inlined at each of the five slots the
// interpreter below made the shade kernels 6 600 instructions long and 9 % of their stall samples instruction
// fetches, profiles/r02k_full_c5.txt; as an out-of-line function the call spilled the path state around it and the
// camera-ray shade kernel lost what it had won, profiles/r02n.)
//
// Folded chains (DevTexBaked, ptb_scene_dev.cuh) run straight through; the rest are interpreted: position wrappers
// forward getFloat (texture.h:78-81, transform_texture.h:30-33), filters do not override it, so under a filter the
// chain runs as getColor and the mean is taken above.
__device__ __forceinline__ float4 texEval(const DevTexBaked *__restrict__ baked, const int2 *__restrict__ texops, const float *__restrict__ texdata,
                                          const DevImage *__restrict__ images, int root, float x, float y, float z, bool bilinear)
{
    V3 pos = mk(x, y, z), color;
    float f;
    const float4 *rec = reinterpret_cast<const float4 *>(baked + root);
    const float4 head = __ldg(rec), r0 = __ldg(rec + 1), r1 = __ldg(rec + 2), r2 = __ldg(rec + 3), mul = __ldg(rec + 4); // (all five at once)
    const uint32_t fl = __float_as_uint(head.x);
    if(fl & PTB_TB_VALID)
    {
        const DevImage *im = images + __float_as_int(head.y);
        if(fl & PTB_TB_AFFINE) // TransformedTexture, include/texture.h:74-81
            pos = Ar<true>::apply(r0, r1, r2, pos);
        if(fl & PTB_TB_SPHERICAL)
            pos = texSpherical(pos);
        else if(fl & PTB_TB_MIRRORBALL)
            pos = texMirrorBall(pos);
        const float4 px = (fl & PTB_TB_SKYBOX) ? skyboxLookup(im, pos, bilinear) : imageLookup(*im, pos, bilinear);
        color = (fl & PTB_TB_ALPHA) ? mk(px.w, px.w, px.w) : mk(px.x, px.y, px.z); // include/image_texture.h:45-66,135-176
        if(fl & PTB_TB_LOG_FIRST)
            color = mk(texLog(color.x), texLog(color.y), texLog(color.z));
        if(fl & PTB_TB_MUL) // include/filter_texture.h:43-46
            color = mk(color.x * mul.x, color.y * mul.y, color.z * mul.z);
        if((fl & (PTB_TB_LOG | PTB_TB_LOG_FIRST)) == PTB_TB_LOG) // include/filter_texture.h:68-71
            color = mk(texLog(color.x), texLog(color.y), texLog(color.z));
        f = (fl & PTB_TB_ALPHA) && !(fl & (PTB_TB_MUL | PTB_TB_LOG)) ? px.w : meanRgb(color);
        return make_float4(color.x, color.y, color.z, f);
    }
    int i = root;
    int2 op = __ldg(&texops[i]);
    bool filtered = false;
    while(!PTB_TEX_IS_LEAF(op.x))
    {
        if(op.x == PTB_TEX_AFFINE) // TransformedTexture, include/texture.h:74-81
        {
            const float *m = texdata + op.y;
            float4 r0 = make_float4(__ldg(m), __ldg(m + 1), __ldg(m + 2), __ldg(m + 3));
            float4 r1 = make_float4(__ldg(m + 4), __ldg(m + 5), __ldg(m + 6), __ldg(m + 7));
            float4 r2 = make_float4(__ldg(m + 8), __ldg(m + 9), __ldg(m + 10), __ldg(m + 11));
            pos = Ar<true>::apply(r0, r1, r2, pos);
        }
        else if(op.x == PTB_TEX_MIRRORBALL)
            pos = texMirrorBall(pos);
        else if(op.x == PTB_TEX_SPHERICAL)
            pos = texSpherical(pos);
        else
            filtered = true;
        op = __ldg(&texops[++i]);
    }
    float4 px;
    bool alpha = false;
    if(op.x == PTB_TEX_CONST)
    {
        const float *d = texdata + op.y;
        px = make_float4(__ldg(d), __ldg(d + 1), __ldg(d + 2), 0.f);
    }
    else if(op.x == PTB_TEX_IMAGE || op.x == PTB_TEX_IMAGE_ALPHA)
    {
        px = imageLookup(images[op.y], pos, bilinear);
        alpha = op.x == PTB_TEX_IMAGE_ALPHA;
    }
    else
    {
        px = skyboxLookup(images + op.y, pos, bilinear);
        alpha = op.x == PTB_TEX_SKYBOX_ALPHA;
    }
    if(alpha) // getColor = Color(alpha), getFloat = alpha (include/image_texture.h:45-66,135-176)
    {
        color = mk(px.w, px.w, px.w);
        f = px.w;
    }
    else
    {
        color = mk(px.x, px.y, px.z);
        f = meanRgb(color);
    }
    if(filtered)
    {
        for(int k = i - 1; k >= root; k--)
        {
            int2 w = __ldg(&texops[k]);
            if(w.x == PTB_TEX_MULTIPLY) // include/filter_texture.h:43-46
            {
                const float *d = texdata + w.y;
                color = mk(color.x * __ldg(d), color.y * __ldg(d + 1), color.z * __ldg(d + 2));
            }
            else if(w.x == PTB_TEX_LOG) // include/filter_texture.h:68-71
                color = mk(texLog(color.x), texLog(color.y), texLog(color.z));
        }
        f = meanRgb(color);
    }
    return make_float4(color.x, color.y, color.z, f);
}

// Evaluates texture chain `root` at pos: colour (Texture::getColor) and float (Texture::getFloat) values.
__device__ __forceinline__ void evalTexture(const DevScene &sc, int root, V3 pos, bool wantFloat, bool bilinear, V3 &color, float &f)
{
    (void)wantFloat;
    const float4 v = texEval(sc.texBaked, sc.texops, sc.texdata, sc.images, root, pos.x, pos.y, pos.z, bilinear);
    color = mk(v.x, v.y, v.z), f = v.w;
}

__device__ __forceinline__ V3 texColor(const DevScene &sc, int root, V3 pos, bool bilinear)
{
    V3 c;
    float f;
    evalTexture(sc, root, pos, false, bilinear, c, f);
    return c;
}
__device__ __forceinline__ float texFloat(const DevScene &sc, int root, V3 pos, bool bilinear)
{
    V3 c;
    float f;
    evalTexture(sc, root, pos, true, bilinear, c, f);
    return f;
}

} // namespace ptb
