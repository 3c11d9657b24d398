// ptb200_scene.hpp — host-side C++ scene API of the B200 render path.
//
// Same class names, constructors, ownership rules and error behaviour as the reference's
// scene API (reference include/{vector3d,transform,ray,misc,color,span,object,sphere,plane,
// union,difference,intersection,material,texture,image_texture,filter_texture,
// transform_texture,image}.h), so scene-building code such as makeWorld()
// (reference src/test.cpp:107-145) compiles against either.  What differs is what sits behind
// it: instead of a tree of CPU SpanIterators, every Object / Material / Texture knows how to
// *flatten* itself into the linear SoA CSG program of include/ptb200.h, which the CUDA
// wavefront integrator executes.  There is no CPU render path in this library.
//
// Ownership: combinators and TransformedObject own and delete their
// children; objects borrow `const Material *`; a Material owns its five textures; wrapper
// textures own their delegate; Image is a shared, reference-counted handle.
#ifndef PTB200_SCENE_HPP_INCLUDED
#define PTB200_SCENE_HPP_INCLUDED

#include <assert.h>
#include <math.h>
#include <stdint.h>

#include <cmath>
#include <iostream>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "ptb200.h"

// ---- image errors and image handles (global namespace, as reference include/image.h) -------

class ImageLoadError : public std::runtime_error
{
public:
    explicit ImageLoadError(const std::string &what) : std::runtime_error(what) {}
};

class ImageStoreError : public std::runtime_error
{
public:
    explicit ImageStoreError(const std::string &what) : std::runtime_error(what) {}
};

namespace PathTrace
{

// reference include/misc.h:7-8
const float eps = 1e-3;
const float max_value = 1e20;

template <typename T>
class AutoDestruct
{
    T *held;

public:
    AutoDestruct(T *v) : held(v) {}
    ~AutoDestruct() { delete held; }
};

// reference include/vector3d.h:15-34 — closed [min,max] uniform from a 32-bit generator.
template <typename T>
class uniform_real_distribution
{
    T lo, hi;

public:
    uniform_real_distribution(T lo, T hi) : lo(lo), hi(hi) {}
    template <typename RG>
    T operator()(RG &rg) const
    {
        T u = rg();
        u -= (int64_t)rg.min();
        u /= (int64_t)rg.max() - (int64_t)rg.min();
        u *= hi - lo;
        u += lo;
        return u;
    }
};

// reference include/vector3d.h:36-219.  Host-side only: used to describe scenes.  Every
// operation is a single correctly-rounded binary32 op in the reference's order, so values that
// end up in the flattened scene are bit-identical to what the reference would hold.
class Vector3D
{
public:
    float x, y, z;
    Vector3D() : x(0), y(0), z(0) {}
    Vector3D(float x, float y, float z) : x(x), y(y), z(z) {}
    Vector3D(float v) : x(v), y(v), z(v) {}

    friend Vector3D operator+(const Vector3D &a, const Vector3D &b) { return Vector3D(a.x + b.x, a.y + b.y, a.z + b.z); }
    friend Vector3D operator-(const Vector3D &a, const Vector3D &b) { return Vector3D(a.x - b.x, a.y - b.y, a.z - b.z); }
    friend Vector3D operator*(const Vector3D &a, const Vector3D &b) { return Vector3D(a.x * b.x, a.y * b.y, a.z * b.z); }
    friend Vector3D operator/(const Vector3D &a, const Vector3D &b) { return Vector3D(a.x / b.x, a.y / b.y, a.z / b.z); }
    friend Vector3D operator*(const Vector3D &a, float s) { return Vector3D(a.x * s, a.y * s, a.z * s); }
    friend Vector3D operator*(float s, const Vector3D &a) { return a * s; }
    friend Vector3D operator/(const Vector3D &a, float s) { return Vector3D(a.x / s, a.y / s, a.z / s); }
    Vector3D operator-() const { return Vector3D(-x, -y, -z); }
    friend bool operator==(const Vector3D &a, const Vector3D &b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
    friend bool operator!=(const Vector3D &a, const Vector3D &b) { return !(a == b); }
    const Vector3D &operator+=(const Vector3D &r) { return *this = *this + r; }
    const Vector3D &operator-=(const Vector3D &r) { return *this = *this - r; }
    const Vector3D &operator*=(const Vector3D &r) { return *this = *this * r; }
    const Vector3D &operator/=(const Vector3D &r) { return *this = *this / r; }
    const Vector3D &operator*=(float r) { return *this = *this * r; }
    const Vector3D &operator/=(float r) { return *this = *this / r; }

    friend float dot(const Vector3D &a, const Vector3D &b)
    {
        Vector3D p = a * b;
        return p.x + p.y + p.z;
    }
    friend float abs_squared(const Vector3D &v) { return dot(v, v); }
    friend float abs(const Vector3D &v) { return std::sqrt(abs_squared(v)); }
    friend Vector3D normalize(const Vector3D &v)
    {
        float len = abs(v);
        return v / (len == 0 ? 1.0f : len);
    }

    // reference include/vector3d.h:164-185 — rejection-sampled point in a ball.
    template <typename RG>
    static Vector3D rand(RG &r, float max = 1, float min = 0)
    {
        assert(max >= 0);
        assert(min <= max || min == 0);
        if(max == 0)
            return Vector3D();
        uniform_real_distribution<float> side(-max, max);
        Vector3D p;
        float len;
        do
        {
            p.x = side(r);
            p.y = side(r);
            p.z = side(r);
            len = abs(p);
        } while(len > max || (min > 0 && len == 0));
        if(min > 0)
            p *= ((max + len * (max - min)) / len);
        return p;
    }
    // reference include/vector3d.h:186-214
    Vector3D reflect(Vector3D normal) const
    {
        normal = normalize(normal);
        return *this - 2 * dot(*this, normal) * normal;
    }
    float refractStrength(float relative_ior, Vector3D normal) const
    {
        if(relative_ior < eps || relative_ior > 1 / eps || normal == Vector3D(0) || *this == Vector3D(0))
            return 0;
        normal = normalize(normal);
        float c = dot(normalize(*this), normal);
        float k = 1 - relative_ior * relative_ior * (1 - c * c);
        if(k <= 0)
            return 0;
        return (float)::sqrt(::sqrt((double)k)); // the reference's unqualified sqrt binds to ::sqrt(double)
    }
    Vector3D refract(float relative_ior, Vector3D normal) const
    {
        if(relative_ior < eps || relative_ior > 1 / eps || normal == Vector3D(0) || *this == Vector3D(0))
            return Vector3D(0);
        normal = normalize(normal);
        Vector3D in = normalize(*this);
        float c = dot(in, normal);
        float k = 1 - relative_ior * relative_ior * (1 - c * c);
        if(k < 0)
            return Vector3D(0);
        return normalize(relative_ior * in - (relative_ior * c + std::sqrt(k)) * normal);
    }
    friend std::ostream &operator<<(std::ostream &o, const Vector3D &v) { return o << "<" << v.x << ", " << v.y << ", " << v.z << ">"; }
};

typedef Vector3D Color; // reference include/color.h:8

// reference include/ray.h:9-23 — direction is NOT normalised.
struct Ray
{
    Vector3D origin, dir;
    Ray(Vector3D origin, Vector3D dir) : origin(origin), dir(dir) { assert(this->dir != Vector3D(0)); }
    Vector3D getPoint(float t) const { return origin + t * dir; }
};

// reference include/transform.h:16-422 — 3x4 affine, row-vector convention:
//   apply(v) = (v.x*x00 + v.y*x10 + v.z*x20 + x30, ...)
// Members keep the reference's names and declaration order (it is public data).
class Matrix
{
public:
    float x00, x10, x20, x30;
    float x01, x11, x21, x31;
    float x02, x12, x22, x32;

    Matrix() : x00(1), x10(0), x20(0), x30(0), x01(0), x11(1), x21(0), x31(0), x02(0), x12(0), x22(1), x32(0) {}
    Matrix(float x00, float x10, float x20, float x30, float x01, float x11, float x21, float x31, float x02, float x12, float x22, float x32)
        : x00(x00), x10(x10), x20(x20), x30(x30), x01(x01), x11(x11), x21(x21), x31(x31), x02(x02), x12(x12), x22(x22), x32(x32)
    {
    }

    float get(const int x, const int y) const
    {
        if(x < 0 || x > 3 || y < 0 || y > 2)
            return x == y ? 1 : 0;
        return (&x00)[y * 4 + x];
    }
    void set(const int x, const int y, float value)
    {
        if(x < 0 || x > 3 || y < 0 || y > 2)
            return;
        (&x00)[y * 4 + x] = value;
    }

    static Matrix identity() { return Matrix(); }

    // reference include/transform.h:207-229 (cos/sin in double, narrowed to float)
    static Matrix rotate(const Vector3D axis, const double angle)
    {
        Vector3D u = normalize(axis);
        float c = cos(angle), s = sin(angle);
        float v = 1 - c;
        float xx = u.x * u.x, xy = u.x * u.y, xz = u.x * u.z;
        float yy = u.y * u.y, yz = u.y * u.z, zz = u.z * u.z;
        return Matrix(xx + (1 - xx) * c, xy * v - u.z * s, xz * v + u.y * s, 0,
                      xy * v + u.z * s, yy + (1 - yy) * c, yz * v - u.x * s, 0,
                      xz * v - u.y * s, yz * v + u.x * s, zz + (1 - zz) * c, 0);
    }
    static Matrix rotateX(double angle) { return rotate(Vector3D(1, 0, 0), angle); }
    static Matrix rotateY(double angle) { return rotate(Vector3D(0, 1, 0), angle); }
    static Matrix rotateZ(double angle) { return rotate(Vector3D(0, 0, 1), angle); }
    static Matrix translate(Vector3D p) { return Matrix(1, 0, 0, p.x, 0, 1, 0, p.y, 0, 0, 1, p.z); }
    static Matrix translate(float x, float y, float z) { return Matrix(1, 0, 0, x, 0, 1, 0, y, 0, 0, 1, z); }
    static Matrix scale(float x, float y, float z) { return Matrix(x, 0, 0, 0, 0, y, 0, 0, 0, 0, z, 0); }
    static Matrix scale(Vector3D s) { return scale(s.x, s.y, s.z); }
    static Matrix scale(float s) { return scale(s, s, s); }

    // reference include/transform.h:342-347
    float determinant() const
    {
        return x00 * (x11 * x22 - x12 * x21) + x10 * (x02 * x21 - x01 * x22) + x20 * (x01 * x12 - x02 * x11);
    }
    // reference include/transform.h:350-383 — same cofactor order so results are bit-identical.
    Matrix inverse() const
    {
        float det = determinant();
        if(det == 0.0f)
            throw std::domain_error("can't invert singular matrix");
        float f = 1.0f / det;
        float t0 = -x10 * x21 * x32 + x11 * x20 * x32 + x10 * x22 * x31 - x12 * x20 * x31 - x11 * x22 * x30 + x12 * x21 * x30;
        float t1 = x00 * x21 * x32 - x01 * x20 * x32 - x00 * x22 * x31 + x02 * x20 * x31 + x01 * x22 * x30 - x02 * x21 * x30;
        float t2 = -x00 * x11 * x32 + x01 * x10 * x32 + x00 * x12 * x31 - x02 * x10 * x31 - x01 * x12 * x30 + x02 * x11 * x30;
        return Matrix((x11 * x22 - x12 * x21) * f, (x12 * x20 - x10 * x22) * f, (x10 * x21 - x11 * x20) * f, t0 * f,
                      (x02 * x21 - x01 * x22) * f, (x00 * x22 - x02 * x20) * f, (x01 * x20 - x00 * x21) * f, t1 * f,
                      (x01 * x12 - x02 * x11) * f, (x02 * x10 - x00 * x12) * f, (x00 * x11 - x01 * x10) * f, t2 * f);
    }
    friend Matrix invert(const Matrix &m) { return m.inverse(); }

    // reference include/transform.h:391-406 — "this, then rt".
    Matrix concat(Matrix rt) const
    {
        return Matrix(x00 * rt.x00 + x01 * rt.x10 + x02 * rt.x20,
                      x10 * rt.x00 + x11 * rt.x10 + x12 * rt.x20,
                      x20 * rt.x00 + x21 * rt.x10 + x22 * rt.x20,
                      x30 * rt.x00 + x31 * rt.x10 + x32 * rt.x20 + rt.x30,
                      x00 * rt.x01 + x01 * rt.x11 + x02 * rt.x21,
                      x10 * rt.x01 + x11 * rt.x11 + x12 * rt.x21,
                      x20 * rt.x01 + x21 * rt.x11 + x22 * rt.x21,
                      x30 * rt.x01 + x31 * rt.x11 + x32 * rt.x21 + rt.x31,
                      x00 * rt.x02 + x01 * rt.x12 + x02 * rt.x22,
                      x10 * rt.x02 + x11 * rt.x12 + x12 * rt.x22,
                      x20 * rt.x02 + x21 * rt.x12 + x22 * rt.x22,
                      x30 * rt.x02 + x31 * rt.x12 + x32 * rt.x22 + rt.x32);
    }
    Vector3D apply(Vector3D v) const
    {
        return Vector3D(v.x * x00 + v.y * x10 + v.z * x20 + x30, v.x * x01 + v.y * x11 + v.z * x21 + x31, v.x * x02 + v.y * x12 + v.z * x22 + x32);
    }
    Vector3D applyNoTranslate(Vector3D v) const
    {
        return Vector3D(v.x * x00 + v.y * x10 + v.z * x20, v.x * x01 + v.y * x11 + v.z * x21, v.x * x02 + v.y * x12 + v.z * x22);
    }
    void store(float out[12]) const
    {
        for(int i = 0; i < 12; i++)
            out[i] = (&x00)[i];
    }
};

inline Vector3D transform(const Matrix &m, Vector3D v) { return m.apply(v); }
inline Ray transform(const Matrix &m, Ray r) { return Ray(m.apply(r.origin), m.applyNoTranslate(r.dir)); }
inline bool operator==(const Matrix &a, const Matrix &b)
{
    for(int i = 0; i < 12; i++)
        if((&a.x00)[i] != (&b.x00)[i])
            return false;
    return true;
}
inline bool operator!=(const Matrix &a, const Matrix &b) { return !(a == b); }

} // namespace PathTrace

// ---- flattener ------------------------------------------------------------------------------

namespace ptb200
{
class Image_data;
class DeviceScene;

// Collects the linear scene program while an Object graph is being visited.
class Flattener
{
public:
    Flattener();
    ~Flattener();
    // CSG side
    void pushTransform(const PathTrace::Matrix &m); // throws std::domain_error on a singular matrix
    void popTransform();
    void addSphere(PathTrace::Vector3D center, float r_squared, int material_index);
    void addPlane(PathTrace::Vector3D normal, float d, int material_index);
    void addOp(int kind);
    // material / texture side
    int findMaterial(const void *key) const; // -1 if this Material* has not been flattened yet
    std::vector<const void *> materialKeys() const; // the key behind each material index
    int beginMaterial(const void *key, float ior);
    void beginTexture(int material_index, int slot);
    void addTexOp(int kind, const float *data, int n_data);
    void addTexImageOp(int kind, const std::shared_ptr<Image_data> *images, int n_images);
    // result; pointers stay valid until the Flattener is destroyed or modified
    PtbSceneDesc desc() const;

private:
    struct Impl;
    Impl *impl;
    Flattener(const Flattener &);
    Flattener &operator=(const Flattener &);
};

// Pixel storage shared by Image handles.
class Image_data
{
public:
    std::vector<float> rgba;
    unsigned w, h;
    Image_data(unsigned w, unsigned h) : rgba((size_t)4 * w * h, 0.0f), w(w), h(h) {}
};
} // namespace ptb200

class MutableImage;

// reference include/image.h:48-99 — immutable, shared RGBA32F image.
class Image
{
public:
    explicit Image(std::string fileName, std::string format = ""); // .hdr/.pic (Radiance RLE), .png; throws ImageLoadError
    Image() {}
    PathTrace::Color getPixel(int x, int y) const
    {
        if(!data || y < 0 || (unsigned)y >= data->h || x < 0 || (unsigned)x >= data->w)
            return PathTrace::Color();
        const float *p = &data->rgba[4 * (x + (size_t)y * data->w)];
        return PathTrace::Color(p[0], p[1], p[2]);
    }
    float getPixelAlpha(int x, int y) const
    {
        if(!data || y < 0 || (unsigned)y >= data->h || x < 0 || (unsigned)x >= data->w)
            return 0;
        return data->rgba[4 * (x + (size_t)y * data->w) + 3];
    }
    unsigned width() const { return data->w; }
    unsigned height() const { return data->h; }
    operator bool() const { return (bool)data; }
    bool operator!() const { return !data; }
    friend bool operator==(Image l, Image r) { return l.data == r.data; }
    friend bool operator!=(Image l, Image r) { return l.data != r.data; }
    const std::shared_ptr<ptb200::Image_data> &shared() const { return data; }

private:
    std::shared_ptr<ptb200::Image_data> data;
    friend class MutableImage;
};

// reference include/image.h:101-213
class MutableImage
{
    std::vector<float> px;
    unsigned w, h;

public:
    MutableImage(unsigned w, unsigned h) : px((size_t)4 * w * h, 0.0f), w(w), h(h) { assert(w > 0 && h > 0); }
    explicit MutableImage(Image img)
    {
        if(!img)
            throw std::runtime_error("can't create MutableImage from empty Image");
        w = img.width();
        h = img.height();
        px = img.data->rgba;
    }
    PathTrace::Color getPixel(int x, int y) const
    {
        if(y < 0 || (unsigned)y >= h || x < 0 || (unsigned)x >= w)
            return PathTrace::Color();
        const float *p = &px[4 * (x + (size_t)y * w)];
        return PathTrace::Color(p[0], p[1], p[2]);
    }
    PathTrace::Color getPixelAlpha(int x, int y) const
    {
        if(y < 0 || (unsigned)y >= h || x < 0 || (unsigned)x >= w)
            return PathTrace::Color();
        return px[4 * (x + (size_t)y * w) + 3];
    }
    void setPixel(int x, int y, PathTrace::Color c, float alpha = 1)
    {
        if(y < 0 || (unsigned)y >= h || x < 0 || (unsigned)x >= w)
            return;
        float *p = &px[4 * (x + (size_t)y * w)];
        p[0] = c.x;
        p[1] = c.y;
        p[2] = c.z;
        p[3] = alpha;
    }
    unsigned width() const { return w; }
    unsigned height() const { return h; }
    float *pixels() { return px.data(); }
    const float *pixels() const { return px.data(); }
    operator Image() const
    {
        Image out;
        out.data = std::make_shared<ptb200::Image_data>(w, h);
        out.data->rgba = px;
        return out;
    }
    void writeHDR(std::string fileName) const; // Radiance RLE writer; throws ImageStoreError
};

namespace PathTrace
{

// ---- textures (reference include/texture.h, image_texture.h, filter_texture.h,
//      transform_texture.h).  getColor/getFloat evaluate the flattened texture program on the
//      GPU (one point per call; meant for inspection, the renderer evaluates them in-kernel). ---

class Texture
{
public:
    virtual Color getColor(Vector3D pos) const;
    virtual float getFloat(Vector3D pos) const;
    virtual Texture *duplicate() const = 0;
    virtual Texture *transform(const Matrix &) const { return NULL; }
    virtual ~Texture() {}
    // Appends this texture's program (outermost op first, leaf last).
    virtual void flattenTexture(ptb200::Flattener &f) const = 0;
};

class ColorTexture : public Texture
{
    Color color;

public:
    ColorTexture(Color color) : color(color) {}
    ColorTexture(float r, float g, float b) : color(r, g, b) {}
    ColorTexture(float v) : color(v) {}
    virtual Texture *duplicate() const { return new ColorTexture(color); }
    virtual Texture *transform(const Matrix &) const { return new ColorTexture(color); }
    virtual void flattenTexture(ptb200::Flattener &f) const
    {
        float d[3] = {color.x, color.y, color.z};
        f.addTexOp(PTB_TEX_CONST, d, 3);
    }
};

class TransformedTexture : public Texture
{
    Matrix m;
    Texture *t;

public:
    TransformedTexture(const Matrix &m, Texture *t) : m(m), t(t) {}
    virtual ~TransformedTexture() { delete t; }
    virtual Texture *duplicate() const { return new TransformedTexture(m, t->duplicate()); }
    virtual Texture *transform(const Matrix &outer) const { return new TransformedTexture(this->m.concat(outer), t->duplicate()); }
    virtual void flattenTexture(ptb200::Flattener &f) const
    {
        float d[12];
        m.store(d);
        f.addTexOp(PTB_TEX_AFFINE, d, 12);
        t->flattenTexture(f);
    }
};

inline Texture *transform(const Matrix &m, Texture *t)
{
    Texture *r = t->transform(m);
    return r ? r : new TransformedTexture(m, t->duplicate());
}

class ImageTexture : public Texture
{
    Image image;

public:
    ImageTexture(Image image) : image(image) {}
    virtual Texture *duplicate() const { return new ImageTexture(image); }
    virtual void flattenTexture(ptb200::Flattener &f) const { f.addTexImageOp(PTB_TEX_IMAGE, &image.shared(), 1); }
};

class ImageAlphaTexture : public Texture
{
    Image image;

public:
    ImageAlphaTexture(Image image) : image(image) {}
    virtual Texture *duplicate() const { return new ImageAlphaTexture(image); }
    virtual void flattenTexture(ptb200::Flattener &f) const { f.addTexImageOp(PTB_TEX_IMAGE_ALPHA, &image.shared(), 1); }
};

class ImageSkyboxTexture : public Texture
{
    Image face[6]; // top, bottom, left, right, front, back

public:
    ImageSkyboxTexture(Image top, Image bottom, Image left, Image right, Image front, Image back)
    {
        face[0] = top, face[1] = bottom, face[2] = left, face[3] = right, face[4] = front, face[5] = back;
    }
    virtual Texture *duplicate() const { return new ImageSkyboxTexture(face[0], face[1], face[2], face[3], face[4], face[5]); }
    virtual void flattenTexture(ptb200::Flattener &f) const
    {
        std::shared_ptr<ptb200::Image_data> s[6];
        for(int i = 0; i < 6; i++)
            s[i] = face[i].shared();
        f.addTexImageOp(PTB_TEX_SKYBOX, s, 6);
    }
};

class ImageSkyboxAlphaTexture : public Texture
{
    Image face[6];

public:
    ImageSkyboxAlphaTexture(Image top, Image bottom, Image left, Image right, Image front, Image back)
    {
        face[0] = top, face[1] = bottom, face[2] = left, face[3] = right, face[4] = front, face[5] = back;
    }
    virtual Texture *duplicate() const { return new ImageSkyboxAlphaTexture(face[0], face[1], face[2], face[3], face[4], face[5]); }
    virtual void flattenTexture(ptb200::Flattener &f) const
    {
        std::shared_ptr<ptb200::Image_data> s[6];
        for(int i = 0; i < 6; i++)
            s[i] = face[i].shared();
        f.addTexImageOp(PTB_TEX_SKYBOX_ALPHA, s, 6);
    }
};

class FilterTexture : public Texture
{
protected:
    Texture *const t;
    FilterTexture(Texture *t) : t(t) {}

public:
    virtual ~FilterTexture() { delete t; }
};

class MultiplyTexture : public FilterTexture
{
    Color factor;

public:
    MultiplyTexture(Color factor, Texture *t) : FilterTexture(t), factor(factor) {}
    virtual Texture *duplicate() const { return new MultiplyTexture(factor, t->duplicate()); }
    virtual void flattenTexture(ptb200::Flattener &f) const
    {
        float d[3] = {factor.x, factor.y, factor.z};
        f.addTexOp(PTB_TEX_MULTIPLY, d, 3);
        t->flattenTexture(f);
    }
};

class LogTexture : public FilterTexture
{
public:
    LogTexture(Texture *t) : FilterTexture(t) {}
    virtual Texture *duplicate() const { return new LogTexture(t->duplicate()); }
    virtual void flattenTexture(ptb200::Flattener &f) const
    {
        f.addTexOp(PTB_TEX_LOG, NULL, 0);
        t->flattenTexture(f);
    }
};

class TransformTexture : public Texture
{
protected:
    Texture *const t;
    TransformTexture(Texture *t) : t(t) {}

public:
    virtual ~TransformTexture() { delete t; }
};

class MirrorBallSkymapTexture : public TransformTexture
{
public:
    MirrorBallSkymapTexture(Texture *t) : TransformTexture(t) {}
    virtual Texture *duplicate() const { return new MirrorBallSkymapTexture(t->duplicate()); }
    virtual void flattenTexture(ptb200::Flattener &f) const
    {
        f.addTexOp(PTB_TEX_MIRRORBALL, NULL, 0);
        t->flattenTexture(f);
    }
};

class SphericalCoordinatesSkymapTexture : public TransformTexture
{
public:
    SphericalCoordinatesSkymapTexture(Texture *t) : TransformTexture(t) {}
    virtual Texture *duplicate() const { return new SphericalCoordinatesSkymapTexture(t->duplicate()); }
    virtual void flattenTexture(ptb200::Flattener &f) const
    {
        f.addTexOp(PTB_TEX_SPHERICAL, NULL, 0);
        t->flattenTexture(f);
    }
};

// ---- material (reference include/material.h:10-42) -----------------------------------------

struct Material
{
    Texture *reflect;
    Texture *scatter_coefficient; // 0 = specular ... 1 = diffuse
    Texture *emissive;
    Texture *transmit;
    float ior;
    Texture *transmit_reflect_coefficient; // 0 = reflect ... 1 = transmit
    Material(Texture *reflect = new ColorTexture(1), Texture *scatter_coefficient = new ColorTexture(1), Texture *emissive = new ColorTexture(0),
             Texture *transmit = new ColorTexture(0), float ior = 1, Texture *transmit_reflect_coefficient = new ColorTexture(0))
        : reflect(reflect), scatter_coefficient(scatter_coefficient), emissive(emissive), transmit(transmit), ior(ior),
          transmit_reflect_coefficient(transmit_reflect_coefficient)
    {
    }
    ~Material()
    {
        delete reflect;
        delete scatter_coefficient;
        delete emissive;
        delete transmit;
        delete transmit_reflect_coefficient;
    }
    Material *duplicate() const
    {
        return new Material(reflect->duplicate(), scatter_coefficient->duplicate(), emissive->duplicate(), transmit->duplicate(), ior,
                            transmit_reflect_coefficient->duplicate());
    }
    // Index of this material in the flattened scene (flattening it on first use).
    int flattenMaterial(ptb200::Flattener &f) const;

private:
    Material(const Material &);
    const Material &operator=(const Material &);
};

inline Material *transform(const Matrix &m, const Material *mat)
{
    return new Material(transform(m, mat->reflect), transform(m, mat->scatter_coefficient), transform(m, mat->emissive), transform(m, mat->transmit),
                        mat->ior, transform(m, mat->transmit_reflect_coefficient));
}

// ---- spans (reference include/span.h) --------------------------------------------------------

class Span
{
public:
    float start;
    Vector3D startNormal;
    const Material *startMaterial;
    float end;
    Vector3D endNormal;
    const Material *endMaterial;
    Span() : start(0), startMaterial(NULL), end(0), endMaterial(NULL) {}
    Span(float start, Vector3D startNormal, const Material *startMaterial, float end, Vector3D endNormal, const Material *endMaterial)
        : start(start), startNormal(startNormal), startMaterial(startMaterial), end(end), endNormal(endNormal), endMaterial(endMaterial)
    {
    }
    virtual ~Span() {}
    bool isEmpty() const { return end <= start; }
    operator bool() const { return end > start; }
    bool operator!() const { return end <= start; }
};

// The pull-iterator protocol of reference include/span.h:129-171.
class SpanIterator
{
public:
    virtual const Span &operator*() const = 0;
    virtual const Span *operator->() const = 0;
    virtual bool isAtEnd() const = 0;
    virtual void next() = 0;
    virtual void init(const Ray &ray) = 0;
    virtual ~SpanIterator() {}
    bool operator!() const { return isAtEnd(); }
    operator bool() const { return !isAtEnd(); }
    void operator++(int) { next(); }
    const SpanIterator &operator++()
    {
        next();
        return *this;
    }

protected:
    SpanIterator() {}

private:
    SpanIterator(const SpanIterator &);
    const SpanIterator &operator=(const SpanIterator &);
};

// ---- objects (reference include/object.h, sphere.h, plane.h, union.h, difference.h,
//      intersection.h) --------------------------------------------------------------------------

class Object
{
public:
    Object() {}
    // A GPU-backed iterator: init(ray) runs the CSG kernel for that one ray on the device and
    // the iterator then walks the returned root span list.  Caller owns the result.
    virtual SpanIterator *makeSpanIterator() const;
    virtual Object *transform(const Matrix &) const { return NULL; }
    virtual ~Object() {}
    virtual Object *duplicate() const = 0;
    // Appends this object's postfix CSG program.
    virtual void flatten(ptb200::Flattener &f) const = 0;

private:
    Object(const Object &);
    const Object &operator=(const Object &);
};

class TransformedObject : public Object
{
    Matrix m;
    Object *o;

public:
    TransformedObject(const Matrix &m, Object *o) : m(m), o(o) {}
    virtual ~TransformedObject() { delete o; }
    virtual Object *transform(const Matrix &outer) const { return new TransformedObject(this->m.concat(outer), o->duplicate()); }
    virtual Object *duplicate() const { return new TransformedObject(m, o->duplicate()); }
    virtual void flatten(ptb200::Flattener &f) const
    {
        f.pushTransform(m);
        o->flatten(f);
        f.popTransform();
    }
};

inline Object *transform(const Matrix &m, Object *o)
{
    Object *r = o->transform(m);
    return r ? r : new TransformedObject(m, o->duplicate());
}

class Sphere : public Object
{
    Vector3D center;
    const Material *material;
    float r, r_squared;

public:
    Sphere(Vector3D center, float r, const Material *material) : center(center), material(material), r(r), r_squared(r * r) {}
    virtual Object *duplicate() const { return new Sphere(center, r, material); }
    virtual void flatten(ptb200::Flattener &f) const { f.addSphere(center, r_squared, material->flattenMaterial(f)); }
};

class Plane : public Object
{
    const Vector3D normal;
    const float d;
    const Material *const material;

public:
    Plane(Vector3D normal, float d, const Material *material) : normal(normal), d(d), material(material) {}
    Plane(Vector3D normal, Vector3D pos, const Material *material) : normal(normal), d(-dot(normal, pos)), material(material) {}
    virtual Object *duplicate() const { return new Plane(normal, d, material); }
    virtual void flatten(ptb200::Flattener &f) const { f.addPlane(normal, d, material->flattenMaterial(f)); }
};

// The three combinators share everything but the op code.  transform() reproduces the
// reference's behaviour of building both operands from `a` (reference
// include/union.h:19-22, difference.h:19-22, intersection.h:19-22).
template <int OpKind, typename Self>
class CsgPair : public Object
{
protected:
    Object *const a;
    Object *const b;
    CsgPair(Object *a, Object *b) : a(a), b(b) {}

public:
    virtual ~CsgPair()
    {
        delete a;
        delete b;
    }
    virtual Object *duplicate() const { return new Self(a->duplicate(), b->duplicate()); }
    virtual Object *transform(const Matrix &m) const { return new Self(PathTrace::transform(m, a), PathTrace::transform(m, a)); }
    virtual void flatten(ptb200::Flattener &f) const
    {
        a->flatten(f);
        b->flatten(f);
        f.addOp(OpKind);
    }
};

class Union : public CsgPair<PTB_OP_UNION, Union>
{
public:
    Union(Object *a, Object *b) : CsgPair<PTB_OP_UNION, Union>(a, b) {}
};
class Difference : public CsgPair<PTB_OP_DIFFERENCE, Difference>
{
public:
    Difference(Object *a, Object *b) : CsgPair<PTB_OP_DIFFERENCE, Difference>(a, b) {}
};
class Intersection : public CsgPair<PTB_OP_INTERSECTION, Intersection>
{
public:
    Intersection(Object *a, Object *b) : CsgPair<PTB_OP_INTERSECTION, Intersection>(a, b) {}
};

} // namespace PathTrace

#endif // PTB200_SCENE_HPP_INCLUDED
