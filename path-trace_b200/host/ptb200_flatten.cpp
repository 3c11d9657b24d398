// ptb200_flatten.cpp — the scene flattener: Object / Material / Texture graph -> the linear SoA
// CSG program of include/ptb200.h.
//
// Nested TransformedObject wrappers are kept as a chain of separate matrices (not concatenated)
// so the device applies them one after another in the reference's rounding order
// (reference include/object.h:66-70 applies transform(m, ray) once per wrapper level).
#include "ptb200_scene.hpp"

#include <map>

namespace ptb200
{

struct Flattener::Impl
{
    std::vector<PtbPrim> prims;
    std::vector<PtbXform> xforms;
    std::vector<PtbChain> chains;
    std::vector<int32_t> chainXf;
    std::vector<uint32_t> ops;
    std::vector<PtbMaterial> materials;
    std::vector<PtbTexOp> texops;
    std::vector<float> texdata;
    std::vector<PtbImageDesc> images;
    std::vector<std::shared_ptr<Image_data>> imageRefs; // keeps pixel storage alive
    std::map<const void *, int> materialIndex;
    std::map<const Image_data *, int> imageIndex;
    std::vector<int> xformStack; // indices into xforms, outermost first
    int currentChain;            // chain id of xformStack, -2 = not interned yet
    Impl() : currentChain(-1) {}

    int internChain()
    {
        if(currentChain != -2)
            return currentChain;
        if(xformStack.empty())
            return currentChain = -1;
        PtbChain c;
        c.first = (int32_t)chainXf.size();
        c.count = (int32_t)xformStack.size();
        chainXf.insert(chainXf.end(), xformStack.begin(), xformStack.end());
        chains.push_back(c);
        return currentChain = (int)chains.size() - 1;
    }
    int internImage(const std::shared_ptr<Image_data> &img)
    {
        if(!img)
            throw std::runtime_error("empty Image used in a texture");
        std::map<const Image_data *, int>::iterator it = imageIndex.find(img.get());
        if(it != imageIndex.end())
            return it->second;
        PtbImageDesc d;
        d.rgba = img->rgba.data();
        d.width = (int32_t)img->w;
        d.height = (int32_t)img->h;
        images.push_back(d);
        imageRefs.push_back(img);
        return imageIndex[img.get()] = (int)images.size() - 1;
    }
};

Flattener::Flattener() : impl(new Impl) {}
Flattener::~Flattener() { delete impl; }

void Flattener::pushTransform(const PathTrace::Matrix &m)
{
    PtbXform x;
    m.store(x.m);
    m.inverse().store(x.inv); // throws std::domain_error when singular, as invert(m) in the
                              // reference's TransformedSpanIterator ctor (include/object.h:46)
    impl->xforms.push_back(x);
    impl->xformStack.push_back((int)impl->xforms.size() - 1);
    impl->currentChain = -2;
}

void Flattener::popTransform()
{
    impl->xformStack.pop_back();
    impl->currentChain = -2;
}

void Flattener::addSphere(PathTrace::Vector3D center, float r_squared, int material_index)
{
    PtbPrim p;
    p.p[0] = center.x, p.p[1] = center.y, p.p[2] = center.z, p.p[3] = r_squared;
    p.type = PTB_PRIM_SPHERE;
    p.material = material_index;
    p.chain = impl->internChain();
    p.reserved = 0;
    impl->prims.push_back(p);
    impl->ops.push_back(PTB_OP(PTB_OP_PUSH_PRIM, impl->prims.size() - 1));
}

void Flattener::addPlane(PathTrace::Vector3D normal, float d, int material_index)
{
    PtbPrim p;
    p.p[0] = normal.x, p.p[1] = normal.y, p.p[2] = normal.z, p.p[3] = d;
    p.type = PTB_PRIM_PLANE;
    p.material = material_index;
    p.chain = impl->internChain();
    p.reserved = 0;
    impl->prims.push_back(p);
    impl->ops.push_back(PTB_OP(PTB_OP_PUSH_PRIM, impl->prims.size() - 1));
}

void Flattener::addOp(int kind) { impl->ops.push_back(PTB_OP(kind, 0)); }

int Flattener::findMaterial(const void *key) const
{
    std::map<const void *, int>::const_iterator it = impl->materialIndex.find(key);
    return it == impl->materialIndex.end() ? -1 : it->second;
}

std::vector<const void *> Flattener::materialKeys() const
{
    std::vector<const void *> keys(impl->materials.size(), (const void *)NULL);
    for(std::map<const void *, int>::const_iterator it = impl->materialIndex.begin(); it != impl->materialIndex.end(); ++it)
        keys[it->second] = it->first;
    return keys;
}

int Flattener::beginMaterial(const void *key, float ior)
{
    PtbMaterial m;
    for(int i = 0; i < 5; i++)
        m.tex[i] = -1;
    m.ior = ior;
    m.reserved[0] = m.reserved[1] = 0;
    impl->materials.push_back(m);
    return impl->materialIndex[key] = (int)impl->materials.size() - 1;
}

void Flattener::beginTexture(int material_index, int slot) { impl->materials[material_index].tex[slot] = (int32_t)impl->texops.size(); }

void Flattener::addTexOp(int kind, const float *data, int n_data)
{
    PtbTexOp op;
    op.kind = kind;
    op.arg = (int32_t)impl->texdata.size();
    impl->texdata.insert(impl->texdata.end(), data, data + n_data);
    impl->texops.push_back(op);
}

void Flattener::addTexImageOp(int kind, const std::shared_ptr<Image_data> *images, int n_images)
{
    // A skybox needs its six faces at consecutive indices: re-register them together (cheap
    // descriptors, the pixel storage is shared) unless they already happen to be consecutive.
    int first = -1;
    if(n_images == 1)
        first = impl->internImage(images[0]);
    else
    {
        first = (int)impl->images.size();
        for(int i = 0; i < n_images; i++)
        {
            if(!images[i])
                throw std::runtime_error("empty Image used in a skybox texture");
            PtbImageDesc d;
            d.rgba = images[i]->rgba.data();
            d.width = (int32_t)images[i]->w;
            d.height = (int32_t)images[i]->h;
            impl->images.push_back(d);
            impl->imageRefs.push_back(images[i]);
        }
    }
    PtbTexOp op;
    op.kind = kind;
    op.arg = first;
    impl->texops.push_back(op);
}

PtbSceneDesc Flattener::desc() const
{
    PtbSceneDesc d;
    d.abi_version = PTB200_ABI_VERSION;
    d.n_prims = (int32_t)impl->prims.size();
    d.n_xforms = (int32_t)impl->xforms.size();
    d.n_chains = (int32_t)impl->chains.size();
    d.n_chain_xf = (int32_t)impl->chainXf.size();
    d.n_ops = (int32_t)impl->ops.size();
    d.n_materials = (int32_t)impl->materials.size();
    d.n_texops = (int32_t)impl->texops.size();
    d.n_texdata = (int32_t)impl->texdata.size();
    d.n_images = (int32_t)impl->images.size();
    d.prims = impl->prims.data();
    d.xforms = impl->xforms.data();
    d.chains = impl->chains.data();
    d.chain_xf = impl->chainXf.data();
    d.ops = impl->ops.data();
    d.materials = impl->materials.data();
    d.texops = impl->texops.data();
    d.texdata = impl->texdata.data();
    d.images = impl->images.data();
    return d;
}

} // namespace ptb200

namespace PathTrace
{

int Material::flattenMaterial(ptb200::Flattener &f) const
{
    int index = f.findMaterial(this);
    if(index >= 0)
        return index;
    index = f.beginMaterial(this, ior);
    const Texture *slots[5] = {reflect, scatter_coefficient, emissive, transmit, transmit_reflect_coefficient};
    for(int s = 0; s < 5; s++)
    {
        f.beginTexture(index, s);
        slots[s]->flattenTexture(f);
    }
    return index;
}

} // namespace PathTrace
