// ptb_capi.cu — implementation of the C ABI in include/ptb200.h: scene compilation and upload,
// launch configuration, the wavefront driver.  Host C++; every entry point catches everything
// and returns a status code (nothing throws across the boundary).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "ptb_internal.h"
#include "ptb_csg_fast.cuh"

using namespace ptb;

namespace ptb
{
namespace
{
thread_local std::string g_error;
}
int fail(int code, const std::string &msg)
{
    g_error = msg;
    return code;
}
const std::string &lastError() { return g_error; }
void setLastError(const std::string &msg) { g_error = msg; }
} // namespace ptb

namespace
{
// Scratch blocks are cached per process instead of going back to the driver: cudaMalloc / cudaFree of
// the multi-GB ray queues cost from milliseconds to seconds (cudaFree synchronises the device and
// unmaps) and a renderer that creates and destroys scenes per frame would pay that every frame.
// ptb200_trim() returns the cache to the driver; an allocation that fails trims and retries.
struct CachedBlock
{
    void *p;
    size_t bytes;
    int device;
};
static std::mutex g_cacheLock;
static std::vector<CachedBlock> g_cache;

// Image arrays (cudaArray, RGBA32F) of destroyed scenes, reused by size: cudaMallocArray / cudaFreeArray
// of the 82 MB environment map cost tens of milliseconds per scene.
struct CachedArray
{
    cudaArray_t array;
    int w, h, device;
};
static std::vector<CachedArray> g_arrayCache;
// Pinned staging for host -> device uploads from pageable memory (two halves, used alternately).
static std::mutex g_stagingLock;
static char *g_staging = nullptr;
static const size_t kStagingHalf = size_t(8) << 20;
static cudaEvent_t g_stagingDone[2] = {nullptr, nullptr};
static int stagingInit()
{
    if(!g_staging)
    {
        CU(cudaMallocHost((void **)&g_staging, 2 * kStagingHalf));
        CU(cudaEventCreateWithFlags(&g_stagingDone[0], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&g_stagingDone[1], cudaEventDisableTiming));
    }
    return PTB_OK;
}
} // namespace

namespace ptb
{
void cacheTrim(int device)
{ // device < 0: every device
    std::vector<CachedBlock> drop;
    std::vector<CachedArray> dropArrays;
    {
        std::lock_guard<std::mutex> g(g_cacheLock);
        for(size_t i = 0; i < g_arrayCache.size();)
            if(device < 0 || g_arrayCache[i].device == device)
            {
                dropArrays.push_back(g_arrayCache[i]);
                g_arrayCache[i] = g_arrayCache.back();
                g_arrayCache.pop_back();
            }
            else
                i++;
        for(size_t i = 0; i < g_cache.size();)
            if(device < 0 || g_cache[i].device == device)
            {
                drop.push_back(g_cache[i]);
                g_cache[i] = g_cache.back();
                g_cache.pop_back();
            }
            else
                i++;
    }
    int current = 0;
    cudaGetDevice(&current);
    for(const CachedBlock &b : drop)
    {
        cudaSetDevice(b.device);
        cudaFree(b.p);
    }
    for(const CachedArray &a : dropArrays)
    {
        cudaSetDevice(a.device);
        cudaFreeArray(a.array);
    }
    cudaSetDevice(current);
}

void DeviceBuffer::release()
{
    if(p)
    {
        std::lock_guard<std::mutex> g(g_cacheLock);
        g_cache.push_back(CachedBlock{p, bytes, device});
    }
    p = nullptr, bytes = 0;
}
cudaError_t DeviceBuffer::ensure(size_t n)
{
    if(n <= bytes)
        return cudaSuccess;
    release();
    n = std::max<size_t>(n, size_t(1) << 20); // small blocks too go through the cache: a cudaMalloc / cudaFree pair per scene and buffer stalls for up to 100 ms now and then in a process that holds > 100 GB
    cudaError_t e = cudaGetDevice(&device);
    if(e != cudaSuccess)
        return e;
    { // best fit among the cached blocks of this device, at most twice the size asked for
        std::lock_guard<std::mutex> g(g_cacheLock);
        size_t best = g_cache.size();
        for(size_t i = 0; i < g_cache.size(); i++)
            if(g_cache[i].device == device && g_cache[i].bytes >= n && g_cache[i].bytes / 2 <= n && (best == g_cache.size() || g_cache[i].bytes < g_cache[best].bytes))
                best = i;
        if(best != g_cache.size())
        {
            p = g_cache[best].p, bytes = g_cache[best].bytes;
            g_cache[best] = g_cache.back();
            g_cache.pop_back();
            return cudaSuccess;
        }
    }
    e = cudaMalloc(&p, n);
    if(e != cudaSuccess)
    {
        cudaGetLastError();
        cacheTrim(device);
        e = cudaMalloc(&p, n);
    }
    if(e == cudaSuccess)
        bytes = n;
    else
        p = nullptr;
    return e;
}

// Pinned host blocks for the per-level counters a stats call reads back: kept per process like the device blocks
// (cudaMallocHost / cudaFreeHost per scene are synchronising calls).
static std::vector<std::pair<unsigned *, size_t>> g_hostCounterCache;
static unsigned *hostCountersGet(size_t n, size_t &got)
{
    {
        std::lock_guard<std::mutex> g(g_cacheLock);
        for(size_t i = 0; i < g_hostCounterCache.size(); i++)
            if(g_hostCounterCache[i].second >= n)
            {
                unsigned *p = g_hostCounterCache[i].first;
                got = g_hostCounterCache[i].second;
                g_hostCounterCache[i] = g_hostCounterCache.back();
                g_hostCounterCache.pop_back();
                return p;
            }
    }
    unsigned *p = nullptr;
    got = std::max<size_t>(n, 4096);
    if(cudaMallocHost((void **)&p, got * sizeof(unsigned)) != cudaSuccess)
    {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}
static void hostCountersPut(unsigned *p, size_t n)
{
    std::lock_guard<std::mutex> g(g_cacheLock);
    if(g_hostCounterCache.size() < 16)
        g_hostCounterCache.push_back(std::make_pair(p, n));
    else
        cudaFreeHost(p);
}

RenderScratch::~RenderScratch()
{
    if(hostCounters)
        hostCountersPut(hostCounters, hostCountersLen);
    for(cudaEvent_t e : events)
        cudaEventDestroy(e);
}

// Is this host buffer page-locked (cudaHostAlloc / cudaHostRegister / torch pin_memory)?  Then it is DMA-ed from
// where it lies; a pageable one goes through the pinned staging halves (a CPU copy at ~10 GB/s).
bool hostIsPinned(const void *p)
{
    cudaPointerAttributes attr;
    if(cudaPointerGetAttributes(&attr, p) != cudaSuccess)
    {
        cudaGetLastError();
        return false;
    }
    return attr.type == cudaMemoryTypeHost;
}
// Host -> linear device memory from a pageable buffer, through the pinned staging halves: the CPU copy of
// piece k+1 overlaps the DMA of piece k (the driver's own pageable path ran at < 2 GB/s here).
int uploadLinear(void *dst, const void *src, size_t bytes)
{
    if(hostIsPinned(src))
    {
        CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, nullptr));
        CU(cudaStreamSynchronize(nullptr));
        return PTB_OK;
    }
    std::lock_guard<std::mutex> g(g_stagingLock);
    if(int rc = stagingInit())
        return rc;
    int k = 0;
    for(size_t at = 0; at < bytes; at += kStagingHalf, k ^= 1)
    {
        const size_t n = std::min(kStagingHalf, bytes - at);
        CU(cudaEventSynchronize(g_stagingDone[k])); // the copy that last used this half has finished
        memcpy(g_staging + k * kStagingHalf, (const char *)src + at, n);
        CU(cudaMemcpyAsync((char *)dst + at, g_staging + k * kStagingHalf, n, cudaMemcpyHostToDevice, nullptr));
        CU(cudaEventRecord(g_stagingDone[k], nullptr));
    }
    CU(cudaStreamSynchronize(nullptr));
    return PTB_OK;
}
} // namespace ptb

// ---- device-side ordering of a scene's shared scratch ---------------------------------------------------
int PtbScene::begin(cudaStream_t stream)
{
    if(!lastUse)
        CU(cudaEventCreateWithFlags(&lastUse, cudaEventDisableTiming));
    if(lastUsePending && stream != lastUseStream)
        CU(cudaStreamWaitEvent(stream, lastUse, 0)); // the previous call (another stream) still owns queues, hits, counters, overflow
    return PTB_OK;
}
int PtbScene::end(cudaStream_t stream)
{
    if(!lastUse)
        CU(cudaEventCreateWithFlags(&lastUse, cudaEventDisableTiming));
    CU(cudaEventRecord(lastUse, stream));
    lastUseStream = stream, lastUsePending = true;
    return PTB_OK;
}
void PtbScene::quiesce()
{
    if(lastUse && lastUsePending)
    {
        cudaEventSynchronize(lastUse);
        lastUsePending = false;
    }
}

namespace
{
// The scene's tables are laid out in ONE host image and go to the device with one copy into one block from the
// library's cache: ten cudaMalloc + cudaMemcpy pairs per ptb200_scene_create (and ten cudaFree per destroy) cost
// 5-240 ms a scene, erratically, in a process that holds tens of GB of queues.
struct TableImage
{
    std::vector<char> host;
    template <typename T>
    size_t add(const std::vector<T> &v)
    {
        const size_t at = (host.size() + 255) & ~(size_t)255;
        host.resize(at + std::max<size_t>(v.size(), 1) * sizeof(T), 0);
        if(!v.empty())
            memcpy(host.data() + at, v.data(), v.size() * sizeof(T));
        return at;
    }
    size_t reserve(size_t bytes)
    {
        const size_t at = (host.size() + 255) & ~(size_t)255;
        host.resize(at + bytes, 0);
        return at;
    }
};

// deviceLinear != NULL: the pixels are already on this device as a linear RGBA32F buffer (multi-GPU create)
int createImage(const float *rgba, int w, int h, ImageSlot &slot, const float4 *deviceLinear = nullptr)
{
    if((!rgba && !deviceLinear) || w <= 0 || h <= 0)
        return fail(PTB_EINVAL, "image with no pixels");
    cudaChannelFormatDesc fmt = cudaCreateChannelDesc<float4>();
    int device = 0;
    CU(cudaGetDevice(&device));
    {
        std::lock_guard<std::mutex> g(g_cacheLock);
        for(size_t i = 0; i < g_arrayCache.size(); i++)
            if(g_arrayCache[i].w == w && g_arrayCache[i].h == h && g_arrayCache[i].device == device)
            {
                slot.array = g_arrayCache[i].array;
                g_arrayCache[i] = g_arrayCache.back();
                g_arrayCache.pop_back();
                break;
            }
    }
    if(!slot.array)
        CU(cudaMallocArray(&slot.array, &fmt, (size_t)w, (size_t)h));
    slot.w = w, slot.h = h;
    const size_t rowBytes = (size_t)w * 16;
    if(deviceLinear)
        CU(cudaMemcpy2DToArray(slot.array, 0, 0, deviceLinear, rowBytes, rowBytes, (size_t)h, cudaMemcpyDeviceToDevice));
    else if((size_t)h * rowBytes < (size_t(1) << 20) || rowBytes > kStagingHalf || hostIsPinned(rgba))
        CU(cudaMemcpy2DToArray(slot.array, 0, 0, rgba, rowBytes, rowBytes, (size_t)h, cudaMemcpyHostToDevice));
    else
    { // the caller's buffer is pageable: copy it through pinned staging in 8 MB pieces, the CPU copy of
      // piece k+1 overlapping the DMA of piece k (the driver's own pageable path ran at < 2 GB/s here)
        std::lock_guard<std::mutex> g(g_stagingLock);
        if(int rc = stagingInit())
            return rc;
        const int rowsPerPiece = (int)(kStagingHalf / rowBytes);
        int k = 0;
        for(int y = 0; y < h; y += rowsPerPiece, k ^= 1)
        {
            const int rows = std::min(rowsPerPiece, h - y);
            CU(cudaEventSynchronize(g_stagingDone[k])); // the copy that last used this half has finished
            memcpy(g_staging + k * kStagingHalf, (const char *)rgba + (size_t)y * rowBytes, (size_t)rows * rowBytes);
            CU(cudaMemcpy2DToArrayAsync(slot.array, 0, (size_t)y, g_staging + k * kStagingHalf, rowBytes, rowBytes, (size_t)rows, cudaMemcpyHostToDevice, nullptr));
            CU(cudaEventRecord(g_stagingDone[k], nullptr));
        }
        CU(cudaStreamSynchronize(nullptr));
    }
    cudaResourceDesc res;
    memset(&res, 0, sizeof(res));
    res.resType = cudaResourceTypeArray;
    res.res.array.array = slot.array;
    cudaTextureDesc td;
    memset(&td, 0, sizeof(td));
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeBorder; // black outside, as Image::getPixel
    td.filterMode = cudaFilterModePoint;                           // nearest texel
    td.readMode = cudaReadModeElementType;
    td.normalizedCoords = 0;
    CU(cudaCreateTextureObject(&slot.point, &res, &td, nullptr));
    td.filterMode = cudaFilterModeLinear;
    CU(cudaCreateTextureObject(&slot.linear, &res, &td, nullptr));
    slot.w = w, slot.h = h;
    return PTB_OK;
}

int syncImageTable(PtbScene *s)
{
    std::vector<DevImage> table(s->images.size());
    for(size_t i = 0; i < table.size(); i++)
    {
        table[i].tex = s->images[i].point;
        table[i].texLinear = s->images[i].linear;
        table[i].w = s->images[i].w;
        table[i].h = s->images[i].h;
    }
    if((int)table.size() > s->imageCapacity)
    {
        int cap = std::max<int>(16, (int)table.size() * 2);
        DevImage *p = nullptr;
        CU(cudaMalloc((void **)&p, cap * sizeof(DevImage)));
        s->allocs.push_back(p);
        s->dImages = p;
        s->imageCapacity = cap;
    }
    if(!table.empty())
        CU(cudaMemcpy(s->dImages, table.data(), table.size() * sizeof(DevImage), cudaMemcpyHostToDevice));
    s->dev.images = s->dImages;
    s->dev.nImages = (int)table.size();
    return PTB_OK;
}

// Checks a texture chain: wrappers then exactly one leaf, all arguments in range.
int validateTexture(const PtbSceneDesc *d, int root)
{
    if(root < 0 || root >= d->n_texops)
        return fail(PTB_EINVAL, "texture root out of range");
    for(int i = root;; i++)
    {
        if(i >= d->n_texops)
            return fail(PTB_EINVAL, "texture chain runs past the end of texops");
        const PtbTexOp &op = d->texops[i];
        switch(op.kind)
        {
        case PTB_TEX_CONST:
        case PTB_TEX_MULTIPLY:
            if(op.arg < 0 || op.arg + 3 > d->n_texdata)
                return fail(PTB_EINVAL, "texture data offset out of range");
            break;
        case PTB_TEX_AFFINE:
            if(op.arg < 0 || op.arg + 12 > d->n_texdata)
                return fail(PTB_EINVAL, "texture matrix offset out of range");
            break;
        case PTB_TEX_IMAGE:
        case PTB_TEX_IMAGE_ALPHA:
            if(op.arg < 0)
                return fail(PTB_EINVAL, "negative image index");
            break;
        case PTB_TEX_SKYBOX:
        case PTB_TEX_SKYBOX_ALPHA:
            if(op.arg < 0)
                return fail(PTB_EINVAL, "negative image index");
            break;
        case PTB_TEX_LOG:
        case PTB_TEX_MIRRORBALL:
        case PTB_TEX_SPHERICAL:
            break;
        default:
            return fail(PTB_EINVAL, "unknown texture op");
        }
        if(PTB_TEX_IS_LEAF(op.kind))
            return PTB_OK;
    }
}
} // namespace

// ---- scene ------------------------------------------------------------------------------------
extern "C" const char *ptb200_last_error(void) { return lastError().c_str(); }
extern "C" int ptb200_abi_version(void) { return PTB200_ABI_VERSION; }
extern "C" int ptb200_trim(void)
{
    cacheTrim(-1);
    return PTB_OK;
}
extern "C" int ptb200_device_count(void)
{
    int n = 0;
    if(cudaGetDeviceCount(&n) != cudaSuccess)
        return 0;
    return n;
}

// ---- fast-program compiler ------------------------------------------------------------------------
// Item of the postfix sequence the fast stream is emitted from.
struct FastItem
{
    int kind;  // PTB_OP_*
    int prim;  // PUSH_PRIM: the primitive
    int clusterEnd; // first leaf of a plane cluster: sequence index of the Union that completes it, else -1
};

// Compiles the (validated) postfix program for the fast evaluator, ptb_csg_fast.cuh: chains folded
// into one matrix per leaf in double precision, planes moved to world space, the operator after a
// leaf stored as the leaf's action, further operators packed behind it, a SKIP word before the right
// operand of every Difference / Intersection, and a guarded CLUSTER record in front of every maximal
// subtree that is a Union of planes only.
//
// Plane clusters.  Every plane yields one half-line, (-inf, t] or [t, +inf).  When every lower end lies
// strictly below every upper start, UnionSpanIterator (src/union.cpp:84-134) reduces any tree of such
// half-lines to {(-inf, max lower end], [min upper start, +inf)} whatever the tree's shape: lists of that
// form merge into a list of that form (the two lower spans tie on their start, so the right operand
// absorbs the left and keeps the larger end, :124-133; of the two upper spans the earlier start absorbs
// the other, :113-123).  The device folds that in registers and, if the whole warp is in that case, acts
// as one leaf and jumps over the subtree's own words.  When a lower end reaches an upper start the
// reference's result depends on the tree (an extended span is not re-merged with later spans of its own
// list), so the warp falls through to the subtree's words, which follow the record unchanged.
static void compileFastProgram(const PtbSceneDesc *d, std::vector<uint4> &out, int &maxStack, int &poolBound, FastTree *tree = nullptr)
{
    auto f2u = [](float f) {
        uint32_t u;
        memcpy(&u, &f, 4);
        return u;
    };
    auto u2f = [](uint32_t u) {
        float f;
        memcpy(&f, &u, 4);
        return (double)f;
    };
    std::vector<FastItem> seq;
    for(int i = 0; i < d->n_ops; i++)
    {
        uint32_t kind = PTB_OP_KIND(d->ops[i]);
        seq.push_back(FastItem{(int)kind, kind == PTB_OP_PUSH_PRIM ? (int)PTB_OP_ARG(d->ops[i]) : -1, -1});
    }
    { // maximal subtrees made of planes and Unions only, at least two planes
        struct Sub
        {
            int first;
            bool pure;
        };
        std::vector<Sub> stack, sub(seq.size());
        for(size_t i = 0; i < seq.size(); i++)
        {
            if(seq[i].kind == PTB_OP_PUSH_PRIM)
                stack.push_back(Sub{(int)i, d->prims[seq[i].prim].type == PTB_PRIM_PLANE});
            else
            {
                Sub b = stack.back();
                stack.pop_back();
                Sub &a = stack.back();
                a.pure = a.pure && b.pure && seq[i].kind == PTB_OP_UNION;
            }
            sub[i] = stack.back();
        }
        int covered = (int)seq.size(); // items from here on already belong to a cluster
        for(int i = (int)seq.size() - 1; i >= 0; i--)
            if(i < covered && seq[i].kind != PTB_OP_PUSH_PRIM && sub[i].pure && (i - sub[i].first) / 2 + 1 <= 255)
            {
                seq[sub[i].first].clusterEnd = i;
                covered = sub[i].first;
            }
    }
    const int nItems = (int)seq.size();
    { // deepest operand stack and worst-case pool use (every primitive hit, nothing fused) of THIS sequence
        std::vector<int> sizes;
        for(int i = 0; i < nItems; i++)
        {
            int total = 0;
            for(int v : sizes)
                total += v;
            if(seq[i].kind == PTB_OP_PUSH_PRIM)
            {
                sizes.push_back(1);
                poolBound = std::max(poolBound, total + sizes.back());
            }
            else
            {
                int lb = sizes.back();
                sizes.pop_back();
                int la = sizes.back();
                sizes.back() = la + lb;
                poolBound = std::max(poolBound, total + la + lb); // merge output is appended, then moved down
            }
            maxStack = std::max(maxStack, (int)sizes.size());
        }
    }
    // where does each operand start?  skipEnd[j] = index of the Difference/Intersection whose
    // right operand starts at sequence position j
    std::vector<int> skipEnd(nItems, -1), first;
    for(int i = 0; i < nItems; i++)
    {
        if(seq[i].kind == PTB_OP_PUSH_PRIM)
        {
            first.push_back(i);
            continue;
        }
        int b = first.back();
        first.pop_back();
        if(seq[i].kind != PTB_OP_UNION)
            skipEnd[b] = i;
    }
    // world -> object matrix of a primitive: M = M_k ... M_1 (chain_xf lists the outermost wrapper,
    // applied first, first)
    auto foldChain = [&](const PtbPrim &p, double M[3][4]) {
        double I[3][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}};
        memcpy(M, I, sizeof(I));
        int nchain = p.chain >= 0 ? d->chains[p.chain].count : 0;
        for(int k = 0; k < nchain; k++)
        {
            const float *m = d->xforms[d->chain_xf[d->chains[p.chain].first + k]].m;
            double N[3][4];
            for(int r = 0; r < 3; r++)
                for(int c = 0; c < 4; c++)
                {
                    double v = c == 3 ? (double)m[4 * r + 3] : 0.0;
                    for(int j = 0; j < 3; j++)
                        v += (double)m[4 * r + j] * M[j][c];
                    N[r][c] = v;
                }
            memcpy(M, N, sizeof(N));
        }
        return nchain;
    };
    auto worldPlane = [&](const PtbPrim &p) { // n' = M^T n, D' = D + n.t
        double M[3][4];
        foldChain(p, M);
        double n[3], D = p.p[3];
        for(int j = 0; j < 3; j++)
            n[j] = (double)p.p[0] * M[0][j] + (double)p.p[1] * M[1][j] + (double)p.p[2] * M[2][j];
        D += (double)p.p[0] * M[0][3] + (double)p.p[1] * M[1][3] + (double)p.p[2] * M[2][3];
        return make_uint4(f2u((float)n[0]), f2u((float)n[1]), f2u((float)n[2]), f2u((float)D));
    };
    auto tagsOf = [](int prim) {
        uint32_t tag = (uint32_t)prim << 1;
        return tag | (tag << 16);
    };
    std::vector<size_t> skipWord(nItems, 0); // SKIP word to patch when item i has been emitted
    std::vector<size_t> leafWord(nItems, (size_t)-1); // header word of every transformed-sphere leaf
    if(tree)
    { // what the generator of the register-resident evaluator needs (ptb_specialize.cu): the postfix sequence
      // itself and where each leaf's (and cluster's) record sits in the stream
        tree->items.resize(nItems);
        for(int i = 0; i < nItems; i++)
            tree->items[i] = FastTreeItem{seq[i].kind, seq[i].prim, seq[i].clusterEnd, -1, -1};
    }
    size_t header = 0; // word that can still take operators
    bool sealed = true, freshLeaf = false;
    // the open cluster record: it takes the operators that follow its subtree, exactly those the
    // subtree's last leaf word takes
    size_t cluster = 0, clusterWords = 0;
    int clusterEnd = -1;
    bool clusterOpen = false, clusterFresh = false;
    auto addOperator = [&](size_t w, bool fresh, uint32_t code) { // to a leaf / OPS word with room
        if(fresh)
        {
            out[w].x = (out[w].x & ~FH_MISS_SHIFT) | ((code + FA_UNION) << 4); // the leaf's action
            if(code == 2u)
                out[w].x |= FH_TAIL; // a missed leaf under Intersection drops the list below
            return;
        }
        uint32_t n = (out[w].x >> 8) & 15u;
        out[w].x |= code << (12 + 2 * n);
        out[w].x = (out[w].x & ~0xF00u) | ((n + 1) << 8) | FH_TAIL;
    };
    for(int i = 0; i < nItems; i++)
    {
        const int kind = seq[i].kind;
        if(kind != PTB_OP_PUSH_PRIM)
        {
            uint32_t code = kind == PTB_OP_UNION ? 0u : kind == PTB_OP_DIFFERENCE ? 1u : 2u;
            bool fresh = freshLeaf && !sealed;
            if(!fresh && (sealed || ((out[header].x >> 8) & 15u) >= PTB_FAST_MAX_POST))
            {
                out.push_back(make_uint4(FK_OPS, 0, 0, 0));
                header = out.size() - 1;
                sealed = false;
                clusterOpen = false;
            }
            addOperator(header, fresh, code);
            if(clusterOpen)
            {
                addOperator(cluster, clusterFresh, code);
                clusterFresh = false;
            }
            freshLeaf = false;
            if(kind != PTB_OP_UNION)
            { // the right operand ends here: close the skipped region
                size_t w = skipWord[i];
                out[w].y = (uint32_t)(out.size() - (w + 1));
                sealed = true;
                clusterOpen = false;
            }
            if(i == clusterEnd)
            { // the subtree is complete: its words end here, and what follows applies to the record too
                out[cluster].w = (uint32_t)(out.size() - (cluster + clusterWords));
                clusterOpen = clusterFresh = !(out[header].x & FK_CONTROL);
                if(!clusterOpen)
                    sealed = true; // the last Union sits in an OPS word the record jumps over
                clusterEnd = -1;
            }
            continue;
        }
        clusterOpen = false;
        if(skipEnd[i] >= 0)
        {
            out.push_back(make_uint4(FK_SKIP, 0, 0, 0));
            skipWord[skipEnd[i]] = out.size() - 1;
        }
        const PtbPrim &p = d->prims[seq[i].prim];
        if(seq[i].clusterEnd >= 0)
        { // header {x, tags, extra planes, words to jump}, plane, then {{tags}, plane} per extra plane
            clusterEnd = seq[i].clusterEnd;
            cluster = out.size();
            if(tree)
                tree->items[i].clusterWord = (int)cluster;
            out.push_back(make_uint4(FK_PLANE | FH_MISS_SHIFT, tagsOf(seq[i].prim), 0, 0));
            out.push_back(worldPlane(p));
            for(int k = i + 1; k < clusterEnd; k++)
                if(seq[k].kind == PTB_OP_PUSH_PRIM)
                {
                    out.push_back(make_uint4(tagsOf(seq[k].prim), 0, 0, 0));
                    out.push_back(worldPlane(d->prims[seq[k].prim]));
                    out[cluster].z++;
                }
            clusterWords = out.size() - cluster;
        }
        if(p.type == PTB_PRIM_SPHERE)
        {
            double M[3][4];
            int nchain = foldChain(p, M);
            uint32_t negR2 = f2u(-p.p[3]);
            if(nchain == 0)
            {
                double c0 = (double)p.p[0] * p.p[0] + (double)p.p[1] * p.p[1] + (double)p.p[2] * p.p[2] - (double)p.p[3];
                out.push_back(make_uint4(FK_SPHERE | FH_MISS_SHIFT, tagsOf(seq[i].prim), negR2, f2u((float)c0)));
                header = out.size() - 1;
                out.push_back(make_uint4(f2u(-p.p[0]), f2u(-p.p[1]), f2u(-p.p[2]), 0));
            }
            else
            {
                double c0 = -(double)p.p[3];
                for(int r = 0; r < 3; r++)
                    c0 += (M[r][3] - (double)p.p[r]) * (M[r][3] - (double)p.p[r]);
                out.push_back(make_uint4(FK_SPHERE_XF | FH_MISS_SHIFT, tagsOf(seq[i].prim), negR2, f2u((float)c0)));
                header = out.size() - 1;
                leafWord[i] = header;
                for(int r = 0; r < 3; r++)
                    out.push_back(make_uint4(f2u((float)M[r][0]), f2u((float)M[r][1]), f2u((float)M[r][2]), f2u((float)(M[r][3] - (double)p.p[r]))));
            }
        }
        else
        {
            out.push_back(make_uint4(FK_PLANE | FH_MISS_SHIFT, tagsOf(seq[i].prim), 0, 0));
            header = out.size() - 1;
            out.push_back(worldPlane(p));
        }
        if(tree)
            tree->items[i].leafWord = (int)header;
        sealed = false;
        freshLeaf = true;
    }
    out.push_back(make_uint4(FK_END, 0, 0, 0));
    // flag union-only tails, then pairs of consecutive transformed spheres (see ptb_csg_fast.cuh)
    auto recordLength = [&](size_t w) { return (out[w].x & FK_SPHERE_XF) ? 4u : (out[w].x & FK_PLANE) ? 2u + 2u * out[w].z : (out[w].x & FK_SPHERE) ? 2u : 1u; };
    for(size_t w = 0; w < out.size(); w += recordLength(w))
    {
        uint32_t x = out[w].x, n = (x >> 8) & 15u;
        if(!(x & FH_TAIL) || n == 0)
            continue;
        bool allUnion = (x & FK_CONTROL) || ((x >> 4) & 3u) != FA_INTERSECTION;
        for(uint32_t k = 0; k < n; k++)
            if((x >> (12 + 2 * k)) & 3u)
                allUnion = false;
        if(allUnion)
            out[w].x |= FH_UNION_TAIL;
    }
    std::vector<size_t> runs; // header word of every run's first leaf
    { // Sphere runs: a complete Union-only subtree of 3..15 transformed spheres whose records are adjacent and
      // which no operator follows directly (it is a left operand, or the root).  If no ray of a warp of CAMERA
      // rays hits any of them, all the subtree does is push one empty list: the evaluator tests the run's
      // discriminants in one loop with one vote and jumps over it (ptb_csg_fast.cuh, FH_RUN); otherwise it
      // runs the leaves one by one as before.  The same leaf tests decide, so nothing changes in the results.
        struct Sub
        {
            int first;
            bool pure;
        };
        std::vector<Sub> stack, sub(seq.size());
        for(size_t i = 0; i < seq.size(); i++)
        {
            if(seq[i].kind == PTB_OP_PUSH_PRIM)
            {
                const PtbPrim &p = d->prims[seq[i].prim];
                stack.push_back(Sub{(int)i, p.type == PTB_PRIM_SPHERE && p.chain >= 0 && d->chains[p.chain].count > 0});
            }
            else
            {
                Sub b = stack.back();
                stack.pop_back();
                Sub &a = stack.back();
                a.pure = a.pure && b.pure && seq[i].kind == PTB_OP_UNION;
            }
            sub[i] = stack.back();
        }
        int covered = (int)seq.size();
        for(int i = (int)seq.size() - 1; i >= 0; i--)
        {
            if(!(i < covered && seq[i].kind != PTB_OP_PUSH_PRIM && sub[i].pure))
                continue;
            const int f = sub[i].first, n = (i - f) / 2 + 1;
            covered = f; // maximal subtrees only
            if(n < 3 || n > 15 || (i + 1 < (int)seq.size() && seq[i + 1].kind != PTB_OP_PUSH_PRIM))
                continue;
            bool adjacent = leafWord[f] != (size_t)-1 && !(out[leafWord[f]].x & FH_TAIL);
            for(int k = f, leaf = 0; k <= i && adjacent; k++)
                if(seq[k].kind == PTB_OP_PUSH_PRIM)
                    adjacent = leafWord[k] == leafWord[f] + 4u * (size_t)leaf++;
            if(adjacent)
            {
                out[leafWord[f]].x |= FH_RUN | ((uint32_t)n << 12);
                runs.push_back(leafWord[f]);
            }
        }
    }
    for(size_t w = 0; w < out.size();)
    {
        if((out[w].x & FK_SPHERE_XF) && !(out[w].x & FH_TAIL) && w + 4 < out.size() && (out[w + 4].x & FK_SPHERE_XF))
        {
            out[w].x |= FH_PAIR;
            w += 8;
        }
        else
            w += recordLength(w);
    }
    { // Bounding spheres of the runs' leaves, behind FK_END (the evaluators never run into them; a run's first header
      // carries their word index in bits 16-29).  A leaf is {x : |M x + t'|^2 <= r^2}: an ellipsoid centred at
      // C = -M^-1 t' that lies within R = r / sigma_min(M) of C.  Used by camera rays only (origin 0): a word is
      // {C, sqrt(|C|^2 - R^2)}.  PTB_FAST_BOUNDS=0 leaves them out.
        const char *e = getenv("PTB_FAST_BOUNDS");
        const bool enabled = !e || atoi(e) != 0;
        for(size_t at = 0; enabled && at < runs.size(); at++)
        {
            const size_t w = runs[at];
            const uint32_t n = (out[w].x >> 12) & 15u;
            const size_t table = out.size();
            if(table + n > 0x3FFFu)
                break;
            std::vector<uint4> words;
            for(uint32_t k = 0; k < n; k++)
            {
                const uint4 *rec = &out[w + 4u * k];
                double A[3][3], t[3], B[3][3];
                for(int r = 0; r < 3; r++)
                {
                    A[r][0] = u2f(rec[1 + r].x), A[r][1] = u2f(rec[1 + r].y), A[r][2] = u2f(rec[1 + r].z), t[r] = u2f(rec[1 + r].w);
                }
                const double det = A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) - A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
                                   A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
                const double r2 = -(double)u2f(rec[0].z);
                if(!(fabs(det) > 1e-12) || !(r2 > 0))
                    break;
                for(int i = 0; i < 3; i++)
                    for(int j = 0; j < 3; j++)
                    { // adjugate / det
                        const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
                        B[j][i] = (A[i1][j1] * A[i2][j2] - A[i1][j2] * A[i2][j1]) / det;
                    }
                double C[3];
                for(int i = 0; i < 3; i++)
                    C[i] = -(B[i][0] * t[0] + B[i][1] * t[1] + B[i][2] * t[2]);
                // sigma_max(B)^2 = largest eigenvalue of S = B^T B: power iteration from each axis (one of them has a
                // component of at least 1/sqrt 3 along the leading eigenvector), the result rounded up; the trace
                // bounds it from above.
                double S[3][3], lambda = 0;
                for(int i = 0; i < 3; i++)
                    for(int j = 0; j < 3; j++)
                        S[i][j] = B[0][i] * B[0][j] + B[1][i] * B[1][j] + B[2][i] * B[2][j];
                for(int start = 0; start < 3; start++)
                {
                    double v[3] = {start == 0 ? 1.0 : 0.0, start == 1 ? 1.0 : 0.0, start == 2 ? 1.0 : 0.0}, len = 0;
                    for(int it = 0; it < 200; it++)
                    {
                        double u[3] = {S[0][0] * v[0] + S[0][1] * v[1] + S[0][2] * v[2], S[1][0] * v[0] + S[1][1] * v[1] + S[1][2] * v[2], S[2][0] * v[0] + S[2][1] * v[1] + S[2][2] * v[2]};
                        len = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
                        if(!(len > 0))
                            break;
                        v[0] = u[0] / len, v[1] = u[1] / len, v[2] = u[2] / len;
                    }
                    lambda = std::max(lambda, len);
                }
                const double trace = S[0][0] + S[1][1] + S[2][2];
                const double sigma2 = std::min(trace, lambda * 1.01);
                const double R2 = r2 * sigma2 * 1.002 + 1e-9;
                // the word: C and sqrt(|C|^2 - R^2) rounded down by a relative 1e-5 (far more than the rounding of the
                // three-term dot product it is compared with); 0 when the origin lies inside: every line meets it
                const double K = C[0] * C[0] + C[1] * C[1] + C[2] * C[2] - R2;
                words.push_back(make_uint4(f2u((float)C[0]), f2u((float)C[1]), f2u((float)C[2]), f2u((float)(K > 0 ? sqrt(K) * (1.0 - 1e-5) : 0.0))));
            }
            if(words.size() != n)
                continue; // a singular matrix: this run keeps the plain discriminant loop
            out.insert(out.end(), words.begin(), words.end());
            out[w].x |= (uint32_t)table << 16;
        }
    }
}

int ptb::sceneCreateImpl(const PtbSceneDesc *d, PtbScene *s, std::vector<uint4> *fastOnly, const std::vector<ImageSource> *sources, FastTree *treeOnly)
{
    if(!d || d->abi_version != PTB200_ABI_VERSION)
        return fail(PTB_EINVAL, "scene description has a different abi_version");
    if(d->n_prims <= 0 || d->n_ops <= 0 || d->n_materials <= 0)
        return fail(PTB_EINVAL, "empty scene");
    if(d->n_prims > 32767)
        return fail(PTB_EINVAL, "more than 32767 primitives (boundary tags are 16 bits)");
    if(!fastOnly)
    {
        int nDev = 0;
        if(cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0)
            return fail(PTB_ENODEV, "no CUDA device: libptb200 has no CPU fallback");
        CU(cudaGetDevice(&s->device));
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, s->device));
        s->smCount = prop.multiProcessorCount;
        s->smemOptin = prop.sharedMemPerBlockOptin;
    }

    // validate tables
    for(int i = 0; i < d->n_chains; i++)
        if(d->chains[i].first < 0 || d->chains[i].count <= 0 || d->chains[i].first + d->chains[i].count > d->n_chain_xf)
            return fail(PTB_EINVAL, "transform chain out of range");
    for(int i = 0; i < d->n_chain_xf; i++)
        if(d->chain_xf[i] < 0 || d->chain_xf[i] >= d->n_xforms)
            return fail(PTB_EINVAL, "transform index out of range");
    for(int i = 0; i < d->n_prims; i++)
    {
        const PtbPrim &p = d->prims[i];
        if(p.type != PTB_PRIM_SPHERE && p.type != PTB_PRIM_PLANE)
            return fail(PTB_EINVAL, "unknown primitive type");
        if(p.material < 0 || p.material >= d->n_materials)
            return fail(PTB_EINVAL, "primitive material out of range");
        if(p.chain < -1 || p.chain >= d->n_chains)
            return fail(PTB_EINVAL, "primitive chain out of range");
    }
    for(int i = 0; i < d->n_materials; i++)
        for(int k = 0; k < 5; k++)
            if(int rc = validateTexture(d, d->materials[i].tex[k]))
                return rc;

    // compile the postfix program into the 16-byte word stream of ptb_csg.cuh and find the
    // deepest list stack and the worst-case span-pool use (every primitive hit, no merging)
    std::vector<uint4> prog;
    std::vector<int> sizes; // worst-case list length of each stack entry
    int maxStack = 0, poolBound = 1, lastChain = -2;
    for(int i = 0; i < d->n_ops; i++)
    {
        uint32_t op = d->ops[i];
        uint32_t kind = PTB_OP_KIND(op);
        if(kind == PTB_OP_PUSH_PRIM)
        {
            uint32_t prim = PTB_OP_ARG(op);
            if((int)prim >= d->n_prims)
                return fail(PTB_EINVAL, "PUSH_PRIM names a primitive out of range");
            const PtbPrim &p = d->prims[prim];
            int nchain = p.chain >= 0 ? d->chains[p.chain].count : 0;
            bool reuse = p.chain >= 0 && p.chain == lastChain;
            prog.push_back(make_uint4(p.type == PTB_PRIM_SPHERE ? DK_SPHERE : DK_PLANE, prim, (uint32_t)nchain, reuse ? PTB_LEAF_REUSE_RAY : 0u));
            uint4 g;
            memcpy(&g, p.p, 16);
            prog.push_back(g);
            if(!reuse)
                for(int k = 0; k < nchain; k++)
                {
                    const PtbXform &x = d->xforms[d->chain_xf[d->chains[p.chain].first + k]];
                    for(int r = 0; r < 3; r++)
                    {
                        uint4 row;
                        memcpy(&row, x.m + 4 * r, 16);
                        prog.push_back(row);
                    }
                }
            lastChain = p.chain;
            sizes.push_back(1);
            int total = 0;
            for(int v : sizes)
                total += v;
            poolBound = std::max(poolBound, total);
        }
        else if(kind == PTB_OP_UNION || kind == PTB_OP_DIFFERENCE || kind == PTB_OP_INTERSECTION)
        {
            if(sizes.size() < 2)
                return fail(PTB_EINVAL, "CSG operator with fewer than two operands on the stack");
            prog.push_back(make_uint4(kind == PTB_OP_UNION ? DK_UNION : kind == PTB_OP_DIFFERENCE ? DK_DIFFERENCE : DK_INTERSECTION, 0, 0, 0));
            int total = 0;
            for(int v : sizes)
                total += v;
            int lb = sizes.back();
            sizes.pop_back();
            int la = sizes.back();
            sizes.pop_back();
            poolBound = std::max(poolBound, total + la + lb); // merge output is appended, then moved down
            sizes.push_back(la + lb);
        }
        else
            return fail(PTB_EINVAL, "unknown CSG op");
        maxStack = std::max(maxStack, (int)sizes.size());
    }
    if(sizes.size() != 1)
        return fail(PTB_EINVAL, "CSG program does not leave exactly one object on the stack");
    std::vector<uint4> fast;
    FastTree tree;
    compileFastProgram(d, fast, maxStack, poolBound, &tree);
    if(maxStack > PTB_FAST_MAX_STACK)
        return fail(PTB_EINVAL, "CSG tree needs more than 32 pending operands (right-deep nesting); rebalance it");
    if(poolBound > 65535)
        return fail(PTB_EINVAL, "scene too large for 16-bit span-pool offsets");
    if(fastOnly)
    { // ptb200_specialize_probe: validation and the compiled stream only, no device
        *fastOnly = fast;
        if(treeOnly)
            *treeOnly = tree;
        return PTB_OK;
    }
    s->fastHost = fast;
    s->fastTree = tree;
    prog.push_back(make_uint4(DK_END, 0, 0, 0));

    std::vector<float4> geo(d->n_prims);
    std::vector<int4> meta(d->n_prims);
    for(int i = 0; i < d->n_prims; i++)
    {
        const PtbPrim &p = d->prims[i];
        geo[i] = make_float4(p.p[0], p.p[1], p.p[2], p.p[3]);
        meta[i] = make_int4(p.type, p.material, p.chain >= 0 ? d->chains[p.chain].first : 0, p.chain >= 0 ? d->chains[p.chain].count : 0);
    }
    std::vector<float4> xf((size_t)d->n_xforms * 6);
    for(int i = 0; i < d->n_xforms; i++)
    {
        memcpy(&xf[6 * (size_t)i], d->xforms[i].m, 48);
        memcpy(&xf[6 * (size_t)i + 3], d->xforms[i].inv, 48);
    }
    std::vector<int> chainXf(d->chain_xf, d->chain_xf + d->n_chain_xf);
    std::vector<DevMaterial> mats(d->n_materials);
    for(int i = 0; i < d->n_materials; i++)
    {
        for(int k = 0; k < 5; k++)
            mats[i].tex[k] = d->materials[i].tex[k];
        mats[i].ior = d->materials[i].ior;
        mats[i].iorIn = (float)(1.0 / d->materials[i].ior); // `ior = 1.0 / material->ior`, path-trace.h:79
        mats[i].constMask = 0;
        mats[i].emissive = mats[i].reflect = mats[i].transmit = make_float4(0.f, 0.f, 0.f, 0.f);
        mats[i].scatterF = mats[i].trcF = mats[i].pad0 = mats[i].pad1 = 0.f;
        for(int k = 0; k < 5; k++)
        {
            const PtbTexOp &op = d->texops[d->materials[i].tex[k]];
            if(op.kind != PTB_TEX_CONST)
                continue;
            const float *c = d->texdata + op.arg;
            const float4 v = make_float4(c[0], c[1], c[2], (c[0] + c[1] + c[2]) * (1.0f / 3.0f)); // meanRgb, ptb_texture.cuh
            mats[i].constMask |= 1 << k;
            if(k == PTB_MAT_EMISSIVE)
                mats[i].emissive = v;
            else if(k == PTB_MAT_REFLECT)
                mats[i].reflect = v;
            else if(k == PTB_MAT_TRANSMIT)
                mats[i].transmit = v;
            else if(k == PTB_MAT_SCATTER)
                mats[i].scatterF = v.w;
            else
                mats[i].trcF = v.w;
        }
        // A vertex on such a material ends its path after the emission term whatever the geometry: no refraction
        // (transmit_reflect_coefficient is the constant 0, path-trace.h:110-111) and a zero reflectance, whose child
        // contributes exactly 0 (:161).  Every sky material is like that (src/test.cpp:99-105); the shade kernels then
        // skip the surface normal (ptb_shade.cuh).
        const int both = (1 << PTB_MAT_TRC) | (1 << PTB_MAT_REFLECT);
        if((mats[i].constMask & both) == both && !(mats[i].trcF > 0.f) && mats[i].reflect.x == 0.f && mats[i].reflect.y == 0.f && mats[i].reflect.z == 0.f)
            mats[i].constMask |= PTB_MAT_TERMINAL;
    }
    std::vector<int2> texops(d->n_texops);
    for(int i = 0; i < d->n_texops; i++)
        texops[i] = make_int2(d->texops[i].kind, d->texops[i].arg);
    std::vector<float> texdata(d->texdata, d->texdata + d->n_texdata);
    // Fold the chains the materials start (DevTexBaked, ptb_scene_dev.cuh); validateTexture has bounded every chain.
    std::vector<DevTexBaked> baked((size_t)std::max(1, d->n_texops));
    memset(baked.data(), 0, baked.size() * sizeof(DevTexBaked));
    const bool foldTextures = !getenv("PTB_TEX_FOLD") || atoi(getenv("PTB_TEX_FOLD")) != 0; // (tests: 0 = every chain through the interpreter)
    for(int i = 0; foldTextures && i < d->n_materials; i++)
        for(int k = 0; k < 5; k++)
        {
            const int root = d->materials[i].tex[k];
            DevTexBaked b;
            memset(&b, 0, sizeof(b));
            uint32_t fl = PTB_TB_VALID;
            bool ok = true, mapped = false;
            int j = root;
            for(; ok && !PTB_TEX_IS_LEAF(d->texops[j].kind); j++)
            {
                const PtbTexOp &op = d->texops[j];
                if(op.kind == PTB_TEX_AFFINE)
                {
                    if((fl & PTB_TB_AFFINE) || mapped) // one matrix, above the map
                        ok = false;
                    const float *m = d->texdata + op.arg;
                    b.r0 = make_float4(m[0], m[1], m[2], m[3]), b.r1 = make_float4(m[4], m[5], m[6], m[7]), b.r2 = make_float4(m[8], m[9], m[10], m[11]);
                    fl |= PTB_TB_AFFINE;
                }
                else if(op.kind == PTB_TEX_MIRRORBALL || op.kind == PTB_TEX_SPHERICAL)
                {
                    if(mapped)
                        ok = false;
                    mapped = true;
                    fl |= op.kind == PTB_TEX_SPHERICAL ? PTB_TB_SPHERICAL : PTB_TB_MIRRORBALL;
                }
                else if(op.kind == PTB_TEX_MULTIPLY)
                {
                    if(fl & PTB_TB_MUL)
                        ok = false;
                    const float *c = d->texdata + op.arg;
                    b.mul = make_float4(c[0], c[1], c[2], 0.f);
                    fl |= PTB_TB_MUL; // (a LogTexture met later lies BELOW this one and acts first)
                }
                else if(op.kind == PTB_TEX_LOG)
                {
                    if(fl & PTB_TB_LOG)
                        ok = false;
                    fl |= PTB_TB_LOG | ((fl & PTB_TB_MUL) ? PTB_TB_LOG_FIRST : 0u);
                }
                else
                    ok = false;
            }
            const int leaf = d->texops[j].kind;
            if(!ok || leaf == PTB_TEX_CONST)
                continue; // (a bare ColorTexture is folded into the material record above)
            if(leaf == PTB_TEX_SKYBOX || leaf == PTB_TEX_SKYBOX_ALPHA)
                fl |= PTB_TB_SKYBOX;
            if(leaf == PTB_TEX_IMAGE_ALPHA || leaf == PTB_TEX_SKYBOX_ALPHA)
                fl |= PTB_TB_ALPHA;
            b.flags = fl, b.image = d->texops[j].arg;
            baked[root] = b;
        }

    DevScene &dev = s->dev;
    memset(&dev, 0, sizeof(dev));
    int rc;
    TableImage img;
    const size_t oProg = img.add(prog), oFast = img.add(fast), oGeo = img.add(geo), oMeta = img.add(meta), oChain = img.add(chainXf), oXf = img.add(xf), oMats = img.add(mats),
                 oTexops = img.add(texops), oTexdata = img.add(texdata), oBaked = img.add(baked);
    const int imageSlots = std::max(16, 2 * d->n_images);
    const size_t oImages = img.reserve((size_t)imageSlots * sizeof(DevImage));
    CU(s->tables.ensure(std::max<size_t>(img.host.size(), size_t(1) << 20))); // (>= 1 MiB: blocks that size go back to the cache)
    char *base = (char *)s->tables.p;
    CU(cudaMemcpy(base, img.host.data(), img.host.size(), cudaMemcpyHostToDevice));
    dev.program = (const uint4 *)(base + oProg), dev.fastProgram = (const uint4 *)(base + oFast), dev.primGeo = (const float4 *)(base + oGeo);
    dev.primMeta = (const int4 *)(base + oMeta), dev.chainXf = (const int *)(base + oChain), dev.xforms = (const float4 *)(base + oXf);
    dev.materials = (const DevMaterial *)(base + oMats), dev.texops = (const int2 *)(base + oTexops), dev.texdata = (const float *)(base + oTexdata);
    dev.texBaked = (const DevTexBaked *)(base + oBaked);
    s->dImages = (DevImage *)(base + oImages), s->imageCapacity = imageSlots;
    dev.programWords = (int)prog.size();
    dev.fastWords = (int)fast.size();
    dev.nPrims = d->n_prims, dev.nMaterials = d->n_materials, dev.nTexops = d->n_texops;
    dev.maxStack = maxStack, dev.poolBound = poolBound;
    for(int i = 0; i < d->n_images; i++)
    {
        ImageSlot slot;
        if((rc = createImage(d->images[i].rgba, d->images[i].width, d->images[i].height, slot, sources && (size_t)i < sources->size() ? (*sources)[i].device : nullptr)))
            return rc;
        s->images.push_back(slot);
    }
    return syncImageTable(s);
}

extern "C" void ptb200_scene_destroy(PtbScene *s)
{
    if(!s)
        return;
    cudaSetDevice(s->device);
    s->quiesce(); // an asynchronous accumulate / ray batch may still be reading the textures and the scratch
    if(s->lastUse)
        cudaEventDestroy(s->lastUse);
    for(ImageSlot &im : s->images)
    {
        if(im.point)
            cudaDestroyTextureObject(im.point);
        if(im.linear)
            cudaDestroyTextureObject(im.linear);
        if(im.array)
        {
            cudaArray_t evict = nullptr;
            int evictDevice = s->device;
            {
                std::lock_guard<std::mutex> g(g_cacheLock);
                if(g_arrayCache.size() >= 32)
                { // bounded: the oldest entry goes back to the driver
                    evict = g_arrayCache.front().array, evictDevice = g_arrayCache.front().device;
                    g_arrayCache.erase(g_arrayCache.begin());
                }
                g_arrayCache.push_back(CachedArray{im.array, im.w, im.h, s->device});
            }
            if(evict)
            {
                cudaSetDevice(evictDevice);
                cudaFreeArray(evict);
                cudaSetDevice(s->device);
            }
        }
    }
    for(void *p : s->allocs)
        cudaFree(p);
    delete s;
}

extern "C" int ptb200_scene_create(const PtbSceneDesc *desc, PtbScene **out)
{
    if(!out)
        return fail(PTB_EINVAL, "null output pointer");
    *out = nullptr;
    try
    {
        PtbScene *s = new PtbScene;
        int rc = sceneCreateImpl(desc, s);
        if(rc)
        {
            std::string keep = lastError();
            ptb200_scene_destroy(s);
            setLastError(keep);
            return rc;
        }
        *out = s;
        return PTB_OK;
    }
    catch(std::bad_alloc &)
    {
        return fail(PTB_ENOMEM, "out of host memory");
    }
    catch(std::exception &e)
    {
        return fail(PTB_EINVAL, e.what());
    }
}

extern "C" int ptb200_image_upload(PtbScene *s, const float *rgba, int width, int height, int *handle)
{
    if(!s || !handle)
        return fail(PTB_EINVAL, "null argument");
    std::lock_guard<std::mutex> g(s->lock);
    CU(cudaSetDevice(s->device));
    ImageSlot slot;
    if(int rc = createImage(rgba, width, height, slot))
        return rc;
    s->images.push_back(slot);
    *handle = (int)s->images.size() - 1;
    return syncImageTable(s);
}

// ---- launch configuration -----------------------------------------------------------------------
// Persistent grid: (resident CTAs per SM) x (SM count).  The span pool keeps `cap` slots per
// thread in shared memory; scenes whose worst case needs more get a global overflow area.
// Scratch shared by the calls on a scene only changes hands once the last call that used it has finished.
static cudaError_t ensureScratch(PtbScene *s, DeviceBuffer &b, size_t n)
{
    if(n > b.bytes)
        s->quiesce();
    return b.ensure(n);
}

static int chooseConfig(PtbScene *s, int which, DeviceBuffer &overflow, CsgConfig &cfg)
{
    const DevScene &dev = s->dev;
    cfg.threads = 128;
    cfg.fast = which != 1;
    const int rays = which == 2 ? traceRaysPerThread() : 1;
    int cap = std::min(dev.poolBound, cfg.fast ? 12 : 16);
    if(const char *e = getenv("PTB200_POOL_CAP"))
        cap = std::max(2, std::min(dev.poolBound, atoi(e)));
    cfg.cap = cap;
    if(cfg.fast)
    {
        cfg.stageProgram = (size_t)dev.fastWords * 16 <= 96 * 1024 ? 1 : 0;
        cfg.smemBytes = fastSharedBytes(dev, cfg.cap, cfg.stageProgram, rays);
    }
    else
    {
        cfg.stageProgram = (size_t)dev.programWords * 16 <= 64 * 1024 ? 1 : 0;
        cfg.smemBytes = csgSharedBytes(dev, cfg.threads, cfg.cap, cfg.stageProgram);
    }
    if(cfg.smemBytes > s->smemOptin)
        return fail(PTB_EINVAL, "scene needs more shared memory per CTA than the device offers");
    int perSm = csgBlocksPerSm(which, cfg.threads, cfg.smemBytes, !cfg.stageProgram);
    if(perSm <= 0)
        return fail(PTB_ECUDA, "kernel does not fit on an SM with this shared-memory size");
    if(const char *e = getenv("PTB200_BLOCKS_PER_SM"))
        perSm = std::max(1, std::min(perSm, atoi(e)));
    cfg.blocks = perSm * s->smCount;
    cfg.ovfStride = (size_t)cfg.blocks * cfg.threads;
    cfg.overflow = nullptr;
    if(dev.poolBound > cfg.cap)
    {
        size_t bytes = (size_t)rays * (dev.poolBound - cfg.cap) * 3 * sizeof(float) * cfg.ovfStride;
        CU(ensureScratch(s, overflow, bytes));
        cfg.overflow = (float *)overflow.p;
    }
    return PTB_OK;
}

// ---- scene-specialised kernels: when to use them, and how they are launched -------------------------
// PTB_SPECIALIZE=0 never (unless PTB_FLAG_SPECIALIZE), =1 compile at first use and wait, default: a call
// with at least 2^26 rays to trace starts the compile in the background and keeps interpreting until
// the kernels are there (they give the same results, so a frame may switch between passes).
#ifndef PTB_SPEC_GEN2_MAX_WORDS
#define PTB_SPEC_GEN2_MAX_WORDS 96 /* measured on B200 (profiles/README.md): config 4's 292 words stay interpreted */
#endif
struct SpecConfig
{
    const SpecKernels *k = nullptr;
    CsgConfig cfg;
    bool cameraOnly = false; // a large program: only the camera-ray kernel is worth specialising (see chooseSpec)
};
static int specMode()
{
    static const int mode = [] {
        const char *e = getenv("PTB_SPECIALIZE");
        return e ? atoi(e) : -1;
    }();
    return mode;
}
// work = rays this call is going to trace.  Leaves sp.k NULL when the call should interpret.
static int chooseSpec(PtbScene *s, uint32_t flags, double work, int which, SpecConfig &sp)
{
    sp.k = nullptr;
    const bool force = (flags & PTB_FLAG_SPECIALIZE) != 0;
    // Straight-line code is ~35 instructions (560 bytes) per leaf where the interpreter reads 64 bytes of
    // program: beyond a few dozen leaves the kernels stop fitting the SM's instruction cache and instruction
    // fetch becomes the bound (ncu, C4's 70 leaves: 39 % fewer instructions executed, but 15-36 stall
    // cycles per issue waiting for instructions — profiles/r01m_spec_*.txt), so only small programs are
    // specialised unless the caller insists.
    static const size_t maxWords = [] {
        const char *e = getenv("PTB_SPEC_MAX_WORDS");
        if(e && atoi(e) > 0)
            return (size_t)atoi(e);
        // (the register-resident generator shares one function among subtrees of the same shape, so the 70 leaves
        // of config 4 compile to a dozen leaves of code: its limit is compile time, not the instruction cache)
        const char *g = getenv("PTB_SPEC_GEN");
        return g && atoi(g) == 1 ? (size_t)96 : (size_t)PTB_SPEC_GEN2_MAX_WORDS;
    }();
    // Larger programs (up to 4096 words) are specialised for CAMERA rays only, and only when the generated code
    // tests bounding spheres before its shape-function calls (ptb_specialize.cu): a warp of camera rays — two pixels
    // — then skips most of the scene (config 4: 87 % of the calls; 1 530 instructions a ray against the interpreter's
    // 2 800, profiles/r02w_*).  Their queued rays stay interpreted: 32 unrelated rays skip nothing, and 130 KB of
    // straight-line code lose to the interpreter's loop on instruction fetch.
    const bool large = s->fastHost.size() > maxWords;
    if((flags & (PTB_FLAG_INTERPRET | PTB_FLAG_EXACT)) || (!force && (specMode() == 0 || (large && which != SPEC_TRACE))) || s->fastHost.size() > 4096)
    {
        if(force)
            return fail(PTB_EINVAL, "PTB_FLAG_SPECIALIZE cannot be combined with PTB_FLAG_INTERPRET / PTB_FLAG_EXACT, and needs a program of at most 4096 words");
        return PTB_OK;
    }
    if(s->specSource[which].empty())
        s->specSource[which] = specializeSource(s->fastHost, s->fastTree, which);
    if(large && !force)
    {
        static const bool largeOff = getenv("PTB_SPEC_LARGE") && atoi(getenv("PTB_SPEC_LARGE")) == 0;
        if(largeOff || s->specSource[which].find("PTB_SPEC_BOUNDED_CALLS") == std::string::npos)
        {
            s->specNever[which] = true;
            return PTB_OK;
        }
        sp.cameraOnly = true;
    }
    const int mode = force || specMode() == 1 ? SPEC_WAIT : work >= 67108864.0 ? SPEC_START : SPEC_PEEK;
    std::string err;
    const SpecKernels *k = specializeGet(s->specSource[which], s->device, mode, err);
    if(!k)
    {
        if(force)
            return fail(PTB_ECUDA, "specialised kernels unavailable: " + err);
        return PTB_OK;
    }
    // launch shape: no program in shared memory, pools and start stack only
    const DevScene &dev = s->dev;
    CsgConfig &cfg = sp.cfg;
    cfg.threads = PTB_FAST_THREADS, cfg.fast = 1, cfg.stageProgram = 0;
    // (the register-resident evaluator appends every general merge above its operands and has no overflow area of
    // its own — a ray that outgrows the pool reruns through the interpreter — so it gets 16 slots: 24 KB per CTA)
    static const bool gen1 = getenv("PTB_SPEC_GEN") && atoi(getenv("PTB_SPEC_GEN")) == 1;
    cfg.cap = std::min(dev.poolBound, gen1 ? 12 : 16);
    if(const char *e = getenv("PTB200_POOL_CAP")) // tests: force lists into the global overflow area
        cfg.cap = std::max(2, std::min(dev.poolBound, atoi(e)));
    cfg.smemBytes = fastSharedBytes(dev, cfg.cap, false, 1);
    if(!gen1)
        cfg.smemBytes += (size_t)dev.fastWords * 16; // room for the module's constant table behind pools and stacks (never more words than the program has)
    if(cfg.smemBytes > s->smemOptin)
        return PTB_OK;
    const int perSm = specOccupancy(which == SPEC_TRACE ? k->tracePrimary : k->intersect, cfg.threads, cfg.smemBytes);
    if(perSm <= 0)
        return PTB_OK;
    cfg.blocks = perSm * s->smCount;
    cfg.ovfStride = (size_t)cfg.blocks * cfg.threads;
    cfg.overflow = nullptr;
    if(dev.poolBound > cfg.cap)
    {
        CU(ensureScratch(s, s->specOverflow, (size_t)(dev.poolBound - cfg.cap) * 3 * sizeof(float) * cfg.ovfStride));
        cfg.overflow = (float *)s->specOverflow.p;
    }
    sp.k = k;
    return PTB_OK;
}
#define SPEC_LAUNCH(fn, cfg, stream, ...)                                                                  \
    do                                                                                                     \
    {                                                                                                      \
        void *args_[] = {__VA_ARGS__};                                                                     \
        std::string err_;                                                                                  \
        if(specLaunch(fn, (cfg).blocks, (cfg).threads, (cfg).smemBytes, stream, args_, err_) != 0)         \
            return fail(PTB_ECUDA, err_);                                                                  \
    } while(0)

// ---- ray-batch intersection -----------------------------------------------------------------------
extern "C" int ptb200_intersect_device(PtbScene *s, const void *d_origins4, const void *d_dirs4, size_t n, void *d_hits, void *d_spans, int max_spans, uint32_t flags,
                                       void *stream)
{
    if(!s || !d_origins4 || !d_dirs4 || !d_hits)
        return fail(PTB_EINVAL, "null argument");
    if(d_spans && max_spans <= 0)
        return fail(PTB_EINVAL, "max_spans must be positive when spans are requested");
    std::lock_guard<std::mutex> g(s->lock);
    CU(cudaSetDevice(s->device));
    CsgConfig cfg;
    bool exact = flags & PTB_FLAG_EXACT;
    if(int rc = chooseConfig(s, exact ? 1 : 0, s->scratch.overflow, cfg))
        return rc;
    if(n == 0)
        return PTB_OK;
    SpecConfig sp;
    if(int rc = chooseSpec(s, flags, (double)n, SPEC_INTERSECT, sp))
        return rc;
    if(int rc = s->begin((cudaStream_t)stream))
        return rc;
    if(sp.k)
    {
        DevScene dev = s->dev;
        int cap = sp.cfg.cap;
        float *ovf = sp.cfg.overflow;
        size_t ovfStride = sp.cfg.ovfStride;
        SPEC_LAUNCH((flags & PTB_FLAG_ORIGIN0) ? sp.k->intersectO0 : sp.k->intersect, sp.cfg, (cudaStream_t)stream, &dev, &cap, &ovf, &ovfStride, (void *)&d_origins4, (void *)&d_dirs4, &n,
                    &d_hits, &d_spans, &max_spans);
        return s->end((cudaStream_t)stream);
    }
    launchIntersect(s->dev, cfg, (const float4 *)d_origins4, (const float4 *)d_dirs4, n, (PtbHit *)d_hits, (PtbSpan *)d_spans, max_spans, exact, (flags & PTB_FLAG_ORIGIN0) != 0, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return s->end((cudaStream_t)stream);
}

extern "C" int ptb200_intersect(PtbScene *s, const float *origins, const float *dirs, size_t n, PtbHit *hits, PtbSpan *spans, int max_spans, uint32_t flags)
{
    if(!s || (n && (!origins || !dirs || !hits)))
        return fail(PTB_EINVAL, "null argument");
    if(n == 0)
        return PTB_OK;
    try
    {
        CU(cudaSetDevice(s->device));
        std::vector<float4> o4(n), d4(n);
        for(size_t i = 0; i < n; i++)
        {
            o4[i] = make_float4(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2], 0.f);
            d4[i] = make_float4(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2], 0.f);
        }
        DeviceBuffer bo, bd, bh, bs;
        CU(bo.ensure(n * 16));
        CU(bd.ensure(n * 16));
        CU(bh.ensure(n * sizeof(PtbHit)));
        if(spans)
            CU(bs.ensure(n * (size_t)max_spans * sizeof(PtbSpan)));
        CU(cudaMemcpy(bo.p, o4.data(), n * 16, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(bd.p, d4.data(), n * 16, cudaMemcpyHostToDevice));
        if(spans)
            CU(cudaMemset(bs.p, 0, n * (size_t)max_spans * sizeof(PtbSpan)));
        if(int rc = ptb200_intersect_device(s, bo.p, bd.p, n, bh.p, spans ? bs.p : nullptr, max_spans, flags, nullptr))
            return rc;
        CU(cudaDeviceSynchronize());
        CU(cudaMemcpy(hits, bh.p, n * sizeof(PtbHit), cudaMemcpyDeviceToHost));
        if(spans)
            CU(cudaMemcpy(spans, bs.p, n * (size_t)max_spans * sizeof(PtbSpan), cudaMemcpyDeviceToHost));
        return PTB_OK;
    }
    catch(std::bad_alloc &)
    {
        return fail(PTB_ENOMEM, "out of host memory");
    }
}

// ---- scene-specialised kernels: inspection -------------------------------------------------------------
extern "C" int ptb200_specialize_probe(const PtbSceneDesc *desc, int which, char *source, size_t source_capacity, size_t *source_length, size_t *cubin_bytes)
{
    try
    {
        std::vector<uint4> fast;
        FastTree tree;
        if(int rc = sceneCreateImpl(desc, nullptr, &fast, nullptr, &tree))
            return rc;
        std::string src;
        if(which == 2)
        { // the compiled fast stream itself, one word per line (tests: the CPU emulation of the evaluators runs the
          // interpreter over these words beside the generated code)
            char buf[64];
            for(const uint4 &w : fast)
            {
                snprintf(buf, sizeof(buf), "%08x %08x %08x %08x\n", w.x, w.y, w.z, w.w);
                src += buf;
            }
            cubin_bytes = nullptr;
        }
        else
            src = specializeSource(fast, tree, which);
        if(source_length)
            *source_length = src.size();
        if(source && source_capacity)
        {
            size_t n = std::min(source_capacity - 1, src.size());
            memcpy(source, src.data(), n);
            source[n] = 0;
        }
        if(cubin_bytes)
        {
            std::vector<char> cubin;
            std::string log;
            if(specializeCompile(src, cubin, log) != 0)
                return fail(PTB_ECUDA, log);
            *cubin_bytes = cubin.size();
        }
        return PTB_OK;
    }
    catch(std::bad_alloc &)
    {
        return fail(PTB_ENOMEM, "out of host memory");
    }
}

extern "C" int ptb200_scene_specialized(PtbScene *s, double *compile_ms, int *passes_specialized, int *passes_total)
{
    if(!s)
        return fail(PTB_EINVAL, "null argument");
    std::lock_guard<std::mutex> g(s->lock);
    if(passes_specialized)
        *passes_specialized = s->lastSpecPasses;
    if(passes_total)
        *passes_total = s->lastPasses;
    const int which = s->specSource[SPEC_TRACE].empty() ? SPEC_INTERSECT : SPEC_TRACE; // a scene only used for ray batches has the ray-batch kernels
    if(s->specSource[which].empty() || s->specNever[which])
        return 0;
    std::string err;
    const SpecKernels *k = specializeGet(s->specSource[which], s->device, SPEC_PEEK, err);
    if(compile_ms)
        *compile_ms = k ? k->compileMs : -1.0; // (-1: asked for, not there yet)
    return k ? (s->lastSpecCameraOnly && which == SPEC_TRACE ? 2 : 1) : 0;
}

// ---- inspection entry points ------------------------------------------------------------------------
extern "C" int ptb200_texture_eval(PtbScene *s, int tex_root, const float *positions, size_t n, float *rgb, float *f, uint32_t flags)
{
    if(!s || (n && !positions))
        return fail(PTB_EINVAL, "null argument");
    if(tex_root < 0 || tex_root >= s->dev.nTexops)
        return fail(PTB_EINVAL, "texture root out of range");
    if(n == 0)
        return PTB_OK;
    std::lock_guard<std::mutex> g(s->lock);
    CU(cudaSetDevice(s->device));
    DeviceBuffer bp, bc, bf;
    CU(bp.ensure(n * 12));
    CU(bc.ensure(n * 12));
    CU(bf.ensure(n * 4));
    CU(cudaMemcpy(bp.p, positions, n * 12, cudaMemcpyHostToDevice));
    launchTextureEval(s->dev, tex_root, (const float *)bp.p, n, (float *)bc.p, (float *)bf.p, flags, nullptr);
    CU(cudaGetLastError());
    CU(cudaDeviceSynchronize());
    if(rgb)
        CU(cudaMemcpy(rgb, bc.p, n * 12, cudaMemcpyDeviceToHost));
    if(f)
        CU(cudaMemcpy(f, bf.p, n * 4, cudaMemcpyDeviceToHost));
    return PTB_OK;
}

extern "C" int ptb200_sample_scatter(const float *normal, const float *mirror, float scatter, size_t n, uint64_t seed, uint32_t flags, float *dirs)
{
    if(!normal || !mirror || !dirs)
        return fail(PTB_EINVAL, "null argument");
    int nDev = 0;
    if(cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0)
        return fail(PTB_ENODEV, "no CUDA device: libptb200 has no CPU fallback");
    DeviceBuffer b;
    CU(b.ensure(n * 12));
    launchScatterSample(normal, mirror, scatter, n, seed, flags, (float *)b.p, nullptr);
    CU(cudaGetLastError());
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(dirs, b.p, n * 12, cudaMemcpyDeviceToHost));
    return PTB_OK;
}

extern "C" int ptb200_vector_eval(int op, const float *v, const float *normal, const float *ior, size_t n, float *out)
{
    if(op < 0 || op > 2 || (n && (!v || !normal || !out || (op != 0 && !ior))))
        return fail(PTB_EINVAL, "bad op or null argument");
    int nDev = 0;
    if(cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0)
        return fail(PTB_ENODEV, "no CUDA device: libptb200 has no CPU fallback");
    if(n == 0)
        return PTB_OK;
    DeviceBuffer bv, bn, bi, bo;
    const size_t outFloats = op == 2 ? n : 3 * n;
    CU(bv.ensure(n * 12));
    CU(bn.ensure(n * 12));
    CU(bi.ensure(n * 4));
    CU(bo.ensure(outFloats * 4));
    CU(cudaMemcpy(bv.p, v, n * 12, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(bn.p, normal, n * 12, cudaMemcpyHostToDevice));
    if(ior)
        CU(cudaMemcpy(bi.p, ior, n * 4, cudaMemcpyHostToDevice));
    launchVectorEval(op, (const float *)bv.p, (const float *)bn.p, (const float *)bi.p, n, (float *)bo.p, nullptr);
    CU(cudaGetLastError());
    CU(cudaMemcpy(out, bo.p, outFloats * 4, cudaMemcpyDeviceToHost));
    return PTB_OK;
}

extern "C" int ptb200_rng_words(int which, uint64_t seed, const uint32_t *pixel, const uint32_t *sample, const uint32_t *block, size_t n, uint32_t *out)
{
    if(which < 0 || which > 1 || (n && (!pixel || !sample || !out)))
        return fail(PTB_EINVAL, "bad generator or null argument");
    int nDev = 0;
    if(cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0)
        return fail(PTB_ENODEV, "no CUDA device: libptb200 has no CPU fallback");
    if(n == 0)
        return PTB_OK;
    const size_t words = which == 0 ? 2 * n : 4 * n;
    DeviceBuffer bp, bs, bb, bo;
    CU(bp.ensure(n * 4));
    CU(bs.ensure(n * 4));
    CU(bb.ensure(n * 4));
    CU(bo.ensure(words * 4));
    CU(cudaMemcpy(bp.p, pixel, n * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(bs.p, sample, n * 4, cudaMemcpyHostToDevice));
    if(block)
        CU(cudaMemcpy(bb.p, block, n * 4, cudaMemcpyHostToDevice));
    launchRngWords(which, seed, (const uint32_t *)bp.p, (const uint32_t *)bs.p, block ? (const uint32_t *)bb.p : nullptr, n, (uint32_t *)bo.p, nullptr);
    CU(cudaGetLastError());
    CU(cudaMemcpy(out, bo.p, words * 4, cudaMemcpyDeviceToHost));
    return PTB_OK;
}

extern "C" int ptb200_fp32_peak(double *fma_per_second, double *sm_clock_mhz)
{
    int nDev = 0, dev = 0;
    if(cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0)
        return fail(PTB_ENODEV, "no CUDA device: libptb200 has no CPU fallback");
    CU(cudaGetDevice(&dev));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, dev));
    cudaEvent_t a, b;
    CU(cudaEventCreate(&a));
    CU(cudaEventCreate(&b));
    int blocks = prop.multiProcessorCount * 8;
    launchFmaPeak(blocks, 256, 64, nullptr); // warm-up
    double best = 0;
    for(int rep = 0; rep < 5; rep++)
    {
        CU(cudaEventRecord(a));
        double fmas = launchFmaPeak(blocks, 256, 4096, nullptr);
        CU(cudaEventRecord(b));
        CU(cudaEventSynchronize(b));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, a, b));
        best = std::max(best, fmas / (ms * 1e-3));
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    if(fma_per_second)
        *fma_per_second = best;
    if(sm_clock_mhz)
        *sm_clock_mhz = best / ((double)prop.multiProcessorCount * 128.0) * 1e-6; // implied clock at 128 lanes/SM
    return PTB_OK;
}

// ---- wavefront render driver ------------------------------------------------------------------------
int ptb::checkParams(const PtbRenderParams *p, int &x0, int &y0, int &rw, int &rh)
{
    if(!p)
        return fail(PTB_EINVAL, "null render params");
    if(p->width <= 0 || p->height <= 0 || (long long)p->width * p->height > (1ll << 31) - 1)
        return fail(PTB_EINVAL, "bad image size");
    if(p->spp <= 0 || p->sample_stride <= 0 || p->sample_offset < 0 || p->max_depth < 0 || p->max_depth > 255)
        return fail(PTB_EINVAL, "bad sample count / stride / depth");
    x0 = p->x0, y0 = p->y0;
    int x1 = p->x1, y1 = p->y1;
    if(x0 == 0 && y0 == 0 && x1 == 0 && y1 == 0)
        x1 = p->width, y1 = p->height;
    if(x0 < 0 || y0 < 0 || x1 > p->width || y1 > p->height || x1 <= x0 || y1 <= y0)
        return fail(PTB_EINVAL, "bad pixel rectangle");
    rw = x1 - x0, rh = y1 - y0;
    return PTB_OK;
}

// Caller holds scene->lock and has made scene->device current.
int ptb::renderPasses(PtbScene *s, const PtbRenderParams *p, float4 *accum, PtbStats *stats, cudaStream_t stream, const PixelListArg *points)
{
    int x0, y0, rw, rh;
    if(int rc = checkParams(p, x0, y0, rw, rh))
        return rc;
    if(points)
        x0 = y0 = 0, rw = (int)points->count, rh = 1;
    RenderScratch &sx = s->scratch;
    CsgConfig cfg;
    if(int rc = chooseConfig(s, 2, sx.overflow, cfg))
        return rc;
    const size_t nPix = (size_t)rw * rh;
    int perPass = p->samples_per_pass;
    if(perPass <= 0)
    { // PTB_PATHS_IN_FLIGHT (default 128 Mi: 16 samples per pixel at 3840x2160; measured +11 % from 1 to 8
      // samples per pass and +2 % from 8 to 16 — the launches of a pass are amortised), but never more than a
      // quarter of what the device could give: the free memory plus what this scene's queues already hold
        static const size_t inFlight = [] {
            const char *e = getenv("PTB_PATHS_IN_FLIGHT");
            long long v = e ? atoll(e) : 0;
            return v > 0 ? (size_t)v : (size_t(128) << 20);
        }();
        size_t want = inFlight, freeB = 0, totalB = 0;
        if(cudaMemGetInfo(&freeB, &totalB) == cudaSuccess)
            want = std::min(want, std::max<size_t>(size_t(1) << 20, (freeB + sx.queues.bytes + sx.hits.bytes) / 4 / 136));
        perPass = (int)std::max<size_t>(1, want / nPix);
    }
    perPass = std::min(perPass, p->spp);
    if((size_t)perPass * nPix > 0xFF000000ull) // (the shade kernels add a grid stride of < 2^24 to a work index in 32 bits)
        return fail(PTB_EINVAL, "samples_per_pass x pixels exceeds the 32-bit queue index");
    // the last global sample index must fit the kernels' 32-bit sample counter
    if((long long)p->sample_offset + ((long long)p->spp - 1) * (long long)p->sample_stride > 0x7FFFFFFFll)
        return fail(PTB_EINVAL, "sample_offset + (spp - 1) * sample_stride exceeds INT32_MAX");
    size_t capacity = 0;
    for(;;)
    { // 136 B per path in flight (two 64 B queues + the 8 B hit record); a pass that does not fit is halved
        capacity = nPix * (size_t)perPass;
        if(ensureScratch(s, sx.queues, capacity * 64 * 2) == cudaSuccess && ensureScratch(s, sx.hits, capacity * sizeof(uint2)) == cudaSuccess)
            break;
        cudaGetLastError();
        if(perPass == 1)
            return fail(PTB_ENOMEM, "no device memory for the ray queues of even one sample per pixel per pass");
        perPass = (perPass + 1) / 2;
    }
    const int nPasses = (p->spp + perPass - 1) / perPass;
    const int stages = p->max_depth + 1; // primary + max_depth bounces
    uint2 *hits = (uint2 *)sx.hits.p;
    const int shadePerSm = shadeBlocksPerSm(true), shadePerSmBounce = shadeBlocksPerSm(false);
    if(shadePerSm <= 0 || shadePerSmBounce <= 0)
        return fail(PTB_ECUDA, "shade kernel does not fit on an SM");
    const int shadeBlocks = shadePerSm * s->smCount, shadeBlocksBounce = shadePerSmBounce * s->smCount;
    // Bounce levels from `tailLevel` on hold a few thousand rays at most (C5: 0.2 % of the camera rays at level
    // 3); launched as a trace / shade pair per level they cost ~12 us each for nothing, 26 launches per pass.
    // ONE tail kernel takes the queue of level tailLevel and follows every path to its end (trace + shade in a
    // loop, no queue).  PTB_TAIL_LEVEL: first bounce level the tail kernel owns (0 = off: a pair per level).
    static const int tailEnv = [] {
        const char *e = getenv("PTB_TAIL_LEVEL");
        return e ? atoi(e) : 4; // measured on C5 (865 ms frames): pairs only 853, tail from level 2: 909, 3: 865, 4: 845
    }();
    const int tailLevel = tailEnv > 0 && tailEnv <= p->max_depth && tailBlocksPerSm(!cfg.stageProgram, cfg.smemBytes) > 0 ? tailEnv : p->max_depth + 1;
    const int pairLevels = std::min(p->max_depth, tailLevel - 1); // bounce levels run as trace + shade pairs
    const bool hasTail = tailLevel <= p->max_depth;
    const int tailBlocks = hasTail ? tailBlocksPerSm(!cfg.stageProgram, cfg.smemBytes) * s->smCount : 0;
    CU(ensureScratch(s, sx.counters, (size_t)nPasses * (stages + 1) * sizeof(unsigned)));
    if(int rc = s->begin(stream))
        return rc;
    CU(cudaMemsetAsync(sx.counters.p, 0, (size_t)nPasses * (stages + 1) * sizeof(unsigned), stream));
    RayQueue q[2];
    for(int k = 0; k < 2; k++)
    {
        float4 *base = (float4 *)sx.queues.p + (size_t)k * capacity * 4;
        q[k].q0 = base, q[k].q1 = base + capacity, q[k].q2 = base + 2 * capacity, q[k].q3 = base + 3 * capacity;
    }
    const bool timed = stats != nullptr;
    const int evPerPass = 3 + 2 * pairLevels + (hasTail ? 1 : 0); // start, trace0, shade0, trace/shade of every pair level, tail
    size_t nEvents = timed ? (size_t)nPasses * evPerPass : 0;
    while(sx.events.size() < nEvents)
    {
        cudaEvent_t e;
        CU(cudaEventCreate(&e));
        sx.events.push_back(e);
    }
    RenderPass pass;
    pass.width = points ? points->pitch : p->width, pass.height = p->height;
    pass.pixelList = points ? points->list : nullptr;
    pass.x0 = x0, pass.y0 = y0, pass.rw = rw, pass.rh = rh;
    pass.sampleStride = p->sample_stride;
    pass.maxDepth = p->max_depth;
    pass.sw = p->screen_width, pass.sh = p->screen_height, pass.sd = p->screen_distance;
    pass.twoOverWidth = 2.f / (float)p->width, pass.twoOverHeight = 2.f / (float)p->height;
    pass.byPixels = FastDivisor::make((uint32_t)rw * (uint32_t)rh), pass.byRowWidth = FastDivisor::make((uint32_t)rw);
    pass.seed = p->seed;
    pass.flags = p->flags;
    uint64_t launches = 0;
    int specPasses = 0;
    // Queued (bounce) rays run the specialised kernels too.  (The first generation — pool, bit stack, one out-of-line
    // push-and-operate call per leaf some lane hit — lost on these incoherent rays, C5 211-231 ms against the
    // interpreter's 207-222, and kept them interpreted; the register-resident evaluator has no such call and no
    // warp votes: C5 bounce trace 196 -> 104 ms per frame, profiles/r02c_*.)  PTB_SPEC_BOUNCE=0 keeps them interpreted.
    static const int specBounceEnv = getenv("PTB_SPEC_BOUNCE") ? atoi(getenv("PTB_SPEC_BOUNCE")) : -1;
    static const bool specGen1 = getenv("PTB_SPEC_GEN") && atoi(getenv("PTB_SPEC_GEN")) == 1;
    const bool specBounce = specBounceEnv >= 0 ? specBounceEnv != 0 : (!specGen1 || (p->flags & PTB_FLAG_SPECIALIZE) != 0);
    for(int i = 0; i < nPasses; i++)
    {
        unsigned *c = (unsigned *)sx.counters.p + (size_t)i * (stages + 1);
        pass.samplesThisPass = std::min(perPass, p->spp - i * perPass);
        pass.bySamples = FastDivisor::make((uint32_t)pass.samplesThisPass);
        pass.sampleBase = (int)((long long)p->sample_offset + (long long)i * perPass * p->sample_stride);
        cudaEvent_t *ev = timed ? &sx.events[(size_t)i * evPerPass] : nullptr;
        if(timed)
            CU(cudaEventRecord(ev[0], stream));
        // the specialised kernels may arrive (background compile) between two passes
        SpecConfig sp;
        if(int rc = chooseSpec(s, p->flags, (double)nPix * p->spp, SPEC_TRACE, sp))
            return rc;
        DevScene devArg = s->dev;
        if(sp.k)
        {
            specPasses++;
            s->lastSpecCameraOnly = sp.cameraOnly;
            SPEC_LAUNCH(sp.k->tracePrimary, sp.cfg, stream, &devArg, &sp.cfg.cap, &sp.cfg.overflow, &sp.cfg.ovfStride, &pass, &hits);
        }
        else
            launchTracePrimary(s->dev, cfg, pass, hits, stream);
        if(timed)
            CU(cudaEventRecord(ev[1], stream));
        launchShade(s->dev, shadeBlocks, true, pass, accum, hits, q[0], nullptr, q[0], c, stream);
        if(timed)
            CU(cudaEventRecord(ev[2], stream));
        for(int b = 1; b <= pairLevels; b++)
        {
            if(sp.k && specBounce && !sp.cameraOnly)
            {
                RayQueue qin = q[(b - 1) & 1];
                const unsigned *cin = c + (b - 1);
                SPEC_LAUNCH(sp.k->traceQueue, sp.cfg, stream, &devArg, &sp.cfg.cap, &sp.cfg.overflow, &sp.cfg.ovfStride, &qin, &cin, &hits);
            }
            else
                launchTraceQueue(s->dev, cfg, q[(b - 1) & 1], c + (b - 1), hits, stream);
            if(timed)
                CU(cudaEventRecord(ev[1 + 2 * b], stream));
            launchShade(s->dev, shadeBlocksBounce, false, pass, accum, hits, q[(b - 1) & 1], c + (b - 1), q[b & 1], c + b, stream);
            if(timed)
                CU(cudaEventRecord(ev[2 + 2 * b], stream));
        }
        launches += 2 + 2 * (uint64_t)pairLevels;
        if(hasTail)
        { // rays of level tailLevel are in q[(tailLevel - 1) & 1], counted by c[tailLevel - 1]; rays the tail traces beyond those go to c[tailLevel]
            launchTail(s->dev, cfg, tailBlocks, pass, accum, q[(tailLevel - 1) & 1], c + (tailLevel - 1), c + tailLevel, stream);
            if(timed)
                CU(cudaEventRecord(ev[evPerPass - 1], stream));
            launches += 1;
        }
        CU(cudaGetLastError());
    }
    if(int rc = s->end(stream))
        return rc;
    if(!timed)
        return PTB_OK;
    size_t nCounters = (size_t)nPasses * (stages + 1);
    if(sx.hostCountersLen < nCounters)
    {
        if(sx.hostCounters)
            hostCountersPut(sx.hostCounters, sx.hostCountersLen);
        sx.hostCounters = hostCountersGet(nCounters, sx.hostCountersLen);
        if(!sx.hostCounters)
        {
            sx.hostCountersLen = 0;
            return fail(PTB_ENOMEM, "no page-locked host memory for the ray counters");
        }
    }
    CU(cudaMemcpyAsync(sx.hostCounters, sx.counters.p, nCounters * sizeof(unsigned), cudaMemcpyDeviceToHost, stream));
    CU(cudaStreamSynchronize(stream));
    s->lastUsePending = false;
    memset(stats, 0, sizeof(*stats));
    stats->launches = launches;
    stats->launches_elided = (uint64_t)nPasses * (2 * (uint64_t)(p->max_depth - pairLevels) - (hasTail ? 1 : 0));
    s->lastSpecPasses = specPasses, s->lastPasses = nPasses;
    for(int i = 0; i < nPasses; i++)
    {
        const unsigned *c = sx.hostCounters + (size_t)i * (stages + 1);
        uint64_t primary = nPix * (uint64_t)std::min(perPass, p->spp - i * perPass);
        stats->paths += primary;
        stats->rays_primary += primary;
        // algorithmic HBM traffic of the wavefront: 8 B hit record written + read per ray, 64 B per
        // queued ray written by shade and read back by shade, of which trace re-reads 32 B (o, d)
        stats->queue_bytes += 16ull * primary + 64ull * c[0];
        for(int b = 1; b <= pairLevels; b++)
        {
            stats->rays_bounce += c[b - 1];
            stats->queue_bytes += (32ull + 16ull + 64ull) * c[b - 1] + 64ull * c[b];
        }
        if(hasTail) // the tail reads its queue once (64 B per ray) and traces c[tailLevel - 1] + c[tailLevel] rays
            stats->rays_bounce += (uint64_t)c[tailLevel - 1] + c[tailLevel], stats->queue_bytes += 64ull * c[tailLevel - 1];
        cudaEvent_t *ev = &sx.events[(size_t)i * evPerPass];
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, ev[0], ev[1]));
        stats->ms_trace_primary += ms;
        CU(cudaEventElapsedTime(&ms, ev[0], ev[2]));
        stats->ms_primary += ms;
        CU(cudaEventElapsedTime(&ms, ev[1], ev[2]));
        stats->ms_shade_primary += ms;
        stats->rays_first_bounce += c[0];
        CU(cudaEventElapsedTime(&ms, ev[2], ev[evPerPass - 1]));
        stats->ms_bounce += ms;
        for(int b = 1; b <= pairLevels; b++)
        {
            CU(cudaEventElapsedTime(&ms, ev[2 * b], ev[1 + 2 * b]));
            stats->ms_trace_bounce += ms;
            CU(cudaEventElapsedTime(&ms, ev[1 + 2 * b], ev[2 + 2 * b]));
            stats->ms_shade_bounce += ms;
        }
    }
    stats->rays = stats->rays_primary + stats->rays_bounce;
    float total = 0;
    CU(cudaEventElapsedTime(&total, sx.events[0], sx.events[(size_t)nPasses * evPerPass - 1]));
    stats->ms_total = total;
    return PTB_OK;
}

extern "C" int ptb200_render_accumulate_device(PtbScene *s, const PtbRenderParams *p, void *d_accum4, PtbStats *stats, void *stream)
{
    if(!s || !d_accum4)
        return fail(PTB_EINVAL, "null argument");
    std::lock_guard<std::mutex> g(s->lock);
    if(s->asyncBusy)
        return fail(PTB_EBUSY, "an asynchronous render is running on this scene");
    CU(cudaSetDevice(s->device));
    return renderPasses(s, p, (float4 *)d_accum4, stats, (cudaStream_t)stream);
}

extern "C" int ptb200_resolve_device(const void *d_accum4, void *d_rgba4, size_t n_pixels, void *stream)
{
    if(!d_accum4 || !d_rgba4)
        return fail(PTB_EINVAL, "null argument");
    if(n_pixels == 0)
        return PTB_OK;
    launchResolve((const float4 *)d_accum4, (float4 *)d_rgba4, n_pixels, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return PTB_OK;
}

extern "C" int ptb200_render(PtbScene *s, const PtbRenderParams *p, float *rgba_out, PtbStats *stats)
{
    if(!s || !rgba_out)
        return fail(PTB_EINVAL, "null argument");
    int x0, y0, rw, rh;
    if(int rc = checkParams(p, x0, y0, rw, rh))
        return rc;
    std::lock_guard<std::mutex> g(s->lock);
    if(s->asyncBusy)
        return fail(PTB_EBUSY, "an asynchronous render is running on this scene");
    CU(cudaSetDevice(s->device));
    size_t nPix = (size_t)p->width * p->height;
    CU(ensureScratch(s, s->scratch.accum, nPix * 16));
    float4 *accum = (float4 *)s->scratch.accum.p;
    CU(cudaMemsetAsync(accum, 0, nPix * 16, nullptr));
    PtbStats local;
    if(int rc = renderPasses(s, p, accum, stats ? stats : &local, nullptr))
        return rc;
    cudaEvent_t a, b;
    CU(cudaEventCreate(&a));
    CU(cudaEventCreate(&b));
    CU(cudaEventRecord(a, nullptr));
    launchResolve(accum, accum, nPix, nullptr);
    CU(cudaEventRecord(b, nullptr));
    CU(cudaGetLastError());
    // only the rendered rectangle is written to the caller's buffer
    CU(cudaMemcpy2D(rgba_out + 4 * ((size_t)y0 * p->width + x0), (size_t)p->width * 16, accum + (size_t)y0 * p->width + x0, (size_t)p->width * 16, (size_t)rw * 16,
                    (size_t)rh, cudaMemcpyDeviceToHost));
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    if(stats)
    {
        stats->ms_resolve = ms;
        stats->launches += 1;
    }
    return PTB_OK;
}

// tracePixel for a LIST of pixels (reference include/path-trace.h:187-207, called per pixel at src/test.cpp:450):
// the wavefront runs over the listed pixels only; every pixel gets the samples, RNG streams and hence the value it
// has in a frame render with the same parameters.
extern "C" int ptb200_render_pixels(PtbScene *s, const PtbRenderParams *p, const int32_t *xy, size_t n, float *rgba_out, PtbStats *stats)
{
    if(!s || (n && (!xy || !rgba_out)))
        return fail(PTB_EINVAL, "null argument");
    int x0, y0, rw, rh;
    if(int rc = checkParams(p, x0, y0, rw, rh))
        return rc;
    if(p->width > 65535 || p->height > 65535 || n > 0x7FFFFFFFull)
        return fail(PTB_EINVAL, "image or list too large for a pixel list (16-bit coordinates)");
    if(stats)
        memset(stats, 0, sizeof(*stats));
    if(n == 0)
        return PTB_OK;
    std::vector<uint32_t> packed(n);
    for(size_t i = 0; i < n; i++)
    {
        const int x = xy[2 * i], y = xy[2 * i + 1];
        if(x < 0 || y < 0 || x >= p->width || y >= p->height)
            return fail(PTB_EINVAL, "listed pixel outside the image");
        packed[i] = ((uint32_t)y << 16) | (uint32_t)x;
    }
    std::lock_guard<std::mutex> g(s->lock);
    if(s->asyncBusy)
        return fail(PTB_EBUSY, "an asynchronous render is running on this scene");
    CU(cudaSetDevice(s->device));
    const size_t nPix = (size_t)p->width * p->height;
    CU(ensureScratch(s, s->scratch.accum, nPix * 16));
    DeviceBuffer list, out;
    CU(list.ensure(n * 4));
    CU(out.ensure(n * 16));
    float4 *accum = (float4 *)s->scratch.accum.p;
    CU(cudaMemcpy(list.p, packed.data(), n * 4, cudaMemcpyHostToDevice));
    CU(cudaMemsetAsync(accum, 0, nPix * 16, nullptr));
    PixelListArg arg{(const uint32_t *)list.p, (unsigned)n, p->width};
    PtbStats local;
    if(int rc = renderPasses(s, p, accum, stats ? stats : &local, nullptr, &arg))
        return rc;
    launchGatherPixels(accum, (const uint32_t *)list.p, (unsigned)n, p->width, (float4 *)out.p, nullptr);
    CU(cudaGetLastError());
    CU(cudaMemcpy(rgba_out, out.p, n * 16, cudaMemcpyDeviceToHost));
    if(stats)
        stats->launches += 1;
    return PTB_OK;
}

// Adaptive rendering: RenderBlock::renderSquare's quadtree (reference src/test.cpp:423-499) run
// level by level for all tiles at once.  Level `size` decides every square of that size (interpolate
// or subdivide), the lattice points the subdivided ones ask for are traced as ONE pixel-list render
// through the same wavefront kernels, then the next level follows.
extern "C" int ptb200_render_adaptive(PtbScene *s, const PtbRenderParams *p, const PtbAdaptiveParams *ap, float *rgba_out, uint8_t *source_out, float *lattice_out,
                                      PtbStats *stats, PtbAdaptiveStats *astats)
{
    if(!s || !rgba_out || !ap)
        return fail(PTB_EINVAL, "null argument");
    int x0, y0, rw, rh;
    if(int rc = checkParams(p, x0, y0, rw, rh))
        return rc;
    if(rw != p->width || rh != p->height)
        return fail(PTB_EINVAL, "adaptive rendering takes the whole image, not a rectangle");
    const int B = ap->block_size;
    if(B < 2 || B > 1024 || (B & (B - 1)) != 0)
        return fail(PTB_EINVAL, "block_size must be a power of two in [2, 1024] (getBlockSize, reference src/test.cpp:40-48)");
    if(ap->max_sample_size < 0 || !(ap->min_color_delta >= 0.f))
        return fail(PTB_EINVAL, "bad max_sample_size / min_color_delta");
    const int W = p->width, H = p->height;
    const int tilesX = (W + B - 1) / B, tilesY = (H + B - 1) / B;
    const int pitch = tilesX * B + 1, rows = tilesY * B + 1;
    if(pitch > 65535 || rows > 65535)
        return fail(PTB_EINVAL, "image too large for adaptive rendering (16-bit lattice coordinates)");
    std::lock_guard<std::mutex> g(s->lock);
    if(s->asyncBusy)
        return fail(PTB_EBUSY, "an asynchronous render is running on this scene");
    CU(cudaSetDevice(s->device));
    RenderScratch &sx = s->scratch;
    const size_t nLattice = (size_t)pitch * rows, nPix = (size_t)W * H;
    const size_t maxSquares = std::max<size_t>((size_t)tilesX * tilesY, (size_t)(tilesX * B / 2) * (size_t)(tilesY * B / 2));
    // carve-up of the adaptive block (all 16-byte aligned)
    auto align = [](size_t n) { return (n + 255) & ~(size_t)255; };
    const size_t oValue = 0, oInterp = oValue + align(nLattice * 16), oReq = oInterp + align(nLattice * 16), oSqA = oReq + align(nLattice * 4),
                 oSqB = oSqA + align(maxSquares * 4), oList = oSqB + align(maxSquares * 4), oCnt = oList + align(nLattice * 4), oOut = oCnt + 256,
                 oSrc = oOut + align(nPix * 16), oAlt = oSrc + align(nPix), oTemp = oAlt + align(nLattice * 4),
                 sortBytes = align(adaptiveSortList(nullptr, nullptr, (unsigned)nLattice, nullptr, 0, nullptr)), total = oTemp + sortBytes;
    CU(ensureScratch(s, sx.adaptive, total));
    CU(ensureScratch(s, sx.accum, nLattice * 16));
    char *base = (char *)sx.adaptive.p;
    float4 *accum = (float4 *)sx.accum.p, *value = (float4 *)(base + oValue), *interp = (float4 *)(base + oInterp), *out = (float4 *)(base + oOut);
    unsigned *requested = (unsigned *)(base + oReq), *counters = (unsigned *)(base + oCnt);
    uint32_t *squares[2] = {(uint32_t *)(base + oSqA), (uint32_t *)(base + oSqB)}, *list = (uint32_t *)(base + oList);
    uint8_t *source = (uint8_t *)(base + oSrc);
    CU(cudaMemsetAsync(accum, 0, nLattice * 16, nullptr));
    CU(cudaMemsetAsync(base, 0, oSqA, nullptr)); // value, interp, request flags
    // level 0: the corners of every tile (src/test.cpp:504), traced once for the tiles that share them
    std::vector<uint32_t> host;
    for(int j = 0; j <= tilesY; j++)
        for(int i = 0; i <= tilesX; i++)
            host.push_back(((uint32_t)(j * B) << 16) | (uint32_t)(i * B));
    unsigned nList = (unsigned)host.size();
    CU(cudaMemcpyAsync(list, host.data(), host.size() * 4, cudaMemcpyHostToDevice, nullptr));
    CU(cudaStreamSynchronize(nullptr));
    host.clear();
    for(int j = 0; j < tilesY; j++)
        for(int i = 0; i < tilesX; i++)
            host.push_back(((uint32_t)(j * B) << 16) | (uint32_t)(i * B));
    unsigned nSquares = (unsigned)host.size();
    CU(cudaMemcpyAsync(squares[0], host.data(), host.size() * 4, cudaMemcpyHostToDevice, nullptr));
    CU(cudaStreamSynchronize(nullptr));
    PtbStats sum;
    memset(&sum, 0, sizeof(sum));
    PtbAdaptiveStats as;
    memset(&as, 0, sizeof(as));
    auto traceList = [&](unsigned n) -> int {
        if(n == 0u)
            return PTB_OK;
        PixelListArg arg{list, n, pitch};
        PtbStats st;
        if(int rc = renderPasses(s, p, accum, &st, nullptr, &arg))
            return rc;
        launchAdaptiveResolveList(accum, value, list, n, pitch, nullptr);
        CU(cudaGetLastError());
        sum.paths += st.paths, sum.rays += st.rays, sum.launches += st.launches + 1, sum.queue_bytes += st.queue_bytes;
        sum.ms_total += st.ms_total, sum.ms_primary += st.ms_primary, sum.ms_bounce += st.ms_bounce;
        sum.rays_primary += st.rays_primary, sum.rays_bounce += st.rays_bounce;
        sum.ms_trace_primary += st.ms_trace_primary, sum.ms_trace_bounce += st.ms_trace_bounce;
        sum.ms_shade_primary += st.ms_shade_primary, sum.ms_shade_bounce += st.ms_shade_bounce;
        sum.rays_first_bounce += st.rays_first_bounce, sum.launches_elided += st.launches_elided;
        as.pixels_traced += n;
        return PTB_OK;
    };
    if(int rc = traceList(nList))
        return rc;
    int cur = 0;
    for(int size = B; size >= 2 && nSquares > 0u; size /= 2)
    {
        CU(cudaMemsetAsync(counters, 0, 4 * sizeof(unsigned), nullptr));
        AdaptiveLevel lv;
        lv.width = W, lv.height = H, lv.pitch = pitch, lv.size = size, lv.maxSampleSize = ap->max_sample_size;
        lv.delta2 = ap->min_color_delta * ap->min_color_delta;
        lv.squares = squares[cur], lv.nSquares = nSquares;
        lv.value = value, lv.interp = interp, lv.requested = requested;
        lv.childSquares = squares[cur ^ 1], lv.pixelList = list, lv.counters = counters;
        launchAdaptiveDecide(lv, nullptr);
        CU(cudaGetLastError());
        unsigned c[4] = {0u, 0u, 0u, 0u};
        CU(cudaMemcpy(c, counters, sizeof(c), cudaMemcpyDeviceToHost));
        sum.launches += 1;
        as.levels += 1;
        as.squares_interpolated += c[2];
        as.pixels_interpolated += c[3];
        nSquares = c[0], nList = c[1], cur ^= 1;
        adaptiveSortList(list, (uint32_t *)(base + oAlt), nList, base + oTemp, sortBytes, nullptr); // back into image order: coherent camera rays
        CU(cudaGetLastError());
        if(int rc = traceList(nList))
            return rc;
    }
    launchAdaptiveAssemble(value, interp, pitch, W, H, out, source, nullptr);
    CU(cudaGetLastError());
    sum.launches += 1;
    CU(cudaMemcpy(rgba_out, out, nPix * 16, cudaMemcpyDeviceToHost));
    if(source_out)
        CU(cudaMemcpy(source_out, source, nPix, cudaMemcpyDeviceToHost));
    if(lattice_out)
        CU(cudaMemcpy(lattice_out, value, nLattice * 16, cudaMemcpyDeviceToHost));
    if(stats)
        *stats = sum;
    if(astats)
        *astats = as;
    return PTB_OK;
}

extern "C" int ptb200_display8_device(const void *d_rgba4, void *d_rgbx8, size_t n_pixels, int divide_by_alpha, void *stream)
{
    if(!d_rgba4 || !d_rgbx8)
        return fail(PTB_EINVAL, "null argument");
    if(n_pixels == 0)
        return PTB_OK;
    launchDisplay8((const float4 *)d_rgba4, (uchar4 *)d_rgbx8, n_pixels, divide_by_alpha != 0, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return PTB_OK;
}

extern "C" int ptb200_display8(const float *rgba, size_t n_pixels, int divide_by_alpha, uint8_t *rgbx8)
{
    if(!rgba || !rgbx8)
        return fail(PTB_EINVAL, "null argument");
    if(n_pixels == 0)
        return PTB_OK;
    int nDev = 0;
    if(cudaGetDeviceCount(&nDev) != cudaSuccess || nDev == 0)
        return fail(PTB_ENODEV, "no CUDA device: libptb200 has no CPU fallback");
    DeviceBuffer in, out;
    CU(in.ensure(n_pixels * 16));
    CU(out.ensure(n_pixels * 4));
    CU(cudaMemcpy(in.p, rgba, n_pixels * 16, cudaMemcpyHostToDevice));
    launchDisplay8((const float4 *)in.p, (uchar4 *)out.p, n_pixels, divide_by_alpha != 0, nullptr);
    CU(cudaGetLastError());
    CU(cudaMemcpy(rgbx8, out.p, n_pixels * 4, cudaMemcpyDeviceToHost));
    return PTB_OK;
}

// ---- asynchronous / progressive boundary (BlockRenderer::done / copyToBuffer) -----------------------
static void jobMain(PtbJob *job)
{
    PtbScene *s = job->scene;
    {
        std::lock_guard<std::mutex> g(s->lock);
        cudaSetDevice(s->device);
        cudaStream_t stream = nullptr;
        int rc = cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking) == cudaSuccess ? PTB_OK : PTB_ECUDA;
        if(rc == PTB_OK)
            rc = renderPasses(s, &job->params, job->accum, &job->stats, stream);
        if(rc != PTB_OK)
            job->error = rc == PTB_ECUDA && lastError().empty() ? "cudaStreamCreate failed" : lastError();
        if(stream)
        {
            cudaStreamSynchronize(stream);
            cudaStreamDestroy(stream);
        }
        job->status = rc;
        s->asyncBusy = false;
    }
    __sync_synchronize();
    job->done = 1;
}

extern "C" int ptb200_render_async(PtbScene *s, const PtbRenderParams *p, PtbJob **out)
{
    if(!s || !out)
        return fail(PTB_EINVAL, "null argument");
    *out = nullptr;
    int x0, y0, rw, rh;
    if(int rc = checkParams(p, x0, y0, rw, rh))
        return rc;
    PtbJob *job = nullptr;
    {
        std::lock_guard<std::mutex> g(s->lock);
        if(s->asyncBusy)
            return fail(PTB_EBUSY, "an asynchronous render is already running on this scene");
        CU(cudaSetDevice(s->device));
        job = new PtbJob;
        job->scene = s;
        job->params = *p;
        size_t nPix = (size_t)p->width * p->height;
        if(cudaMalloc((void **)&job->accum, nPix * 16) != cudaSuccess || cudaMemset(job->accum, 0, nPix * 16) != cudaSuccess ||
           cudaStreamCreateWithFlags(&job->readStream, cudaStreamNonBlocking) != cudaSuccess)
        {
            if(job->accum)
                cudaFree(job->accum);
            delete job;
            return fail(PTB_ECUDA, "could not allocate the job's accumulation buffer");
        }
        s->asyncBusy = true;
    }
    job->worker = std::thread(jobMain, job);
    *out = job;
    return PTB_OK;
}

extern "C" int ptb200_job_done(PtbJob *job)
{
    if(!job)
        return fail(PTB_EINVAL, "null job");
    if(!job->done)
        return 0;
    if(job->status != PTB_OK)
        return fail(job->status, job->error);
    return 1;
}

extern "C" int ptb200_job_read(PtbJob *job, float *rgba, int width, int height)
{
    if(!job || !rgba)
        return fail(PTB_EINVAL, "null argument");
    if(width != job->params.width || height != job->params.height)
        return fail(PTB_EINVAL, "buffer size differs from the job's image size");
    CU(cudaSetDevice(job->scene->device));
    size_t nPix = (size_t)width * height;
    job->staging.resize(nPix * 4);
    CU(cudaMemcpyAsync(job->staging.data(), job->accum, nPix * 16, cudaMemcpyDeviceToHost, job->readStream));
    CU(cudaStreamSynchronize(job->readStream));
    const float *a = job->staging.data();
    for(size_t i = 0; i < nPix; i++)
        if(a[4 * i + 3] > 0.f)
        { // only valid pixels are written, as RenderBlock::copyToBuffer (src/test.cpp:359-371)
            float inv = 1.f / a[4 * i + 3];
            rgba[4 * i] = a[4 * i] * inv, rgba[4 * i + 1] = a[4 * i + 1] * inv, rgba[4 * i + 2] = a[4 * i + 2] * inv, rgba[4 * i + 3] = 1.f;
        }
    return PTB_OK;
}

extern "C" int ptb200_job_wait(PtbJob *job, PtbStats *stats)
{
    if(!job)
        return fail(PTB_EINVAL, "null job");
    if(job->worker.joinable())
        job->worker.join();
    if(stats)
        *stats = job->stats;
    if(job->status != PTB_OK)
        return fail(job->status, job->error);
    return PTB_OK;
}

extern "C" void ptb200_job_destroy(PtbJob *job)
{
    if(!job)
        return;
    if(job->worker.joinable())
        job->worker.join();
    cudaSetDevice(job->scene->device);
    if(job->readStream)
        cudaStreamDestroy(job->readStream);
    if(job->accum)
        cudaFree(job->accum);
    delete job;
}
