// ptb_multi.cu — sample-sharded multi-GPU rendering INSIDE the library.
//
// The reference's counterpart of this seam is its tile farm: a C++ main hands tiles to render servers over
// TCP and pastes the returned pixels into one frame (reference src/test.cpp:520-630 client, :681-778
// server).  Here the unit of distribution is the SAMPLE, not the tile: rank g of G renders the global
// camera-sample indices g, g+G, g+2G, ... of every pixel (the RNG is keyed by the global index, so the union
// over ranks is exactly the one-GPU sample set), and ONE exchange step follows — an NCCL reduce-scatter of
// the accumulation frame by row blocks over NVLink, after which every rank divides its own rows by the sample
// count and writes them where the caller wants them (its rows of a host image: G PCIe links instead of one).
//
// Two ways in, both C ABI, no Python needed:
//   ptb200_comm_init_all   one process drives n devices (ncclCommInitAll, one host thread + stream per device
//                          per frame) — what a C++ main that replaces the reference's client loop calls;
//   ptb200_comm_init_rank  one process per GPU (torchrun, MPI): the caller ships the 128-byte id.
// libnccl.so.2 is dlopen()ed on first use (like libnvrtc), so libptb200.so loads without it.
#include <dlfcn.h>
#include <string.h>

#include <algorithm>
#include <thread>

#include "ptb_internal.h"

using namespace ptb;

namespace
{
// ---- the NCCL entry points used (declared by hand: no build-time dependency on nccl.h) ------------------
struct NcclId
{
    char bytes[PTB_COMM_ID_BYTES];
};
typedef int (*ncclGetUniqueId_t)(NcclId *);
typedef int (*ncclCommInitRank_t)(void **, int, NcclId, int);
typedef int (*ncclCommInitAll_t)(void **, int, const int *);
typedef int (*ncclCommDestroy_t)(void *);
typedef int (*ncclReduceScatter_t)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef int (*ncclBroadcast_t)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef int (*ncclGroup_t)(void);
typedef const char *(*ncclGetErrorString_t)(int);
enum
{
    kNcclChar = 0,  // ncclInt8
    kNcclFloat = 7, // ncclFloat32
    kNcclSum = 0
};
struct Nccl
{
    bool tried = false, ok = false;
    std::string why;
    ncclGetUniqueId_t GetUniqueId;
    ncclCommInitRank_t CommInitRank;
    ncclCommInitAll_t CommInitAll;
    ncclCommDestroy_t CommDestroy;
    ncclReduceScatter_t ReduceScatter;
    ncclBroadcast_t Broadcast;
    ncclGroup_t GroupStart, GroupEnd;
    ncclGetErrorString_t GetErrorString;
};
Nccl g_nccl;
std::mutex g_ncclLock;

bool loadNccl()
{
    std::lock_guard<std::mutex> g(g_ncclLock);
    if(g_nccl.tried)
        return g_nccl.ok;
    g_nccl.tried = true;
    void *lib = nullptr;
    for(const char *name : {"libnccl.so.2", "libnccl.so"})
        if((lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL)))
            break;
    if(!lib)
    {
        g_nccl.why = "libnccl.so.2 not found";
        return false;
    }
#define SYM(field, name)                                                                 \
    if(!(g_nccl.field = (decltype(g_nccl.field))dlsym(lib, name)))                       \
    {                                                                                    \
        g_nccl.why = std::string("libnccl: missing symbol ") + name;                     \
        return false;                                                                    \
    }
    SYM(GetUniqueId, "ncclGetUniqueId")
    SYM(CommInitRank, "ncclCommInitRank")
    SYM(CommInitAll, "ncclCommInitAll")
    SYM(CommDestroy, "ncclCommDestroy")
    SYM(ReduceScatter, "ncclReduceScatter")
    SYM(Broadcast, "ncclBroadcast")
    SYM(GroupStart, "ncclGroupStart")
    SYM(GroupEnd, "ncclGroupEnd")
    SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
    g_nccl.ok = true;
    return true;
}
#define NC(call)                                                                                              \
    do                                                                                                        \
    {                                                                                                         \
        int e_ = (call);                                                                                      \
        if(e_ != 0)                                                                                           \
            return fail(PTB_ECUDA, std::string(#call) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(e_) : "NCCL error")); \
    } while(0)
} // namespace

struct PtbComm
{
    int nranks = 0;
    bool ownsDevices = false;          // init_all: this process drives every rank, each on its own stream
    std::vector<int> devices;          // local members: CUDA device ...
    std::vector<int> ranks;            // ... global rank ...
    std::vector<void *> comms;         // ... ncclComm_t ...
    std::vector<cudaStream_t> streams; // ... and (init_all) stream
};

// ---- pure partition rules (no GPU, no NCCL: tested on CPU) --------------------------------------------------
// Rows: equal blocks of ceil(height / nranks) rows (a reduce-scatter needs equal counts; the frame is padded to
// nranks blocks on the device); rank r owns [r * block, min(height, (r + 1) * block)).
extern "C" int ptb200_multi_rows(int height, int nranks, int rank, int *row0, int *row1)
{
    if(height <= 0 || nranks <= 0 || rank < 0 || rank >= nranks || !row0 || !row1)
        return fail(PTB_EINVAL, "bad height / rank / nranks");
    const int block = (height + nranks - 1) / nranks;
    *row0 = std::min(height, rank * block);
    *row1 = std::min(height, (rank + 1) * block);
    return PTB_OK;
}
// Samples: rank r renders the global sample indices r, r + nranks, ... below spp (ragged when nranks does not
// divide spp: the first spp % nranks ranks take one more).
extern "C" int ptb200_multi_shard(int spp, int nranks, int rank, int *count, int *offset, int *stride)
{
    if(spp < 0 || nranks <= 0 || rank < 0 || rank >= nranks || !count || !offset || !stride)
        return fail(PTB_EINVAL, "bad spp / rank / nranks");
    *count = spp > rank ? (spp - rank + nranks - 1) / nranks : 0;
    *offset = rank;
    *stride = nranks;
    return PTB_OK;
}

// ---- communicators ----------------------------------------------------------------------------------------------
extern "C" int ptb200_comm_unique_id(void *id)
{
    if(!id)
        return fail(PTB_EINVAL, "null id");
    if(!loadNccl())
        return fail(PTB_ENODEV, "NCCL unavailable: " + g_nccl.why);
    NcclId nid;
    NC(g_nccl.GetUniqueId(&nid));
    memcpy(id, nid.bytes, sizeof(nid.bytes));
    return PTB_OK;
}

extern "C" int ptb200_comm_init_rank(const void *id, int rank, int nranks, PtbComm **out)
{
    if(!id || !out || nranks <= 0 || rank < 0 || rank >= nranks)
        return fail(PTB_EINVAL, "bad id / rank / nranks");
    *out = nullptr;
    if(!loadNccl())
        return fail(PTB_ENODEV, "NCCL unavailable: " + g_nccl.why);
    int device = 0;
    CU(cudaGetDevice(&device));
    NcclId nid;
    memcpy(nid.bytes, id, sizeof(nid.bytes));
    void *comm = nullptr;
    NC(g_nccl.CommInitRank(&comm, nranks, nid, rank));
    PtbComm *c = new PtbComm;
    c->nranks = nranks;
    c->devices.push_back(device), c->ranks.push_back(rank), c->comms.push_back(comm), c->streams.push_back(nullptr);
    *out = c;
    return PTB_OK;
}

extern "C" int ptb200_comm_init_all(const int *devices, int n, PtbComm **out)
{
    if(!out || n <= 0)
        return fail(PTB_EINVAL, "bad device list");
    *out = nullptr;
    if(!loadNccl())
        return fail(PTB_ENODEV, "NCCL unavailable: " + g_nccl.why);
    int have = 0;
    CU(cudaGetDeviceCount(&have));
    std::vector<int> devs(n);
    for(int i = 0; i < n; i++)
    {
        devs[i] = devices ? devices[i] : i;
        if(devs[i] < 0 || devs[i] >= have)
            return fail(PTB_EINVAL, "device index out of range");
    }
    std::vector<void *> comms(n, nullptr);
    NC(g_nccl.CommInitAll(comms.data(), n, devs.data()));
    PtbComm *c = new PtbComm;
    c->nranks = n, c->ownsDevices = true;
    int current = 0;
    cudaGetDevice(&current);
    for(int i = 0; i < n; i++)
    {
        cudaStream_t st = nullptr;
        cudaSetDevice(devs[i]);
        cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
        c->devices.push_back(devs[i]), c->ranks.push_back(i), c->comms.push_back(comms[i]), c->streams.push_back(st);
    }
    cudaSetDevice(current);
    *out = c;
    return PTB_OK;
}

extern "C" int ptb200_comm_info(const PtbComm *c, int *nranks, int *nlocal, int *first_rank)
{
    if(!c)
        return fail(PTB_EINVAL, "null communicator");
    if(nranks)
        *nranks = c->nranks;
    if(nlocal)
        *nlocal = (int)c->devices.size();
    if(first_rank)
        *first_rank = c->ranks[0];
    return PTB_OK;
}

extern "C" void ptb200_comm_destroy(PtbComm *c)
{
    if(!c)
        return;
    int current = 0;
    cudaGetDevice(&current);
    for(size_t i = 0; i < c->comms.size(); i++)
    {
        cudaSetDevice(c->devices[i]);
        if(c->streams[i])
        {
            cudaStreamSynchronize(c->streams[i]);
            cudaStreamDestroy(c->streams[i]);
        }
        if(g_nccl.ok && c->comms[i])
            g_nccl.CommDestroy(c->comms[i]);
    }
    cudaSetDevice(current);
    delete c;
}

// ---- scene replicated on every member ------------------------------------------------------------------------
// The image pixels cross PCIe ONCE (rank 0, through pinned staging) and reach the other GPUs by ncclBroadcast over
// NVLink: N ranks uploading the same 87 MB at the same time (one-process-per-GPU e2e pattern) cost ~21 ms of PCIe
// contention per scene.  Everything else in the description is a few KB and is uploaded by every rank.
extern "C" int ptb200_scene_create_multi(PtbComm *c, const PtbSceneDesc *desc, PtbScene **scenes)
{
    if(!c || !desc || !scenes)
        return fail(PTB_EINVAL, "null argument");
    const int nlocal = (int)c->devices.size();
    for(int m = 0; m < nlocal; m++)
        scenes[m] = nullptr;
    if(desc->n_images < 0 || (desc->n_images > 0 && !desc->images))
        return fail(PTB_EINVAL, "bad image table");
    int current = 0;
    cudaGetDevice(&current);
    std::vector<std::vector<DeviceBuffer>> linear(nlocal);
    std::vector<std::vector<ImageSource>> sources(nlocal);
    int rc = PTB_OK;
    auto body = [&]() -> int {
        for(int m = 0; m < nlocal; m++)
        {
            linear[m] = std::vector<DeviceBuffer>(desc->n_images);
            sources[m].resize(desc->n_images);
        }
        for(int i = 0; i < desc->n_images; i++)
        {
            const PtbImageDesc &im = desc->images[i];
            if(im.width <= 0 || im.height <= 0)
                return fail(PTB_EINVAL, "image with no pixels");
            const size_t bytes = (size_t)im.width * im.height * 16;
            for(int m = 0; m < nlocal; m++)
            {
                CU(cudaSetDevice(c->devices[m]));
                CU(linear[m][i].ensure(bytes));
                sources[m][i].device = (const float4 *)linear[m][i].p;
                if(c->ranks[m] == 0)
                {
                    if(!im.rgba)
                        return fail(PTB_EINVAL, "rank 0 must supply the image pixels");
                    if(int r = uploadLinear(linear[m][i].p, im.rgba, bytes))
                        return r;
                }
            }
            NC(g_nccl.GroupStart());
            for(int m = 0; m < nlocal; m++)
            {
                CU(cudaSetDevice(c->devices[m]));
                NC(g_nccl.Broadcast(linear[m][i].p, linear[m][i].p, bytes, kNcclChar, 0, c->comms[m], c->streams[m]));
            }
            NC(g_nccl.GroupEnd());
        }
        for(int m = 0; m < nlocal; m++)
        {
            CU(cudaSetDevice(c->devices[m]));
            CU(cudaStreamSynchronize(c->streams[m]));
            PtbScene *s = new PtbScene;
            scenes[m] = s;
            if(int r = sceneCreateImpl(desc, s, nullptr, &sources[m]))
                return r;
            CU(cudaDeviceSynchronize()); // the array copies read the linear buffers, which go back to the cache below
        }
        return PTB_OK;
    };
    try
    {
        rc = body();
    }
    catch(std::bad_alloc &)
    {
        rc = fail(PTB_ENOMEM, "out of host memory");
    }
    if(rc != PTB_OK)
    {
        std::string keep = lastError();
        for(int m = 0; m < nlocal; m++)
            if(scenes[m])
            {
                ptb200_scene_destroy(scenes[m]);
                scenes[m] = nullptr;
            }
        setLastError(keep);
    }
    cudaSetDevice(current);
    return rc;
}

// ---- one frame ------------------------------------------------------------------------------------------------
namespace
{
struct MemberResult
{
    int status = PTB_OK;
    std::string error;
};

// Everything one member (rank) does for a frame, on `stream`.
int renderMember(PtbComm *c, int m, PtbScene *s, const PtbRenderParams *p, float *rgbaHost, float4 *rgbaDevice, PtbStats *stats, cudaStream_t stream)
{
    const int rank = c->ranks[m], G = c->nranks;
    CU(cudaSetDevice(c->devices[m]));
    if(s->device != c->devices[m])
        return fail(PTB_EINVAL, "scene lives on another device than its communicator member");
    const int W = p->width, H = p->height;
    const int block = (H + G - 1) / G;
    const size_t blockPixels = (size_t)block * W;
    int row0 = 0, row1 = 0;
    ptb200_multi_rows(H, G, rank, &row0, &row1);
    PtbRenderParams mine = *p;
    int count = 0, offset = 0, stride = 1;
    ptb200_multi_shard(p->spp, G, rank, &count, &offset, &stride);
    mine.spp = count;
    mine.sample_offset = p->sample_offset + offset * p->sample_stride;
    mine.sample_stride = p->sample_stride * stride;
    std::lock_guard<std::mutex> g(s->lock);
    if(s->asyncBusy)
        return fail(PTB_EBUSY, "an asynchronous render is running on this scene");
    if(blockPixels * G * 16 > s->scratch.multi.bytes)
        s->quiesce();
    CU(s->scratch.multi.ensure(blockPixels * G * 16)); // the frame padded to G equal row blocks
    float4 *accum = (float4 *)s->scratch.multi.p;
    if(int rc = s->begin(stream))
        return rc;
    CU(cudaMemsetAsync(accum, 0, blockPixels * G * 16, stream));
    if(stats)
        memset(stats, 0, sizeof(*stats));
    if(count > 0)
        if(int rc = renderPasses(s, &mine, accum, stats, stream))
            return rc;
    // the ONE exchange step: sum over ranks, rank r keeps row block r (in place)
    float4 *mineBlock = accum + (size_t)rank * blockPixels;
    NC(g_nccl.ReduceScatter(accum, mineBlock, blockPixels * 4, kNcclFloat, kNcclSum, c->comms[m], stream));
    const size_t ownPixels = (size_t)(row1 - row0) * W;
    if(ownPixels)
    {
        launchResolve(mineBlock, mineBlock, ownPixels, stream);
        CU(cudaGetLastError());
        if(rgbaDevice)
            CU(cudaMemcpyAsync(rgbaDevice + (size_t)row0 * W, mineBlock, ownPixels * 16, cudaMemcpyDeviceToDevice, stream));
        if(rgbaHost)
            CU(cudaMemcpyAsync(rgbaHost + 4 * (size_t)row0 * W, mineBlock, ownPixels * 16, cudaMemcpyDeviceToHost, stream));
    }
    if(int rc = s->end(stream))
        return rc;
    if(stats)
        stats->launches += 1;
    if(rgbaHost || stats || c->ownsDevices)
    {
        CU(cudaStreamSynchronize(stream));
        s->lastUsePending = false;
    }
    return PTB_OK;
}
} // namespace

extern "C" int ptb200_render_multi(PtbComm *c, PtbScene *const *scenes, const PtbRenderParams *p, float *rgba_out, void *const *d_rgba_out, PtbStats *stats, void *stream)
{
    if(!c || !scenes || !p)
        return fail(PTB_EINVAL, "null argument");
    int x0, y0, rw, rh;
    if(int rc = checkParams(p, x0, y0, rw, rh))
        return rc;
    if(rw != p->width || rh != p->height)
        return fail(PTB_EINVAL, "ptb200_render_multi renders whole frames (the exchange step moves whole row blocks)");
    const int nlocal = (int)c->devices.size();
    for(int m = 0; m < nlocal; m++)
        if(!scenes[m])
            return fail(PTB_EINVAL, "null scene");
    if(!c->ownsDevices)
        return renderMember(c, 0, scenes[0], p, rgba_out, d_rgba_out ? (float4 *)d_rgba_out[0] : nullptr, stats, (cudaStream_t)stream);
    // one process, n devices: one host thread + stream per device for this frame; synchronous
    int current = 0;
    cudaGetDevice(&current);
    std::vector<MemberResult> res(nlocal);
    std::vector<std::thread> threads;
    for(int m = 0; m < nlocal; m++)
        threads.emplace_back([&, m] {
            res[m].status = renderMember(c, m, scenes[m], p, rgba_out, d_rgba_out ? (float4 *)d_rgba_out[m] : nullptr, stats ? stats + m : nullptr, c->streams[m]);
            if(res[m].status != PTB_OK)
                res[m].error = lastError();
        });
    for(std::thread &t : threads)
        t.join();
    cudaSetDevice(current);
    for(int m = 0; m < nlocal; m++)
        if(res[m].status != PTB_OK)
            return fail(res[m].status, "device " + std::to_string(c->devices[m]) + ": " + res[m].error);
    return PTB_OK;
}
