/* ptb200.h — C ABI of the B200-native render hot path for programmerjake/path-trace.
 *
 * This is the drop-in boundary.  The reference has no FFI: the seam its
 * render path sits behind is the C++ virtual interface
 *     Object::makeSpanIterator()            (reference include/object.h:10-24)
 *     SpanIterator::{init,next,isAtEnd,*}   (reference include/span.h:129-171)
 *     traceRay / tracePixel                 (reference include/path-trace.h:58-207)
 * driven by one caller, RenderBlock / BlockRenderer (reference src/test.cpp:310-518).
 * The host C++ scene classes (path-trace_b200/host/ptb200_scene.hpp: same class names and
 * constructors as the reference) flatten an Object graph into the plain arrays below and call
 * these entry points.  Nothing here throws; every function returns 0 on success or a negative
 * PTB_E* code, with ptb200_last_error() holding a thread-local message.  No torch / C++ types
 * cross this boundary: plain pointers and sizes only.
 *
 * All float arithmetic on the device is IEEE binary32, like the reference.
 */
#ifndef PTB200_H_INCLUDED
#define PTB200_H_INCLUDED

#ifdef __CUDACC_RTC__ /* compiled at run time by NVRTC (scene-specialised kernels): no system headers there */
typedef unsigned char uint8_t;
typedef unsigned short uint16_t;
typedef int int32_t;
typedef unsigned int uint32_t;
typedef long long int64_t;
typedef unsigned long long uint64_t;
#else
#include <stddef.h>
#include <stdint.h>
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define PTB200_ABI_VERSION 2

/* ---- status codes -------------------------------------------------------------------------- */
enum {
    PTB_OK = 0,
    PTB_EINVAL = -1,  /* bad argument / malformed scene program                                  */
    PTB_ECUDA = -2,   /* CUDA runtime error (message has cudaGetErrorString)                     */
    PTB_ENOMEM = -3,  /* host or device allocation failed                                        */
    PTB_ENODEV = -4,  /* no CUDA device: the product path has no CPU fallback and fails loudly   */
    PTB_EBUSY = -5    /* an asynchronous render is still running on this scene                   */
};

/* ---- flattened scene ("linear SoA CSG program") -------------------------------------------- */

/* primitive kinds: reference src/sphere.cpp:31-49, src/plane.cpp:35-63 */
enum { PTB_PRIM_SPHERE = 0, PTB_PRIM_PLANE = 1 };

typedef struct PtbPrim {
    float p[4];       /* sphere: centre.xyz, r*r (float product, reference src/sphere.cpp:10)      */
                      /* plane : normal.xyz (as given, NOT normalised), d                          */
    int32_t type;     /* PTB_PRIM_*                                                                */
    int32_t material; /* index into materials[]                                                    */
    int32_t chain;    /* index into chains[] = stack of TransformedObject wrappers, -1 = none      */
    int32_t reserved;
} PtbPrim;

/* One TransformedObject (reference include/object.h:26-98).  Matrix element order is the
 * reference's Matrix member order x00,x10,x20,x30, x01,x11,x21,x31, x02,x12,x22,x32
 * (include/transform.h:19-21).  m maps the parent-space ray into the child's space
 * (include/transform.h:429-432); inv = m.inverse() maps child normals back
 * (include/span.h:122-127). */
typedef struct PtbXform {
    float m[12];
    float inv[12];
} PtbXform;

/* chain_xf[first .. first+count) lists xform indices, OUTERMOST wrapper first. */
typedef struct PtbChain {
    int32_t first;
    int32_t count;
} PtbChain;

/* Postfix CSG program, one uint32 per op: low 8 bits = kind, high 24 bits = argument.
 * PUSH_PRIM pushes the (0- or 1-span) list of primitive <arg>; UNION / DIFFERENCE /
 * INTERSECTION pop list B then list A and push the merge, using exactly the streaming merge
 * rules of reference src/union.cpp:84-134, src/difference.cpp:84-135,
 * src/intersection.cpp:84-130 (quirks included). */
enum { PTB_OP_PUSH_PRIM = 0, PTB_OP_UNION = 1, PTB_OP_DIFFERENCE = 2, PTB_OP_INTERSECTION = 3 };
#define PTB_OP(kind, arg) ((uint32_t)(kind) | ((uint32_t)(arg) << 8))
#define PTB_OP_KIND(op) ((op) & 0xFFu)
#define PTB_OP_ARG(op) ((op) >> 8)

/* Texture programs.  Every reference wrapper texture has exactly one delegate, so a texture is
 * a linear chain of wrapper ops ending in one leaf op, stored outermost-first in texops[].
 * A texture is named by the index of its first op. */
enum {
    /* leaves */
    PTB_TEX_CONST = 0,        /* ColorTexture              include/texture.h:29-58; data: r,g,b      */
    PTB_TEX_IMAGE = 1,        /* ImageTexture              include/image_texture.h:9-35; arg: image  */
    PTB_TEX_IMAGE_ALPHA = 2,  /* ImageAlphaTexture         include/image_texture.h:37-70             */
    PTB_TEX_SKYBOX = 3,       /* ImageSkyboxTexture        include/image_texture.h:72-115; arg: first
                                 of 6 consecutive images top,bottom,left,right,front,back            */
    PTB_TEX_SKYBOX_ALPHA = 4, /* ImageSkyboxAlphaTexture   include/image_texture.h:117-181           */
    /* wrappers */
    PTB_TEX_AFFINE = 8,       /* TransformedTexture        include/texture.h:60-90; data: 12 floats  */
    PTB_TEX_MULTIPLY = 9,     /* MultiplyTexture           include/filter_texture.h:30-47; data rgb  */
    PTB_TEX_LOG = 10,         /* LogTexture                include/filter_texture.h:50-73            */
    PTB_TEX_MIRRORBALL = 11,  /* MirrorBallSkymapTexture   include/transform_texture.h:35-59         */
    PTB_TEX_SPHERICAL = 12    /* SphericalCoordinatesSkymapTexture include/transform_texture.h:61-86 */
};
#define PTB_TEX_IS_LEAF(kind) ((kind) < 8)

typedef struct PtbTexOp {
    int32_t kind; /* PTB_TEX_*                                               */
    int32_t arg;  /* offset into texdata[] (floats) or image index           */
} PtbTexOp;

/* Material (reference include/material.h:10-37): five textures + ior. */
enum { PTB_MAT_REFLECT = 0, PTB_MAT_SCATTER = 1, PTB_MAT_EMISSIVE = 2, PTB_MAT_TRANSMIT = 3, PTB_MAT_TRC = 4 };
typedef struct PtbMaterial {
    int32_t tex[5]; /* texture roots, PTB_MAT_* order */
    float ior;
    int32_t reserved[2];
} PtbMaterial;

/* RGBA float image, row 0 first, as reference Image (include/image.h:48-99). */
typedef struct PtbImageDesc {
    const float *rgba;
    int32_t width, height;
} PtbImageDesc;

typedef struct PtbSceneDesc {
    uint32_t abi_version; /* PTB200_ABI_VERSION */
    int32_t n_prims, n_xforms, n_chains, n_chain_xf, n_ops, n_materials, n_texops, n_texdata, n_images;
    const PtbPrim *prims;
    const PtbXform *xforms;
    const PtbChain *chains;
    const int32_t *chain_xf;
    const uint32_t *ops;
    const PtbMaterial *materials;
    const PtbTexOp *texops;
    const float *texdata;
    const PtbImageDesc *images;
} PtbSceneDesc;

/* Opaque.  The scene data is immutable after create.  Calls on one scene are serialised: a scene owns ONE set of
 * device scratch (ray queues, hit records, counters, span-pool overflow), so a call first waits — on the device,
 * cudaStreamWaitEvent — for the previous call on that scene if it ran on another stream, and the host side is
 * guarded by a mutex.  Threads may share a scene; for calls that overlap on the GPU create one scene per stream.
 * ptb200_scene_destroy waits for the last call on the scene to finish. */
typedef struct PtbScene PtbScene;

/* ---- ray-batch intersection (BASELINE config 3) --------------------------------------------- */

enum { PTB_HIT_MISS = 0, PTB_HIT_ENTER = 1 /* hit a span start */, PTB_HIT_EXIT = 2 /* hit a span end */ };

/* First boundary with t >= eps — the scan at reference include/path-trace.h:66-101. */
typedef struct PtbHit {
    float t;          /* -1 on miss                                                               */
    float normal[3];  /* faces against the ray: startNormal, or -endNormal (path-trace.h:75,89)   */
    int32_t kind;     /* PTB_HIT_*                                                                */
    int32_t prim;     /* primitive whose surface was hit, -1 on miss                              */
    int32_t material; /* its material, -1 on miss                                                 */
    int32_t n_spans;  /* length of the root span list for this ray (may exceed max_spans)         */
} PtbHit;

/* One span of the root iterator (reference include/span.h:12-120). */
typedef struct PtbSpan {
    float start, end;
    float start_normal[3];
    float end_normal[3];
    int32_t start_prim, end_prim;         /* primitive that contributed each boundary             */
    int32_t start_material, end_material;
} PtbSpan;

/* flags for ptb200_intersect* and PtbRenderParams.flags */
enum {
    PTB_FLAG_EXACT = 1u,    /* no FMA contraction, IEEE div/sqrt: bit-identical t with the x86-64
                               reference build.  Without it the compiler may fuse a*b+c and use
                               rsqrt-based normalisation (|rel err| < 1e-6, within the 1e-5 gate). */
    PTB_FLAG_BILINEAR = 2u, /* non-reference extra: bilinear image fetch instead of the reference's
                               nearest texel                                        */
    PTB_FLAG_REJECTION = 4u,/* scatter by the reference's literal rejection loops
                               (path-trace.h:142-158) instead of the divergence-free cap sampler   */
    PTB_FLAG_ORIGIN0 = 8u,  /* ptb200_intersect*: the caller promises every origin is (0,0,0), as for
                               all camera rays (path-trace.h:196); selects the camera-ray variant of
                               the fast evaluator, which the renderer's primary stage runs          */
    PTB_FLAG_SPECIALIZE = 16u, /* run the scene-specialised trace kernels (below), compiling them now if
                               need be; an error if that is not possible                              */
    PTB_FLAG_INTERPRET = 32u   /* never run the scene-specialised kernels in this call                */
};

/* ---- rendering ------------------------------------------------------------------------------ */

typedef struct PtbRenderParams {
    int32_t width, height;         /* screenXResolution / screenYResolution of tracePixel          */
    int32_t x0, y0, x1, y1;        /* pixel rectangle to render, [x0,x1) x [y0,y1); 0,0,0,0 = all  */
    int32_t spp;                   /* camera samples per pixel rendered by THIS call               */
    int32_t sample_offset;         /* global index of this call's first sample                     */
    int32_t sample_stride;         /* global index step (multi-GPU: rank g renders g, g+G, ...)    */
    int32_t max_depth;             /* rayDepth (reference src/test.cpp:36: 16)                     */
    float screen_width, screen_height, screen_distance; /* tracePixel's screenWidth/Height/Distance */
    uint64_t seed;
    uint32_t flags;                /* PTB_FLAG_*                                                   */
    int32_t samples_per_pass;      /* camera samples per pixel in flight per wavefront pass; 0=auto */
} PtbRenderParams;

typedef struct PtbStats {
    uint64_t paths;          /* camera samples completed                                           */
    uint64_t rays;           /* ray segments intersected against the whole scene                   */
    uint64_t launches;       /* kernels launched by this call                                      */
    uint64_t queue_bytes;    /* bytes read + written through the HBM ray queues                    */
    float ms_total;          /* device time of the whole call (CUDA events on the render stream)   */
    float ms_primary;        /* ... of the primary level (camera rays: trace + shade kernels)      */
    float ms_bounce;         /* ... of the bounce levels (queue -> trace -> shade -> queue)        */
    float ms_resolve;        /* ... of the framebuffer resolve kernel                              */
    uint64_t rays_primary;   /* rays traced by the primary kernels                                 */
    uint64_t rays_bounce;    /* rays traced by the bounce kernels                                  */
    float ms_trace_primary;  /* ... of the primary TRACE kernel alone (the roofline kernel)        */
    float ms_trace_bounce;   /* ... of the bounce TRACE kernels alone                              */
    /* ABI 2 */
    uint64_t rays_first_bounce; /* survivors of the primary shade = rays of bounce level 1          */
    uint64_t launches_elided;   /* trace/shade launches not made because the level's queue was empty */
    float ms_shade_primary;  /* ... of the primary SHADE kernel alone                              */
    float ms_shade_bounce;   /* ... of the bounce SHADE kernels alone                              */
} PtbStats;

/* ---- entry points --------------------------------------------------------------------------- */

#ifndef __CUDACC_RTC__ /* host entry points: not seen by the run-time compiled device code */
const char *ptb200_last_error(void);
int ptb200_abi_version(void);
int ptb200_device_count(void);
/* Device scratch (ray queues, hit records, accumulation buffers: GBs at 4K) is cached by the library
 * when a scene is destroyed and reused by the next one, because cudaMalloc/cudaFree of such blocks
 * cost up to seconds.  ptb200_trim() gives the cached blocks back to the driver.  (No reference
 * counterpart: the reference allocates per RenderBlock on the heap, src/test.cpp:325-340.) */
int ptb200_trim(void);

/* Copies everything it needs (caller keeps its buffers); the scene lives on the CUDA device that
 * is current at the time of the call.  Validates the program (stack discipline, index ranges).
 * Limits (PTB_EINVAL beyond them; the reference has none): at most 32767 primitives (boundary tags are 16 bits)
 * and at most 32 pending operands on the evaluator's stack — a balanced tree of any size fits (unionArray,
 * src/test.cpp:52-64, needs log2 n), a right-deep chain `new Union(a, new Union(b, ...))` of more than 32
 * objects does not: build it left-deep or balanced (Union is the same set either way).
 * Image pixels that lie in page-locked host memory (cudaHostAlloc / cudaHostRegister) are DMA-ed from there;
 * pageable ones are staged through the library's own pinned buffer (slower, and a CPU copy). */
int ptb200_scene_create(const PtbSceneDesc *desc, PtbScene **out);
void ptb200_scene_destroy(PtbScene *scene);
/* Appends an image after create (texops may name images not yet uploaded at create time). */
int ptb200_image_upload(PtbScene *scene, const float *rgba, int width, int height, int *handle);

/* Host-buffer ray batch: origins/dirs are n x 3 floats.  hits: n records.  spans: optional
 * n x max_spans records (NULL to skip).  Copies in, runs the CSG kernel, copies out. */
int ptb200_intersect(PtbScene *scene, const float *origins, const float *dirs, size_t n,
                     PtbHit *hits, PtbSpan *spans, int max_spans, uint32_t flags);
/* Device-resident variant: rays as two float4 arrays (xyz + pad), results stay on the device.
 * stream is a cudaStream_t (NULL = default stream).  Asynchronous. */
int ptb200_intersect_device(PtbScene *scene, const void *d_origins4, const void *d_dirs4, size_t n,
                            void *d_hits, void *d_spans, int max_spans, uint32_t flags, void *stream);

/* ptb200_intersect_device WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, same kernels (the interpreting ones: PTB_FLAG_SPECIALIZE is refused, a
 * run-time compile allocates) and the same results as ptb200_intersect_device; the kernels' scratch — the span-pool
 * overflow area that span lists longer than the shared-memory pool spill to — is a static device array of
 * PTB_STATIC_SCRATCH_BYTES (16 MiB, held by every GPU that loads the library until exit) instead of a cudaMalloc'ed
 * block, and the grid shrinks (never below one CTA per SM) until it fits.  PTB_EINVAL with a message that names both
 * sizes when the scene's worst-case span lists need more than that at one CTA per SM (about 73 spans beyond the shared
 * pool per ray on a 148-SM B200).  Calls no cudaMalloc / cudaFree / allocating library routine; asynchronous on
 * <stream>.  Every call on a GPU shares the one array: calls that overlap on that GPU race on it — serialise them
 * (one stream, or events).  Nothing in the array needs clearing between calls: the kernels write a slot before they
 * read it.  _static_scratch_bytes returns PTB_STATIC_SCRATCH_BYTES. */
#define PTB_STATIC_SCRATCH_BYTES (16u << 20)
int ptb200_intersect_device_static(PtbScene *scene, const void *d_origins4, const void *d_dirs4, size_t n, void *d_hits, void *d_spans, int max_spans, uint32_t flags,
                                   void *stream);
size_t ptb200_static_scratch_bytes(void);

/* Evaluates texture program <tex_root> at n positions (n x 3 floats, host): the reference's
 * Texture::getColor -> rgb (n x 3) and Texture::getFloat -> f (n); either output may be NULL. */
int ptb200_texture_eval(PtbScene *scene, int tex_root, const float *positions, size_t n,
                        float *rgb, float *f, uint32_t flags);

/* Synchronous frame render.  rgba_out: host, width*height*4 floats, mean radiance over spp in
 * rgb and 1 in alpha (pixels outside the rectangle are left untouched). */
int ptb200_render(PtbScene *scene, const PtbRenderParams *params, float *rgba_out, PtbStats *stats);
/* tracePixel for a list of n pixels (xy: n pairs x, y): rgba_out gets n x 4 floats, mean radiance over spp and
 * alpha 1.  Each pixel has the value a frame render with the same params gives it (same samples, same RNG
 * streams); params->x0..y1 are ignored.  The reference calls tracePixel per pixel (include/path-trace.h:187-207,
 * src/test.cpp:450); this is that call for a batch. */
int ptb200_render_pixels(PtbScene *scene, const PtbRenderParams *params, const int32_t *xy, size_t n, float *rgba_out, PtbStats *stats);
/* Adds this call's UNNORMALISED radiance sums (rgb) and completed-sample counts (alpha) into a
 * device buffer of width*height float4.  Used for sample-sharded multi-GPU rendering: every rank
 * accumulates its slice, one NCCL sum-reduce follows, ptb200_resolve_device divides.  Synchronous
 * with respect to <stream> only if stats != NULL. */
int ptb200_render_accumulate_device(PtbScene *scene, const PtbRenderParams *params, void *d_accum4,
                                    PtbStats *stats, void *stream);
/* rgba[i] = (accum.rgb / accum.a, 1) where accum.a > 0, else 0.  In place allowed. */
int ptb200_resolve_device(const void *d_accum4, void *d_rgba4, size_t n_pixels, void *stream);

/* Adaptive rendering: the pixel-skipping quadtree of the reference's tile renderer,
 * RenderBlock::renderSquare / interpolateSquare / colorCloseEnough (src/test.cpp:423-499), opt-in
 * (it changes the image: skipped pixels are bilinear blends of traced ones).  The image is cut into
 * block_size x block_size tiles whose corners are traced; a square whose four corners are pairwise
 * within min_color_delta (relative, |a-b|^2 <= delta^2 |a|^2) and whose edge is <= max_sample_size is
 * interpolated, otherwise its five edge/centre midpoints are traced and its four quarters follow.
 * The decisions, the interpolation arithmetic and the final image are exactly those of the reference
 * for the same traced values; the walk is level-synchronous over all tiles instead of depth-first per
 * tile, and a lattice point shared by two tiles is traced once (the reference traces it per tile).
 * Reference values at 1920x1080: block_size 32, max_sample_size 9, min_color_delta 0.003
 * (src/test.cpp:38,50).  params->x0..y1 must name the whole image.  source_out (nullable,
 * width*height bytes): 1 = traced pixel, 2 = interpolated pixel.  lattice_out (nullable): the traced
 * lattice, rows x pitch x 4 floats with pitch = ceil(width/block_size)*block_size + 1 and rows likewise
 * from height — mean radiance and alpha 1 at every traced point (those on and beyond the right and
 * bottom image edge included, which the reference traces too), zeros elsewhere. */
typedef struct PtbAdaptiveParams {
    int32_t block_size;      /* blockSize: power of two                                            */
    int32_t max_sample_size; /* maximumSampleSize: largest square edge that may be interpolated    */
    float min_color_delta;   /* minimumColorDelta                                                   */
    int32_t reserved;
} PtbAdaptiveParams;
typedef struct PtbAdaptiveStats {
    uint64_t pixels_traced;        /* lattice points traced (incl. those beyond the right/bottom edge) */
    uint64_t pixels_interpolated;  /* visible pixels filled by interpolateSquare                       */
    uint64_t squares_interpolated;
    int32_t levels;                /* quadtree levels decided                                          */
    int32_t reserved;
} PtbAdaptiveStats;
#ifndef __CUDACC_RTC__
int ptb200_render_adaptive(PtbScene *scene, const PtbRenderParams *params, const PtbAdaptiveParams *adaptive,
                           float *rgba_out, uint8_t *source_out, float *lattice_out, PtbStats *stats,
                           PtbAdaptiveStats *adaptive_stats);
#endif

/* 8-bit display conversion of the reference's main loop (src/test.cpp:992-995):
 * byte = max(0, min(255, (int)floor(256 * c / count))) per channel, in float, in that order; output
 * is r,g,b,255 per pixel.  divide_by_alpha != 0: count = the pixel's alpha (an accumulation buffer;
 * alpha 0 gives black); else count = 1 (a resolved image).  _device: both buffers on the device. */
int ptb200_display8_device(const void *d_rgba4, void *d_rgbx8, size_t n_pixels, int divide_by_alpha, void *stream);
int ptb200_display8(const float *rgba, size_t n_pixels, int divide_by_alpha, uint8_t *rgbx8);

/* Progressive / asynchronous boundary mirroring BlockRenderer::done()/copyToBuffer()
 * (reference src/test.cpp:310-323).  One job at a time per scene. */
typedef struct PtbJob PtbJob;
int ptb200_render_async(PtbScene *scene, const PtbRenderParams *params, PtbJob **job);
int ptb200_job_done(PtbJob *job);                       /* 1 finished, 0 running, <0 error        */
/* Copies the pixels that already have at least one completed sample into rgba (host,
 * width*height*4); others are left untouched.  May be called while the job runs. */
int ptb200_job_read(PtbJob *job, float *rgba, int width, int height);
int ptb200_job_wait(PtbJob *job, PtbStats *stats);
void ptb200_job_destroy(PtbJob *job);

/* ---- multi-GPU: sample-sharded rendering with ONE NCCL exchange step ---------------------------------
 * Replaces the reference's tile farm seam (NetRenderBlock client, src/test.cpp:520-630; server(),
 * :681-778): there a C++ main hands 32x32 tiles to render servers over TCP; here rank g of G renders the
 * global camera-sample indices g, g+G, ... of EVERY pixel (the RNG is keyed by the global index, so the
 * union over ranks is exactly the one-GPU sample set), then one ncclReduceScatter of the accumulation frame
 * by row blocks over NVLink, and every rank divides and delivers its own rows.  libnccl.so.2 is loaded on
 * first use.  PTB_ENODEV when NCCL is missing.
 *
 * _comm_init_all: ONE process drives n devices (devices == NULL: 0..n-1), ncclCommInitAll, one host thread
 *   and stream per device per frame; calls on such a communicator are synchronous.
 * _comm_unique_id / _comm_init_rank: one process per GPU (torchrun, MPI); rank 0 makes the 128-byte id, the
 *   caller ships it; the member lives on the CUDA device current at the call.
 * _scene_create_multi: the scene on every LOCAL member, scenes[0..nlocal).  Image pixels cross PCIe once (rank
 *   0; other ranks may pass rgba == NULL in PtbImageDesc) and reach the other GPUs by ncclBroadcast.  Collective.
 * _render_multi: one whole frame of params->spp samples over all ranks.  Collective.  rgba_out (host,
 *   width*height*4 floats, nullable): every local member writes ITS rows (ptb200_multi_rows) — the whole image in
 *   init_all mode, this rank's row block in init_rank mode.  d_rgba_out (nullable): per local member a device
 *   image of width*height float4 on that member's device; its own rows are written.  stats (nullable): nlocal
 *   entries.  stream: init_rank mode only (NULL = default); asynchronous there unless rgba_out or stats is given.
 * _multi_rows / _multi_shard: the partition rules (pure functions, no GPU): rank r owns rows
 *   [r*ceil(H/G), min(H, (r+1)*ceil(H/G))) and samples r, r+G, ... below spp. */
#define PTB_COMM_ID_BYTES 128
typedef struct PtbComm PtbComm;
#ifndef __CUDACC_RTC__
int ptb200_comm_unique_id(void *id);
int ptb200_comm_init_rank(const void *id, int rank, int nranks, PtbComm **out);
int ptb200_comm_init_all(const int *devices, int n, PtbComm **out);
int ptb200_comm_info(const PtbComm *comm, int *nranks, int *nlocal, int *first_rank);
void ptb200_comm_destroy(PtbComm *comm);
int ptb200_scene_create_multi(PtbComm *comm, const PtbSceneDesc *desc, PtbScene **scenes);
int ptb200_render_multi(PtbComm *comm, PtbScene *const *scenes, const PtbRenderParams *params, float *rgba_out, void *const *d_rgba_out, PtbStats *stats, void *stream);
int ptb200_multi_rows(int height, int nranks, int rank, int *row0, int *row1);
int ptb200_multi_shard(int spp, int nranks, int rank, int *count, int *offset, int *stride);
#endif

/* Scene-specialised trace kernels.  The CSG program of a scene is also printed as straight-line CUDA
 * (every header, matrix and plane a literal) and compiled for sm_100a at run time with NVRTC; the trace
 * kernels built from it compute exactly what the interpreting ones do with about half the instructions.
 * By default a render / ray batch with >= 2^26 rays to trace starts that compile on a background thread
 * and switches over when it is done (between passes, if need be); compiled kernels are kept per process
 * and per distinct scene program (at most 64 programs per process, one background compile at a time: further
 * scenes stay interpreted).  Programs of more than 96 words are specialised for camera rays only, and only when
 * the generated code can test bounding spheres before its calls (csrc/ptb_specialize.cu); PTB_SPEC_LARGE=0 keeps
 * them interpreted.  Environment: PTB_SPECIALIZE=0 never, =1 compile and wait at first
 * use.  Flags: PTB_FLAG_SPECIALIZE / PTB_FLAG_INTERPRET per call.  No reference counterpart (the
 * reference walks a tree of virtual SpanIterators, include/object.h:10-24).
 * _probe needs no GPU: validates desc, writes the generated source (NUL-terminated, truncated to
 * source_capacity; *source_length = full length) and, when cubin_bytes != NULL, compiles it and reports
 * the size of the sm_100a cubin.
 * _specialized: 1 when the scene's specialised kernels are compiled and loaded (2: and, the program being
 * large, the last render ran them for camera rays only and interpreted its queued rays), else 0 (with
 * *compile_ms = -1 while a compile that a render call asked for is still under way); also reports
 * how many wavefront passes of the last render call that asked for stats ran them. */
int ptb200_specialize_probe(const PtbSceneDesc *desc, int which /* 0: render trace kernels, 1: ray-batch kernels, 2: the compiled program stream as hex text (tests) */, char *source, size_t source_capacity,
                            size_t *source_length, size_t *cubin_bytes);
int ptb200_scene_specialized(PtbScene *scene, double *compile_ms, int *passes_specialized, int *passes_total);

/* Tile wire text of the reference's TCP render farm, host only; sockets stay with the
 * caller.  Server side (src/test.cpp:678-716): _request_parse reads the client's " x y size" line with
 * the reference's validity checks; _encode writes every pixel of tile [x,x+size) x [y,y+size) of a
 * width x height RGBA frame that has a value (alpha > 0) and is not flagged in sent[] (size*size bytes,
 * updated; NULL = send all) as "P<x>,<y>=<r>,<g>,<b>\n" (printf %i / %g, RenderBlock::copyToSocket
 * :372-391), then "E\n", then "F\n" when final != 0 (:715); *length = bytes written (without the NUL);
 * out == NULL only measures.  Client side (NetRenderBlock::run, :596-630): _decode consumes such text
 * into a size x size tile (rgb_tile: 3 floats per pixel, valid: 1 per pixel received), skipping 'S',
 * 'E' and newlines, PTB_EINVAL on a pixel outside the tile or an unknown record; *finished = 1 at 'F';
 * an incomplete last record is left unconsumed (*consumed = bytes used). */
int ptb200_tile_request_parse(const char *line, int screen_width, int screen_height, int *x, int *y, int *size);
int ptb200_tile_encode(const float *rgba, int width, int height, int x, int y, int size, uint8_t *sent, int final, char *out, size_t capacity, size_t *length);
int ptb200_tile_decode(const char *text, size_t length, int x, int y, int size, float *rgb_tile, uint8_t *valid, int *finished, size_t *consumed);

/* Draws n scatter directions (n x 3 floats, host) from the lobe sampler of the shading kernel for
 * one unit normal, un-normalised mirror vector and scatter coefficient (reference
 * include/path-trace.h:140-158); (0,0,0) marks a dropped vertex.  For statistical checks. */
int ptb200_sample_scatter(const float *normal, const float *mirror, float scatter, size_t n, uint64_t seed, uint32_t flags, float *dirs);

/* Vector3D::reflect / refract / refractStrength (reference include/vector3d.h:186-214) as the shade kernel
 * computes them, for known-answer tests.  v, normal: n x 3 floats; ior: n floats (op 1, 2).
 * op 0 reflect -> out n x 3; 1 refract -> out n x 3; 2 refractStrength -> out n. */
int ptb200_vector_eval(int op, const float *v, const float *normal, const float *ior, size_t n, float *out);
/* The random words the render kernels draw in place of the reference's shared LCG (DefaultRandomEngine,
 * include/path-trace.h:21-54; parity is statistical), for statistical tests.  which 0: the
 * sub-pixel jitter hash of camera sample (pixel[i], sample[i]) -> out 2 words per entry; which 1: Philox4x32-10
 * block block[i] (NULL: block 0) of that sample's stream -> out 4 words per entry.  A word w maps to [0,1] as
 * float(w) / 2^32, the reference's uniform_real_distribution (include/vector3d.h:26-33). */
int ptb200_rng_words(int which, uint64_t seed, const uint32_t *pixel, const uint32_t *sample, const uint32_t *block, size_t n, uint32_t *out);

/* FP32 FMA-chain microbenchmark on the current device: measured lane-FMA/s (the FP32 issue-peak
 * denominator of the roofline). */
int ptb200_fp32_peak(double *fma_per_second, double *sm_clock_mhz);
#endif /* !__CUDACC_RTC__ */

#ifdef __cplusplus
}
#endif
#endif /* PTB200_H_INCLUDED */
