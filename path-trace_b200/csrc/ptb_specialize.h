// ptb_specialize.h — scene-specialised trace kernels (NVRTC), see ptb_specialize.cu.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

namespace ptb
{

struct SpecKernels
{
    void *module = nullptr;       // CUmodule
    void *tracePrimary = nullptr; // CUfunction: camera rays
    void *traceQueue = nullptr;   // queued rays
    void *intersect = nullptr;    // ray batches (ptb200_intersect*, fast policy)
    void *intersectO0 = nullptr;  // ... with PTB_FLAG_ORIGIN0
    int regsTrace = 0;
    size_t cubinBytes = 0;
    double compileMs = 0.0;
};

// The CUDA source generated for a compiled fast stream (ptb_csg_fast.cuh): the evaluator plus either the
// two trace kernels of the renderer or the two ray-batch kernels (separate programs: half the compile time).
enum
{
    SPEC_TRACE = 0,
    SPEC_INTERSECT = 1
};
// The postfix CSG sequence the fast stream was compiled from, with the stream position of every leaf record: what
// the generator of the register-resident evaluator walks (it needs the TREE; the stream only has its flattening).
struct FastTreeItem
{
    int kind;        // PTB_OP_*
    int prim;        // PUSH_PRIM: the primitive
    int clusterEnd;  // first leaf of a plane cluster: sequence index of the Union that completes it, else -1
    int leafWord;    // PUSH_PRIM: word index of the leaf's own header in the fast stream
    int clusterWord; // first leaf of a plane cluster: word index of the cluster record
};
struct FastTree
{
    std::vector<FastTreeItem> items;
};
// PTB_SPEC_GEN=1 selects the first-generation generator (one labelled block per stream record, pool + bit stack);
// the default is the register-resident evaluator (ptb_csg_reg.cuh).
std::string specializeSource(const std::vector<uint4> &fastProgram, const FastTree &tree, int which);
// NVRTC: source -> sm_100a cubin.  Needs no GPU.  0 on success; log has the compiler's output.
int specializeCompile(const std::string &source, std::vector<char> &cubin, std::string &log);
// The kernels compiled from `source` (specializeSource), loaded into the current context of `device`.
// Compiled once per distinct source for the life of the process.  mode: SPEC_PEEK only looks;
// SPEC_START starts the compile on a background thread if nobody has yet and returns at once;
// SPEC_WAIT compiles (or waits for a running compile).  NULL while compiling or when impossible (err
// then says why): the caller keeps the interpreting kernels, whose results are the same.
enum
{
    SPEC_PEEK = 0,
    SPEC_START = 1,
    SPEC_WAIT = 2
};
const SpecKernels *specializeGet(const std::string &source, int device, int mode, std::string &err);
int specOccupancy(void *fn, int threads, size_t smemBytes);
int specLaunch(void *fn, int blocks, int threads, size_t smemBytes, cudaStream_t stream, void **args, std::string &err);

} // namespace ptb
