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
std::string specializeSource(const std::vector<uint4> &fastProgram, int which);
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
