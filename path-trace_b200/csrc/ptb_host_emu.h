// ptb_host_emu.h — TEST SUPPORT: lets the device headers of the CSG evaluators (ptb_math.cuh, ptb_csg.cuh,
// ptb_csg_fast.cuh, ptb_csg_reg.cuh) and the code GENERATED for one scene compile as plain host C++, one lane at a
// time, so that tests/test_spec_emulation.py can run the generated register-resident evaluator against the
// interpreter on the CPU (no GPU in the build container).  Only defined under PTB_HOST_EMU; the product never
// sets it.  One emulated thread (threadIdx.x = 0), a "warp" of one lane: votes return their own predicate.
// MUFU approximations (rcp / sqrt / rsqrt) become correctly rounded host operations — both evaluators go through
// the same helpers, so they still have to agree bit for bit with each other.
#pragma once
#ifdef PTB_HOST_EMU
#include <math.h>
#include <stdint.h>
#include <string.h>

#define __device__
#define __host__
#define __global__
#define __constant__ static const
#define __forceinline__ inline
#define __noinline__
#define __restrict__
#define __launch_bounds__(...)

struct uint2 { uint32_t x, y; };
struct uint3 { uint32_t x, y, z; };
struct uint4 { uint32_t x, y, z, w; };
struct int2 { int x, y; };
struct int4 { int x, y, z, w; };
struct float4 { float x, y, z, w; };
struct uchar4 { unsigned char x, y, z, w; };
static inline uint2 make_uint2(uint32_t x, uint32_t y) { uint2 v = {x, y}; return v; }
static inline uint3 make_uint3(uint32_t x, uint32_t y, uint32_t z) { uint3 v = {x, y, z}; return v; }
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { uint4 v = {x, y, z, w}; return v; }
static inline float4 make_float4(float x, float y, float z, float w) { float4 v = {x, y, z, w}; return v; }
typedef unsigned long long cudaTextureObject_t;
typedef void *cudaStream_t;

struct PtbEmuDim { unsigned x, y, z; };
static PtbEmuDim threadIdx = {0, 0, 0}; // (a harness may step it to emulate a CTA-wide copy loop)
static const PtbEmuDim blockIdx = {0, 0, 0}, blockDim = {128, 1, 1}, gridDim = {1, 1, 1};

// shared memory of the one emulated CTA: "shared addresses" are byte offsets into it
static uint4 ptbEmuShared[16384];
#define PTB_EMU_SMEM ((char *)ptbEmuShared)

static inline uint32_t __float_as_uint(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static inline int __float_as_int(float f) { int u; memcpy(&u, &f, 4); return u; }
static inline float __int_as_float(int u) { float f; memcpy(&f, &u, 4); return f; }
static inline float __uint2float_rn(uint32_t u) { return (float)u; }
static inline uint32_t __umulhi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
static inline int __popc(uint32_t v) { return __builtin_popcount(v); }
static inline bool __any_sync(unsigned, bool p) { return p; }
static inline bool __all_sync(unsigned, bool p) { return p; }
static inline unsigned __ballot_sync(unsigned, bool p) { return p ? 1u : 0u; }
static inline unsigned __reduce_or_sync(unsigned, unsigned v) { return v; }
static inline void __syncthreads() {}
static inline void __syncwarp() {}
template <class T> static inline T __ldg(const T *p) { return *p; }
// compile with -ffp-contract=off: these must stay separately rounded operations
static inline float __fmul_rn(float a, float b) { return a * b; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fsub_rn(float a, float b) { return a - b; }
static inline float __fdiv_rn(float a, float b) { return a / b; }
static inline float __fsqrt_rn(float a) { return sqrtf(a); }
static inline float __fdividef(float a, float b) { return a / b; }
static inline float rsqrtf(float a) { return 1.0f / sqrtf(a); }
#endif // PTB_HOST_EMU
