// ptb_specialize.cu — scene-specialised trace kernels: the compiled fast stream of ONE scene printed
// as straight-line CUDA, compiled at run time with NVRTC for sm_100a and loaded through the driver API.
// Host code only.  libnvrtc and libcuda are dlopen()ed on first use, so libptb200.so still loads (and
// fails loudly with PTB_ENODEV) on a machine without a driver; when either is missing or the compile
// fails the renderer keeps the interpreting kernels — same results, more instructions per ray.
#include <dlfcn.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <algorithm>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "ptb200.h"
#include "ptb_specialize.h"
#include "ptb_csg_fast.cuh"

#include "ptb_embedded.inc" // kEmbeddedNames[], kEmbeddedSources[], kEmbeddedCount: the device headers as strings

namespace ptb
{

namespace
{
// ---- the two libraries ------------------------------------------------------------------------------
typedef int (*nvrtcCreateProgram_t)(void **, const char *, const char *, int, const char *const *, const char *const *);
typedef int (*nvrtcCompileProgram_t)(void *, int, const char *const *);
typedef int (*nvrtcGetSize_t)(void *, size_t *);
typedef int (*nvrtcGetData_t)(void *, char *);
typedef int (*nvrtcDestroyProgram_t)(void **);
typedef int (*cuModuleLoadData_t)(void **, const void *);
typedef int (*cuModuleGetFunction_t)(void **, void *, const char *);
typedef int (*cuLaunchKernel_t)(void *, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned, void *, void **, void **);
typedef int (*cuFuncSetAttribute_t)(void *, int, int);
typedef int (*cuFuncGetAttribute_t)(int *, int, void *);
typedef int (*cuOccupancy_t)(int *, void *, int, size_t);
typedef int (*cuGetErrorString_t)(int, const char **);

struct Libs
{
    bool tried = false, ok = false, triedDriver = false, okDriver = false;
    std::string why;
    nvrtcCreateProgram_t nvrtcCreateProgram;
    nvrtcCompileProgram_t nvrtcCompileProgram;
    nvrtcGetSize_t nvrtcGetCUBINSize, nvrtcGetProgramLogSize;
    nvrtcGetData_t nvrtcGetCUBIN, nvrtcGetProgramLog;
    nvrtcDestroyProgram_t nvrtcDestroyProgram;
    cuModuleLoadData_t cuModuleLoadData;
    cuModuleGetFunction_t cuModuleGetFunction;
    cuLaunchKernel_t cuLaunchKernel;
    cuFuncSetAttribute_t cuFuncSetAttribute;
    cuFuncGetAttribute_t cuFuncGetAttribute;
    cuOccupancy_t cuOccupancy;
    cuGetErrorString_t cuGetErrorString;
};
Libs g_libs;
std::mutex g_lock;

#define SYM(lib, field, name)                                                                          \
    if(!(g_libs.field = (decltype(g_libs.field))dlsym(lib, name)))                                     \
    {                                                                                                  \
        g_libs.why = std::string("missing symbol ") + name;                                            \
        return false;                                                                                  \
    }
// NVRTC alone is enough to compile (no GPU, no driver needed)
bool loadNvrtc()
{
    if(g_libs.tried)
        return g_libs.ok;
    g_libs.tried = true;
    void *rtc = nullptr;
    for(const char *name : {"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12"})
        if((rtc = dlopen(name, RTLD_NOW | RTLD_GLOBAL)))
            break;
    if(!rtc)
    {
        g_libs.why = "libnvrtc.so.12 not found";
        return false;
    }
    SYM(rtc, nvrtcCreateProgram, "nvrtcCreateProgram")
    SYM(rtc, nvrtcCompileProgram, "nvrtcCompileProgram")
    SYM(rtc, nvrtcGetCUBINSize, "nvrtcGetCUBINSize")
    SYM(rtc, nvrtcGetCUBIN, "nvrtcGetCUBIN")
    SYM(rtc, nvrtcGetProgramLogSize, "nvrtcGetProgramLogSize")
    SYM(rtc, nvrtcGetProgramLog, "nvrtcGetProgramLog")
    SYM(rtc, nvrtcDestroyProgram, "nvrtcDestroyProgram")
    g_libs.ok = true;
    return true;
}
bool loadDriver()
{
    if(g_libs.triedDriver)
        return g_libs.okDriver;
    g_libs.triedDriver = true;
    void *cu = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
    if(!cu)
    {
        g_libs.why = "libcuda.so.1 not found";
        return false;
    }
    SYM(cu, cuModuleLoadData, "cuModuleLoadData")
    SYM(cu, cuModuleGetFunction, "cuModuleGetFunction")
    SYM(cu, cuLaunchKernel, "cuLaunchKernel")
    SYM(cu, cuFuncSetAttribute, "cuFuncSetAttribute")
    SYM(cu, cuFuncGetAttribute, "cuFuncGetAttribute")
    SYM(cu, cuOccupancy, "cuOccupancyMaxActiveBlocksPerMultiprocessor")
    SYM(cu, cuGetErrorString, "cuGetErrorString")
    g_libs.okDriver = true;
    return true;
}
#undef SYM

std::string hex(uint32_t v)
{
    char buf[16];
    snprintf(buf, sizeof(buf), "0x%08xu", v);
    return buf;
}
std::string u4(const uint4 &v) { return "make_uint4(" + hex(v.x) + ", " + hex(v.y) + ", " + hex(v.z) + ", " + hex(v.w) + ")"; }
std::string label(size_t w) { return "W" + std::to_string(w); }

} // namespace

// One labelled block per record of the fast stream (formats: ptb_csg_fast.cuh, compileFastProgram in
// ptb_capi.cu).  Every block is the body of one trip of evalFast's loop with the record's words as
// literals; PTB_ADVANCE becomes fall-through or goto.
static std::string specializeSourceGen1(const std::vector<uint4> &prog, int which)
{
    std::string s;
    if(const char *e = getenv("PTB_SPEC_MIN_CTAS")) // experiment: resident CTAs per SM the trace kernels are compiled for; default 8 = 64 registers (C5 camera-ray trace: 6 -> 288 ms, 7 -> 272, 8 -> 267, 10 -> 306 with spills, 12 -> 318)
        s += "#define PTB_TRACE_MIN_CTAS " + std::to_string(atoi(e)) + "\n";
    s += "#include \"ptb200.h\"\n#include \"ptb_csg_spec.cuh\"\nnamespace ptb\n{\n";
    s += "template <bool O0>\n__device__ __forceinline__ void evalSpec(const FastCtx &c, const V3 o, const V3 d, int &topOut)\n{\n";
    s += "    FastState st;\n    st.bits = 0u, st.top = 0, st.sp = c.stk;\n";
    s += "    bool hit = false, two = false;\n    float lt0 = 0.f, lt1 = 0.f, ut0 = 0.f, a = 0.f, b = 0.f, e = 0.f;\n    uint32_t ltags = 0u, utags = 0u;\n";
    size_t pairSecond = (size_t)-1; // record that a pair block may enter with its discriminant computed
    for(size_t w = 0; w < prog.size();)
    {
        const uint4 h = prog[w];
        const std::string HX = hex(h.x);
        s += label(w) + ":\n";
        if(h.x & FK_SPHERE_XF)
        {
            const std::string disc = "xfDisc<O0>(" + u4(h) + ", " + u4(prog[w + 1]) + ", " + u4(prog[w + 2]) + ", " + u4(prog[w + 3]) + ", o, d, ";
            const std::string push = "    specSphere(" + HX + ", " + hex(h.y) + ", a, b, e, st);\n";
            if(h.x & FH_PAIR)
            { // both discriminants together; when no ray of the warp hits the first, the second is this trip's leaf
                const uint4 g = prog[w + 4];
                s += "    {\n        float a1, b1, e1;\n        " + disc + "a, b, e);\n";
                s += "        xfDisc<O0>(" + u4(g) + ", " + u4(prog[w + 5]) + ", " + u4(prog[w + 6]) + ", " + u4(prog[w + 7]) + ", o, d, a1, b1, e1);\n";
                s += "        if(!__any_sync(0xFFFFFFFFu, e > PTB_EPS))\n        {\n            st.bits <<= " + std::to_string((h.x >> 6) & 1u) + "u;\n";
                s += "            a = a1, b = b1, e = e1;\n            goto " + label(w + 4) + "_have;\n        }\n    }\n";
                s += push;
                pairSecond = w + 4;
            }
            else
            {
                s += "    " + disc + "a, b, e);\n";
                if(w == pairSecond)
                    s += label(w) + "_have:\n"; // entered with a, b, e already computed by the pair before
                s += "    if(specMissExit(" + HX + ", e > PTB_EPS, st))\n        goto " + label(w + 4) + ";\n" + push;
            }
            w += 4;
        }
        else if(h.x & FK_PLANE)
        {
            const uint32_t k = h.z, len = 2u + 2u * k, after = (uint32_t)w + len + (k ? h.w : 0u);
            s += "    {\n        float loEnd = -3.0e38f, upStart = 3.0e38f;\n        uint32_t loTags = 0u, upTags = 0u;\n        bool full = false;\n";
            s += "        planeFold<O0>(" + u4(prog[w + 1]) + ", " + hex(h.y) + ", o, d, loEnd, upStart, loTags, upTags, full);\n";
            for(uint32_t j = 0; j < k; j++)
                s += "        planeFold<O0>(" + u4(prog[w + 3 + 2 * j]) + ", " + hex(prog[w + 2 + 2 * j].x) + ", o, d, loEnd, upStart, loTags, upTags, full);\n";
            if(k)
                s += std::string("        if(!__all_sync(0xFFFFFFFFu, !full && (") + (k == 1u ? "true" : "loEnd < upStart") + ")))\n            goto " + label(w + len) +
                     "; // the subtree's own words\n";
            s += "        planeFinish(loEnd, upStart, loTags, upTags, full, hit, two, lt0, lt1, ut0, ltags, utags);\n    }\n";
            s += "    if(specMissExit(" + HX + ", hit, st))\n        goto " + label(after) + ";\n";
            s += "    specPushOperate(" + HX + ", hit, two, lt0, lt1, ut0, ltags, utags, st);\n";
            if(after != w + len)
                s += "    goto " + label(after) + ";\n";
            w += len;
        }
        else if(h.x & FK_SPHERE)
        {
            s += "    sphereDisc<O0>(" + u4(h) + ", " + u4(prog[w + 1]) + ", o, d, a, b, e);\n";
            s += "    if(specMissExit(" + HX + ", e > PTB_EPS, st))\n        goto " + label(w + 2) + ";\n";
            s += "    specSphere(" + HX + ", " + hex(h.y) + ", a, b, e, st);\n";
            w += 2;
        }
        else
        {
            const uint32_t sub = (h.x >> 4) & 3u;
            if(sub == 0u)
                s += "    goto WEND;\n";
            else if(sub == 1u)
                s += "    specPushOperate(" + HX + ", false, false, 0.f, 0.f, 0.f, 0u, 0u, st);\n";
            else
                s += "    if(!__any_sync(0xFFFFFFFFu, (st.bits & 1u) != 0u))\n        goto " + label(w + 1 + h.y) + ";\n";
            w += 1;
        }
    }
    s += "WEND:\n    topOut = st.top;\n}\n} // namespace ptb\n";
    s += "#define PTB_EVAL(GP, O0, R, cx, gprog, o, d, tops) evalSpec<O0>(cx, o[0], d[0], tops[0])\n";
    // (PTB_TRACE_TRIP_SYNC = __syncthreads() keeps the four warps of a CTA in step through the straight-line
    // code: measured -20 % on the C4 camera-ray kernel, whose 54 KB of code thrash the instruction cache,
    // +3 % on C5, whose code fits; large programs are not specialised at all, so it stays off)
    if(getenv("PTB_SPEC_TRIP_SYNC"))
        s += "#define PTB_TRACE_TRIP_SYNC() __syncthreads()\n";
    s += "#include \"ptb_trace.cuh\"\n";
    if(which == SPEC_TRACE)
    {
        s += "extern \"C\" __global__ void __launch_bounds__(PTB_FAST_THREADS, PTB_TRACE_MIN_CTAS) ptbSpecTracePrimary(ptb::DevScene sc, int cap, float *overflow, size_t ovfStride, "
             "ptb::RenderPass pass, uint2 *hits)\n{\n    ptb::specSharedInit(cap, overflow, ovfStride);\n    ptb::tracePrimaryBody<true>(sc, cap, overflow, ovfStride, pass, hits);\n}\n";
        s += "extern \"C\" __global__ void __launch_bounds__(PTB_FAST_THREADS, PTB_TRACE_MIN_CTAS) ptbSpecTraceQueue(ptb::DevScene sc, int cap, float *overflow, size_t ovfStride, "
             "ptb::RayQueue in, const unsigned *inCount, uint2 *hits)\n{\n    ptb::specSharedInit(cap, overflow, ovfStride);\n    ptb::traceQueueBody<true>(sc, cap, overflow, ovfStride, in, inCount, hits);\n}\n";
    }
    else
        for(int o0 = 0; o0 < 2; o0++)
            s += std::string("extern \"C\" __global__ void __launch_bounds__(PTB_FAST_THREADS) ptbSpecIntersect") + (o0 ? "O0" : "") +
                 "(ptb::DevScene sc, int cap, float *overflow, size_t ovfStride, const float4 *origins, const float4 *dirs, size_t n, PtbHit *hits, PtbSpan *spans, int maxSpans)\n{\n"
                 "    ptb::specSharedInit(cap, overflow, ovfStride);\n    ptb::intersectFastBody<true, " +
                 (o0 ? "true" : "false") + ">(sc, cap, overflow, ovfStride, origins, dirs, n, hits, spans, maxSpans);\n}\n";
    return s;
}

// ---- second generation: one list variable per node of the tree, lists of up to two spans in registers ---------------
// (ptb_csg_reg.cuh).  The code is generated from the postfix sequence (the TREE) and takes every leaf's constants
// from the compiled fast stream, so the leaf arithmetic — and with it every t, tag and list length — is the
// interpreter's.
namespace
{
struct GenNode
{
    int item;        // index into FastTree::items
    int left, right; // operator: node ids of its operands; leaf: -1
    int firstItem;   // first sequence item of the subtree
};
struct Gen
{
    const std::vector<uint4> &prog;
    const FastTree &tree;
    std::vector<GenNode> nodes;
    std::string s;
    Gen(const std::vector<uint4> &prog, const FastTree &tree) : prog(prog), tree(tree) {}

    std::vector<int> alias; // node -> node whose list variable holds its result (an operator accumulates into its left operand)
    std::string var(int id) const
    {
        while(id < (int)alias.size() && alias[id] != id)
            id = alias[id];
        return "L" + std::to_string(id);
    }
    void line(int indent, const std::string &text) { s += std::string(4 * (size_t)indent, ' ') + text + "\n"; }
    bool isCluster(int id) const
    {
        const GenNode &n = nodes[id];
        return n.left >= 0 && tree.items[n.firstItem].clusterEnd == n.item && tree.items[n.firstItem].clusterWord >= 0;
    }
    // ---- where a leaf's constants come from: literals, or (inside a shape function) words of the instance's table ----
    bool tableMode = false;
    std::map<size_t, int> tabOff; // word of the prototype's records -> offset in an instance's table
    std::string W(size_t w) const { return tableMode ? "ldsTab128(tb + " + std::to_string(16 * tabOff.at(w)) + "u)" : u4(prog[w]); }

    // ---- shape functions --------------------------------------------------------------------------------------------
    // Subtrees of the same SHAPE (same operators, same kinds of leaves, different constants: the eight groups of
    // six transformed spheres cut by two planes of BASELINE config 4) share one out-of-line function that reads its
    // leaves' constants from a table in constant memory.  The code of a 70-leaf scene then is a dozen leaves long
    // and stays in the instruction cache (the first generation's 54 KB of straight-line code per kernel did not:
    // `no_instruction` 15-36 stall cycles per issue, profiles/r01m_spec_c4.txt).
    std::vector<std::string> sig; // per node
    std::vector<int> leaves;      // per node
    std::map<std::string, int> shapeFn;    // chosen signature -> function number
    std::map<std::string, int> shapeProto; // chosen signature -> the node whose code becomes the function's body
    std::vector<std::string> fnDefs;
    std::vector<bool> fnDone;
    std::vector<size_t> fnBounds; // bounding-sphere words an instance of each function brings (0: the function has no test)
    std::vector<uint4> table; // all instances' constants
    bool inFn = false;

    size_t leafWords(size_t w) const { return (prog[w].x & FK_SPHERE_XF) ? 4u : (prog[w].x & FK_PLANE) ? 2u : 2u; }
    void collectWords(int id, std::vector<size_t> &out) const
    { // the records a subtree's code reads, in the order of its code
        const GenNode &n = nodes[id];
        if(n.left < 0)
        {
            const size_t w = (size_t)tree.items[n.item].leafWord;
            for(size_t k = 0; k < leafWords(w); k++)
                out.push_back(w + k);
            return;
        }
        if(isCluster(id))
        {
            const size_t w = (size_t)tree.items[n.firstItem].clusterWord;
            for(size_t k = 0; k < 2u + 2u * prog[w].z; k++)
                out.push_back(w + k);
        }
        collectWords(n.left, out);
        collectWords(n.right, out);
    }
    void computeShapes()
    {
        sig.resize(nodes.size()), leaves.resize(nodes.size());
        for(size_t id = 0; id < nodes.size(); id++)
        { // (children precede their parents in `nodes`)
            const GenNode &n = nodes[id];
            if(n.left < 0)
            {
                const uint32_t x = prog[(size_t)tree.items[n.item].leafWord].x;
                sig[id] = (x & FK_SPHERE_XF) ? "S" : (x & FK_PLANE) ? "P" : "s";
                leaves[id] = 1;
                continue;
            }
            const int kind = tree.items[n.item].kind;
            sig[id] = std::string(isCluster((int)id) ? "C" : "") + (kind == PTB_OP_UNION ? "U(" : kind == PTB_OP_DIFFERENCE ? "D(" : "I(") + sig[n.left] + "," + sig[n.right] + ")";
            leaves[id] = leaves[n.left] + leaves[n.right];
        }
    }
    // Chooses the shapes: most instances first, larger first among equals; a shape is skipped when one of its
    // instances contains, or lies inside, an instance of a shape already chosen.
    void chooseShapes(int root, int minLeaves)
    {
        std::map<std::string, std::vector<int>> byShape;
        for(size_t id = 0; id < nodes.size(); id++)
            if(leaves[id] >= minLeaves && (int)id != root)
                byShape[sig[id]].push_back((int)id);
        std::vector<std::pair<std::pair<int, int>, std::string>> order;
        for(auto &kv : byShape)
            if(kv.second.size() >= 2)
                order.push_back({{-(int)kv.second.size(), -leaves[kv.second[0]]}, kv.first});
        std::sort(order.begin(), order.end());
        std::vector<int> covered(nodes.size(), 0); // 1: inside a chosen instance, 2: contains one
        std::vector<int> parent(nodes.size(), -1);
        for(size_t id = 0; id < nodes.size(); id++)
            if(nodes[id].left >= 0)
                parent[nodes[id].left] = parent[nodes[id].right] = (int)id;
        std::function<void(int)> markInside = [&](int id) {
            covered[id] |= 1;
            if(nodes[id].left >= 0)
                markInside(nodes[id].left), markInside(nodes[id].right);
        };
        for(auto &o : order)
        {
            const std::vector<int> &inst = byShape[o.second];
            bool ok = true;
            for(int id : inst)
                ok = ok && covered[id] == 0;
            if(!ok)
                continue;
            shapeFn[o.second] = (int)shapeFn.size();
            shapeProto[o.second] = inst[0];
            for(int id : inst)
            {
                markInside(id);
                for(int up = parent[id]; up >= 0; up = parent[up])
                    covered[up] |= 2;
            }
        }
        fnDone.assign(shapeFn.size(), false);
        fnBounds.assign(shapeFn.size(), 0);
        fnDefs.assign(shapeFn.size(), std::string());
    }
    int ensureFn(const std::string &sg)
    {
        const int f = shapeFn.at(sg), proto = shapeProto.at(sg);
        if(fnDone[f])
            return f;
        fnDone[f] = true;
        std::string keep;
        keep.swap(s);
        std::vector<size_t> words;
        collectWords(proto, words);
        tabOff.clear();
        for(size_t k = 0; k < words.size(); k++)
            tabOff[words[k]] = (int)k;
        tableMode = inFn = true;
        s += "// shape " + sg + "\n";
        s += "template <bool O0>\nstatic __device__ __noinline__ uint4 ptbShape" + std::to_string(f) +
             "(uint32_t poolBase, int cap, int top, float ox, float oy, float oz, float dx, float dy, float dz, uint32_t tb, uint32_t bt)\n{\n";
        { // Camera rays, bt != 0 (a call site all lanes reach together): the bounding spheres the instance's result lies
          // within, boundsN words at shared address bt.  No ray of the warp meets any of them: the result is empty.  The
          // test sits HERE, not at the call sites, so that the code every warp runs eight times over is these few dozen
          // instructions and stays in the instruction cache.
            std::vector<uint4> protoBounds;
            const size_t nb = boundsOf(proto, protoBounds) ? protoBounds.size() : 0;
            fnBounds[f] = nb;
            if(nb > 0 && nb <= 32)
            {
                line(1, "if(O0 && bt != 0u)");
                line(1, "{");
                line(2, "const V3 dv = mk(dx, dy, dz);");
                line(2, "const float sd = Ar<false>::sqrt(fmaf(dx, dx, fmaf(dy, dy, dz * dz)));");
                line(2, "uint32_t met = 0u;");
                line(2, "for(uint32_t k = 0u; k < " + std::to_string(nb) + "u; k++)");
                line(3, "met |= rlBoundMet(ldsTab128(bt + 16u * k), dv, sd) ? 1u : 0u;");
                line(2, "if(__reduce_or_sync(0xFFFFFFFFu, met) == 0u)");
                line(3, "return make_uint4(0u, 0u, 0u, 0u);");
                line(1, "}");
            }
            else
                fnBounds[f] = 0;
        }
        line(1, "REval ev;");
        line(1, "ev.poolBase = poolBase, ev.cap = cap, ev.top = top, ev.ovf = false, ev.tab = 0u;");
        line(1, "const V3 o = mk(ox, oy, oz), d = mk(dx, dy, dz);");
        node(proto, 1, true, true);
        const std::string r = var(proto);
        line(1, "return make_uint4((uint32_t)" + r + ".n | (ev.ovf ? 0x80000000u : 0u), __float_as_uint(" + r + ".a.t0), __float_as_uint(" + r + ".a.t1), " + r + ".a.tags);");
        s += "}\n";
        alias[proto] = proto;
        tableMode = inFn = false;
        fnDefs[f] = s;
        s.swap(keep);
        return f;
    }
    // ---- bounding spheres (camera rays) ------------------------------------------------------------------------------
    // {C, sqrt(|C|^2 - R^2)} per sphere leaf: what compileFastProgram stores behind FK_END for the leaves of sphere
    // runs (ptb_capi.cu), computed here for untransformed spheres.  boundsOf() lists the spheres a subtree's result
    // lies within — {} when EVERY line that misses all of them leaves the subtree empty: a Union needs both operands
    // bounded, a Difference its left one, an Intersection either.
    std::map<size_t, uint4> runBound; // header word of a transformed-sphere leaf -> its bounding-sphere word
    int divergent = 0;                // > 0: the code being emitted runs under a per-lane condition
    bool likelyHints = false;         // __builtin_expect on the plane clusters' register form (PTB_SPEC_HINTS=0: none)
    bool leanMerges = false;          // small programs: only the closed forms of two single spans out of line (rlMerge11Call)
    bool coldMerges = false;          // large programs: operators call ONE out-of-line merge (rlMergeSlow, ptb_csg_reg.cuh)
    bool usesBounds = false;
    void indexRunBounds()
    {
        for(const FastTreeItem &it : tree.items)
        {
            if(it.kind != PTB_OP_PUSH_PRIM || it.leafWord < 0)
                continue;
            const size_t w = (size_t)it.leafWord;
            const uint32_t x = prog[w].x, table = (x >> 16) & 0x3FFFu;
            if((x & FK_SPHERE_XF) && (x & (FH_RUN | FH_TAIL)) == FH_RUN && table != 0u)
                for(uint32_t k = 0; k < ((x >> 12) & 15u); k++)
                    runBound[w + 4u * k] = prog[table + k];
        }
    }
    bool boundsOf(int id, std::vector<uint4> &out) const
    {
        const GenNode &n = nodes[id];
        if(n.left < 0)
        {
            const size_t w = (size_t)tree.items[n.item].leafWord;
            const uint4 h = prog[w];
            if(h.x & FK_SPHERE_XF)
            {
                auto it = runBound.find(w);
                if(it == runBound.end())
                    return false;
                out.push_back(it->second);
                return true;
            }
            if(h.x & FK_SPHERE)
            { // header {x, tags, -r^2, c0}, then (-c.xyz, 0)
                auto f = [](uint32_t u) {
                    float v;
                    memcpy(&v, &u, 4);
                    return (double)v;
                };
                const double cx = -f(prog[w + 1].x), cy = -f(prog[w + 1].y), cz = -f(prog[w + 1].z), r2 = -f(h.z) * 1.002 + 1e-9;
                const double K = cx * cx + cy * cy + cz * cz - r2;
                const float c[4] = {(float)cx, (float)cy, (float)cz, (float)(K > 0 ? sqrt(K) * (1.0 - 1e-5) : 0.0)};
                uint4 word;
                memcpy(&word, c, 16);
                out.push_back(word);
                return true;
            }
            return false; // a plane
        }
        const int kind = tree.items[n.item].kind;
        if(kind == PTB_OP_UNION)
            return boundsOf(n.left, out) && boundsOf(n.right, out);
        const size_t mark = out.size();
        if(boundsOf(n.left, out))
            return true;
        out.resize(mark);
        if(kind == PTB_OP_INTERSECTION && boundsOf(n.right, out))
            return true;
        out.resize(mark);
        return false;
    }

    // A subtree of a chosen shape: its constants go to the table, its code is a call.  The callee starts with an
    // empty stack of its own on top of ours, so a result of two or more spans lies at our `top` (ptb_csg_reg.cuh).
    void emitCall(int id, int indent)
    {
        const int f = ensureFn(sig[id]);
        std::vector<size_t> words;
        collectWords(id, words);
        const size_t base = table.size();
        for(size_t w : words)
            table.push_back(prog[w]);
        const std::string L = "L" + std::to_string(id);
        // Camera rays: the callee first votes on the bounding spheres the instance's result lies within (see ensureFn);
        // their words follow the instance's constants in the table.  Only where all lanes arrive together.
        std::vector<uint4> bounds;
        const bool bounded = divergent == 0 && !inFn && getenv("PTB_SPEC_BOUNDS") == nullptr && fnBounds[f] > 0 && boundsOf(id, bounds) && bounds.size() == fnBounds[f];
        const size_t bbase = table.size();
        if(bounded)
        {
            usesBounds = true;
            for(const uint4 &b : bounds)
                table.push_back(b);
        }
        line(indent, "RList " + L + ";");
        line(indent, "{");
        line(indent + 1, "const uint4 r_ = ptbShape" + std::to_string(f) + "<O0>(ev.poolBase, ev.cap, ev.top, o.x, o.y, o.z, d.x, d.y, d.z, ev.tab + " + std::to_string(16 * base) + "u, " +
                             (bounded ? "ev.tab + " + std::to_string(16 * bbase) + "u" : "0u") + ");");
        line(indent + 1, L + ".n = (int)(r_.x & 0x7FFFFFFFu), " + L + ".a = rlSpan(__uint_as_float(r_.y), __uint_as_float(r_.z), r_.w), " + L + ".base = ev.top;");
        line(indent + 1, "ev.ovf = ev.ovf || (r_.x >> 31) != 0u;");
        line(indent + 1, "if(" + L + ".n >= 2)");
        line(indent + 2, "ev.top += " + L + ".n;");
        line(indent, "}");
    }

    // ---- cold blocks ---------------------------------------------------------------------------------------------
    // A plane cluster whose half-lines cross (or a ray lying in one of its planes) is merged leaf by leaf like any
    // other subtree: a hundred instructions per plane that a render hardly ever executes, in the middle of the code
    // every ray runs.  PTB_SPEC_COLD_BLOCKS=1 moves them to functions of their own behind the evaluator.  Off by
    // default — measured on B200 (tools/gpu_r02I.sh): C5 bounce trace 86.7 / 86.7 / 92.5 ms per frame against 89.4,
    // C4 camera trace 332 against 309, C3 8.93 against 9.35 Grays/s: within what the placement of a module in memory
    // does to these kernels from one process to the next (C5 bounce trace: 89 or 100 ms, same code, same box).
    bool coldBlocks = false;
    std::vector<std::string> coldDefs;
    int coldCluster(int id)
    {
        const int c = (int)coldDefs.size();
        coldDefs.push_back(std::string());
        std::string keep;
        keep.swap(s);
        s += "template <bool O0>\nstatic __device__ __noinline__ RMerged ptbCold" + std::to_string(c) +
             "(uint32_t poolBase, int cap, int top, float ox, float oy, float oz, float dx, float dy, float dz)\n{\n";
        line(1, "REval ev;");
        line(1, "ev.poolBase = poolBase, ev.cap = cap, ev.top = top, ev.ovf = false, ev.tab = 0u;");
        line(1, "const V3 o = mk(ox, oy, oz), d = mk(dx, dy, dz);");
        divergent++;
        node(id, 1, false, true); // declares L<id>_sub
        divergent--;
        line(1, "RMerged r;");
        line(1, "r.L = L" + std::to_string(id) + "_sub, r.top = ev.top, r.ovf = ev.ovf;");
        line(1, "return r;");
        s += "}\n";
        coldDefs[c] = s;
        s.swap(keep);
        return c;
    }

    // code of one plane record -> RList (a single plane, or the planes of a cluster folded in registers)
    void planeFoldCalls(int indent, size_t w, uint32_t extra)
    {
        line(indent, "float loEnd = -3.0e38f, upStart = 3.0e38f, lt0, lt1, ut0;");
        line(indent, "uint32_t loTags = 0u, upTags = 0u, ltags, utags;");
        line(indent, "bool full = false, hit, two;");
        line(indent, "planeFold<O0>(" + W(w + 1) + ", " + W(w) + ".y, o, d, loEnd, upStart, loTags, upTags, full);");
        for(uint32_t j = 0; j < extra; j++)
            line(indent, "planeFold<O0>(" + W(w + 3 + 2 * j) + ", " + W(w + 2 + 2 * j) + ".x, o, d, loEnd, upStart, loTags, upTags, full);");
    }
    void leaf(int id, int indent)
    {
        const FastTreeItem &it = tree.items[nodes[id].item];
        const size_t w = (size_t)it.leafWord;
        const uint4 h = prog[w];
        line(indent, "RList " + var(id) + ";");
        line(indent, "{");
        if(h.x & FK_PLANE)
        { // PlaneSpanIterator::init, src/plane.cpp:35-63
            planeFoldCalls(indent + 1, w, 0u);
            line(indent + 1, "planeFinish(loEnd, upStart, loTags, upTags, full, hit, two, lt0, lt1, ut0, ltags, utags);");
            line(indent + 1, var(id) + " = rlCluster(ev, hit, two, lt0, lt1, ut0, ltags, utags);");
        }
        else
        { // SphereSpanIterator::init, src/sphere.cpp:31-49
            line(indent + 1, "float a, b, e, t0 = 0.f, t1 = 0.f;");
            line(indent + 1, "const uint4 k0 = " + W(w) + ";");
            if(h.x & FK_SPHERE_XF)
                line(indent + 1, "xfDisc<O0>(k0, " + W(w + 1) + ", " + W(w + 2) + ", " + W(w + 3) + ", o, d, a, b, e);");
            else
                line(indent + 1, "sphereDisc<O0>(k0, " + W(w + 1) + ", o, d, a, b, e);");
            line(indent + 1, "if(e > PTB_EPS)");
            line(indent + 1, "{");
            line(indent + 2, "const float q = sqrtApprox(e), ia = rcpApprox(a);");
            line(indent + 2, "t0 = (-b - q) * ia, t1 = (-b + q) * ia;");
            line(indent + 1, "}");
            line(indent + 1, var(id) + " = rlLeaf(e > PTB_EPS, t0, t1, k0.y);");
        }
        line(indent, "}");
    }
    // The right operand of the root Union, a leaf or a plane cluster, as registers: nB spans b0, b1 (nB = -1: a
    // cluster whose half-lines cross, to be evaluated leaf by leaf).
    void rootRight(int id, int indent)
    {
        const GenNode &n = nodes[id];
        line(indent, "int nB = -1;");
        line(indent, "SpanRec b0 = rlSpan(0.f, 0.f, 0u), b1 = b0;");
        line(indent, "{");
        const FastTreeItem &first = tree.items[n.firstItem];
        const size_t w = (size_t)(n.left < 0 ? first.leafWord : first.clusterWord);
        const uint4 h = prog[w];
        if(h.x & FK_PLANE)
        {
            const uint32_t extra = n.left < 0 ? 0u : h.z;
            planeFoldCalls(indent + 1, w, extra);
            if(extra >= 2u)
                line(indent + 1, "if(!full && loEnd < upStart)");
            else if(extra == 1u)
                line(indent + 1, "if(!full)");
            line(indent + 1, "{");
            line(indent + 2, "planeFinish(loEnd, upStart, loTags, upTags, full, hit, two, lt0, lt1, ut0, ltags, utags);");
            line(indent + 2, "nB = hit ? (two ? 2 : 1) : 0;");
            line(indent + 2, "b0 = rlSpan(lt0, lt1, ltags), b1 = rlSpan(ut0, PTB_MAX_VALUE, utags);");
            line(indent + 1, "}");
        }
        else
        {
            line(indent + 1, "float a, b, e;");
            if(h.x & FK_SPHERE_XF)
                line(indent + 1, "xfDisc<O0>(" + u4(h) + ", " + u4(prog[w + 1]) + ", " + u4(prog[w + 2]) + ", " + u4(prog[w + 3]) + ", o, d, a, b, e);");
            else
                line(indent + 1, "sphereDisc<O0>(" + u4(h) + ", " + u4(prog[w + 1]) + ", o, d, a, b, e);");
            line(indent + 1, "nB = 0;");
            line(indent + 1, "if(e > PTB_EPS)");
            line(indent + 1, "{");
            line(indent + 2, "const float q = sqrtApprox(e), ia = rcpApprox(a);");
            line(indent + 2, "nB = 1, b0 = rlSpan((-b - q) * ia, (-b + q) * ia, " + hex(h.y) + ");");
            line(indent + 1, "}");
        }
        line(indent, "}");
    }
    void node(int id, int indent, bool asCluster = true, bool inlineThis = false)
    {
        const GenNode &n = nodes[id];
        if(n.left < 0)
        {
            leaf(id, indent);
            return;
        }
        if(!inlineThis && !inFn && !sig.empty() && shapeFn.count(sig[id]))
        {
            emitCall(id, indent);
            return;
        }
        if(asCluster && isCluster(id))
        { // a Union-only subtree of planes: folded in registers while every lower end lies below every upper start
          // (compileFastProgram has the proof), else its leaves are merged one by one like any other subtree
            const size_t w = (size_t)tree.items[n.firstItem].clusterWord;
            const uint32_t extra = prog[w].z;
            line(indent, "RList " + var(id) + ";");
            line(indent, "{");
            planeFoldCalls(indent + 1, w, extra);
            line(indent + 1, std::string(likelyHints ? "if(PTB_LIKELY(!full && (" : "if((!full && (") + (extra == 1u ? "true" : "loEnd < upStart") + ")))");
            line(indent + 1, "{");
            line(indent + 2, "planeFinish(loEnd, upStart, loTags, upTags, full, hit, two, lt0, lt1, ut0, ltags, utags);");
            line(indent + 2, var(id) + " = rlCluster(ev, hit, two, lt0, lt1, ut0, ltags, utags);");
            line(indent + 1, "}");
            line(indent + 1, "else");
            line(indent + 1, "{");
            if(!inFn && coldBlocks)
            { // (hardly ever taken and as long as the rest of a small scene's code: out of line, see coldCluster)
                const int c = coldCluster(id);
                line(indent + 2, "const RMerged r_ = ptbCold" + std::to_string(c) + "<O0>(ev.poolBase, ev.cap, ev.top, o.x, o.y, o.z, d.x, d.y, d.z);");
                line(indent + 2, "L" + std::to_string(id) + " = r_.L, ev.top = r_.top, ev.ovf = ev.ovf || r_.ovf;");
            }
            else
            {
                divergent++;
                node(id, indent + 2, false, true); // declares L<id>_sub
                divergent--;
                line(indent + 2, "L" + std::to_string(id) + " = L" + std::to_string(id) + "_sub;");
            }
            line(indent + 1, "}");
            line(indent, "}");
            return;
        }
        const int kind = tree.items[n.item].kind;
        node(n.left, indent);
        if(kind == PTB_OP_UNION)
            node(n.right, indent);
        else
        { // {} \ B = {} n B = {}: the right operand of a lane whose left one is empty is never evaluated
            const std::string rname = "L" + std::to_string(n.right) + "_r"; // (explicit: var() of the operand changes once it aliases its own left child)
            line(indent, "RList " + rname + " = rlEmpty();");
            line(indent, "if(" + var(n.left) + ".n != 0)");
            line(indent, "{");
            divergent++;
            node(n.right, indent + 1);
            divergent--;
            line(indent + 1, rname + " = " + var(n.right) + ";");
            line(indent, "}");
        }
        const std::string rhs = kind == PTB_OP_UNION ? var(n.right) : "L" + std::to_string(n.right) + "_r";
        const char *fn = kind == PTB_OP_UNION ? "rlUnionInto" : kind == PTB_OP_DIFFERENCE ? "rlDifferenceInto" : "rlIntersectionInto";
        line(indent, std::string(fn) + (coldMerges ? "Cold" : leanMerges ? "Lean" : "") + "(ev, " + var(n.left) + ", " + rhs + ");");
        if(asCluster)
            alias[id] = n.left; // from here on this node's list IS the left operand's variable
        else
            line(indent, "RList L" + std::to_string(id) + "_sub = " + var(n.left) + ";");
    }
};
} // namespace

static std::string specializeSourceGen2(const std::vector<uint4> &prog, const FastTree &tree, int which)
{
    Gen g(prog, tree);
    std::vector<int> stack;
    for(size_t i = 0; i < tree.items.size(); i++)
    {
        GenNode n;
        n.item = (int)i, n.left = n.right = -1, n.firstItem = (int)i;
        if(tree.items[i].kind != PTB_OP_PUSH_PRIM)
        {
            n.right = stack.back();
            stack.pop_back();
            n.left = stack.back();
            stack.pop_back();
            n.firstItem = g.nodes[n.left].firstItem;
        }
        g.nodes.push_back(n);
        stack.push_back((int)g.nodes.size() - 1);
    }
    const int root = stack.back();
    g.alias.resize(g.nodes.size());
    for(size_t i = 0; i < g.alias.size(); i++)
        g.alias[i] = (int)i;
    { // shape functions for large scenes with repeated subtrees (PTB_SPEC_SHAPES=0 never, =1 whatever the size)
        const char *e = getenv("PTB_SPEC_SHAPES");
        const int mode = e ? atoi(e) : -1;
        int nLeaves = 0;
        for(const FastTreeItem &it : tree.items)
            nLeaves += it.kind == PTB_OP_PUSH_PRIM;
        // PTB_SPEC_COLD=0 / 1 / 2: operators inline / through the one out-of-line merge / closed forms of two single
        // spans out of line, whatever the size.  Measured on B200 (tools/gpu_r02J.sh, ms per frame): C5 bounce trace
        // inline 89.4 / 100.3 / 89.5 in three processes (the placement of the module decides), lean 92.0 / 92.2 / 91.9,
        // cold 97; C3 9.33 -> 10.04 Grays/s lean; C4 camera trace cold 313, lean 337.
        g.coldMerges = getenv("PTB_SPEC_COLD") ? atoi(getenv("PTB_SPEC_COLD")) == 1 : nLeaves >= 24;
        g.leanMerges = getenv("PTB_SPEC_COLD") ? atoi(getenv("PTB_SPEC_COLD")) == 2 : !g.coldMerges;
        g.likelyHints = getenv("PTB_SPEC_HINTS") ? atoi(getenv("PTB_SPEC_HINTS")) != 0 : g.coldMerges;
        g.coldBlocks = getenv("PTB_SPEC_COLD_BLOCKS") && atoi(getenv("PTB_SPEC_COLD_BLOCKS")) != 0;
        if(mode != 0 && (mode == 1 || nLeaves >= 24))
        {
            g.indexRunBounds();
            g.computeShapes();
            g.chooseShapes(root, mode == 1 ? 3 : 4);
        }
    }
    std::string s;
    if(const char *e = getenv("PTB_SPEC_MIN_CTAS"))
        s += "#define PTB_TRACE_MIN_CTAS " + std::to_string(atoi(e)) + "\n";
    s += "#include \"ptb200.h\"\n#include \"ptb_csg_reg.cuh\"\nnamespace ptb\n{\n";
    const GenNode &rn = g.nodes[root];
    const bool fusedRoot = which == SPEC_TRACE && rn.left >= 0 && tree.items[rn.item].kind == PTB_OP_UNION && !g.isCluster(root) &&
                           (g.nodes[rn.right].left < 0 || g.isCluster(rn.right));
    if(which == SPEC_TRACE)
    { // the renderer's trace kernels only want the first boundary with t >= eps
        g.s += "template <bool O0>\n__device__ __forceinline__ int specTrace(REval &ev, const V3 o, const V3 d, float &t, uint32_t &tag)\n{\n";
        g.line(1, "const float sd_ = O0 ? Ar<false>::sqrt(fmaf(d.x, d.x, fmaf(d.y, d.y, d.z * d.z))) : 0.f; // |d| for the bounding-sphere tests (unused: dropped)");
        if(fusedRoot)
        { // the root Union is scanned while it is merged; its right operand (a leaf or a plane cluster) stays in registers
            g.node(rn.left, 1);
            g.rootRight(rn.right, 1);
            g.line(1, g.likelyHints ? "if(PTB_LIKELY(nB >= 0))" : "if(nB >= 0)");
            g.line(2, "return rlUnionFirstHit(ev, " + g.var(rn.left) + ", nB, b0, b1, t, tag);");
            if(g.isCluster(rn.right) && prog[(size_t)tree.items[g.nodes[rn.right].firstItem].clusterWord].z >= 2u)
            { // a cluster of three or more planes whose half-lines cross: its leaves one by one, then the root as a list
                if(g.coldBlocks)
                {
                    const int c = g.coldCluster(rn.right);
                    g.line(1, "const RMerged r_ = ptbCold" + std::to_string(c) + "<O0>(ev.poolBase, ev.cap, ev.top, o.x, o.y, o.z, d.x, d.y, d.z);");
                    g.line(1, "ev.top = r_.top, ev.ovf = ev.ovf || r_.ovf;");
                    g.line(1, "const RList Lroot = rlUnion(ev, " + g.var(rn.left) + ", r_.L);");
                }
                else
                {
                    g.node(rn.right, 1, false);
                    g.line(1, "const RList Lroot = rlUnion(ev, " + g.var(rn.left) + ", " + "L" + std::to_string(rn.right) + "_sub);");
                }
                g.line(1, "return rlFirstHit(ev, Lroot, t, tag);");
            }
            else
                g.line(1, "return PTB_HIT_MISS; // (not reached)");
        }
        else
        {
            g.node(root, 1);
            g.line(1, "return rlFirstHit(ev, " + g.var(root) + ", t, tag);");
        }
        g.s += "}\n";
    }
    else
    { // ray batches want the whole root list: pool slots [0, n)
        g.s += "template <bool O0>\n__device__ __forceinline__ int specList(REval &ev, const V3 o, const V3 d)\n{\n";
        g.line(1, "const float sd_ = O0 ? Ar<false>::sqrt(fmaf(d.x, d.x, fmaf(d.y, d.y, d.z * d.z))) : 0.f;");
        g.node(root, 1);
        g.line(1, "return rlToPool(ev, " + g.var(root) + ");");
        g.s += "}\n";
    }
    if(!g.table.empty())
    {
        s += "__constant__ uint4 ptbTab[" + std::to_string(g.table.size()) + "] = {\n";
        for(const uint4 &w : g.table)
            s += "    {" + hex(w.x) + ", " + hex(w.y) + ", " + hex(w.z) + ", " + hex(w.w) + "},\n";
        s += "};\n";
        s += "__device__ __forceinline__ void specStageTable(uint32_t tab)\n{\n    for(uint32_t i = threadIdx.x; i < " + std::to_string(g.table.size()) +
             "u; i += PTB_FAST_THREADS)\n        sts128(tab + 16u * i, ptbTab[i]);\n}\n";
    }
    else
        s += "__device__ __forceinline__ void specStageTable(uint32_t) {}\n";
    if(g.usesBounds)
        s += "// PTB_SPEC_BOUNDED_CALLS: shape-function calls of camera rays are guarded by bounding-sphere votes\n";
    for(const std::string &f : g.fnDefs)
        s += f;
    for(const std::string &f : g.coldDefs)
        s += f;
    s += g.s;
    s += "} // namespace ptb\n";
    s += "#define PTB_SPEC_GEN 2\n#include \"ptb_trace.cuh\"\n";
    if(which == SPEC_TRACE)
    {
        s += "extern \"C\" __global__ void __launch_bounds__(PTB_FAST_THREADS, PTB_TRACE_MIN_CTAS) ptbSpecTracePrimary(ptb::DevScene sc, int cap, float *overflow, size_t ovfStride, "
             "ptb::RenderPass pass, uint2 *hits)\n{\n    ptb::tracePrimaryBody<true>(sc, cap, overflow, ovfStride, pass, hits);\n}\n";
        s += "extern \"C\" __global__ void __launch_bounds__(PTB_FAST_THREADS, PTB_TRACE_MIN_CTAS) ptbSpecTraceQueue(ptb::DevScene sc, int cap, float *overflow, size_t ovfStride, "
             "ptb::RayQueue in, const unsigned *inCount, uint2 *hits)\n{\n    ptb::traceQueueBody<true>(sc, cap, overflow, ovfStride, in, inCount, hits);\n}\n";
    }
    else
        for(int o0 = 0; o0 < 2; o0++)
            s += std::string("extern \"C\" __global__ void __launch_bounds__(PTB_FAST_THREADS) ptbSpecIntersect") + (o0 ? "O0" : "") +
                 "(ptb::DevScene sc, int cap, float *overflow, size_t ovfStride, const float4 *origins, const float4 *dirs, size_t n, PtbHit *hits, PtbSpan *spans, int maxSpans)\n{\n"
                 "    ptb::intersectFastBody<true, " +
                 (o0 ? "true" : "false") + ">(sc, cap, overflow, ovfStride, origins, dirs, n, hits, spans, maxSpans);\n}\n";
    return s;
}

std::string specializeSource(const std::vector<uint4> &prog, const FastTree &tree, int which)
{
    static const int gen = [] {
        const char *e = getenv("PTB_SPEC_GEN");
        return e ? atoi(e) : 2;
    }();
    if(gen == 1 || tree.items.empty())
        return specializeSourceGen1(prog, which);
    return specializeSourceGen2(prog, tree, which);
}

int specializeCompile(const std::string &source, std::vector<char> &cubin, std::string &log)
{
    { // the lock covers loading the library only: a render thread that polls for the kernels while another
      // thread compiles must not wait for the compile (NVRTC compiles separate programs concurrently)
        std::lock_guard<std::mutex> g(g_lock);
        if(!loadNvrtc())
        {
            log = g_libs.why;
            return -1;
        }
    }
    void *prog = nullptr;
    int rc = g_libs.nvrtcCreateProgram(&prog, source.c_str(), "ptb_scene.cu", kEmbeddedCount, kEmbeddedSources, kEmbeddedNames);
    if(rc != 0)
    {
        log = "nvrtcCreateProgram failed (" + std::to_string(rc) + ")";
        return -1;
    }
    // -lineinfo (for ncu source pages) costs compile time: only on request.  --split-compile=0: the four
    // kernels are optimised on as many threads as the host has.
    std::vector<const char *> opts = {"--gpu-architecture=sm_100a", "-std=c++17", "--split-compile=0"};
    if(getenv("PTB_SPEC_LINEINFO"))
        opts.push_back("-lineinfo");
    rc = g_libs.nvrtcCompileProgram(prog, (int)opts.size(), opts.data());
    size_t n = 0;
    if(g_libs.nvrtcGetProgramLogSize(prog, &n) == 0 && n > 1)
    {
        log.resize(n);
        g_libs.nvrtcGetProgramLog(prog, &log[0]);
    }
    if(rc != 0)
    {
        log = "NVRTC compile failed (" + std::to_string(rc) + "): " + log;
        g_libs.nvrtcDestroyProgram(&prog);
        return -1;
    }
    n = 0;
    g_libs.nvrtcGetCUBINSize(prog, &n);
    cubin.resize(n);
    g_libs.nvrtcGetCUBIN(prog, cubin.data());
    g_libs.nvrtcDestroyProgram(&prog);
    if(const char *dump = getenv("PTB_SPEC_DUMP"))
        if(FILE *f = fopen(dump, "wb"))
        { // for cuobjdump -sass / -res-usage
            fwrite(cubin.data(), 1, cubin.size(), f);
            fclose(f);
        }
    return cubin.empty() ? -1 : 0;
}

static std::string cuError(int rc)
{
    const char *msg = nullptr;
    if(g_libs.cuGetErrorString && g_libs.cuGetErrorString(rc, &msg) == 0 && msg)
        return msg;
    return "CUDA driver error " + std::to_string(rc);
}

// ---- cache: one entry per distinct generated source, compiled once per process ------------------------
namespace
{
struct Entry
{
    int state = 0; // 0 compiling, 1 cubin ready, 2 failed
    std::vector<char> cubin;
    std::string log;
    double compileMs = 0.0;
    std::map<int, SpecKernels *> loaded; // per device
};
// never destroyed: a compile thread may still be running when the process exits
std::mutex &cacheLock()
{
    static std::mutex *m = new std::mutex();
    return *m;
}
std::map<std::string, Entry *> &cache()
{
    static std::map<std::string, Entry *> *c = new std::map<std::string, Entry *>();
    return *c;
}
// Compiles in flight.  A process that exits while NVRTC is still working on another thread crashes in
// the library's teardown, so an exit handler — registered AFTER libnvrtc was loaded, hence run before
// its destructors — waits for them (a second or two at most).
int g_inFlight = 0;
void waitForCompiles()
{
    for(;;)
    {
        {
            std::lock_guard<std::mutex> g(cacheLock());
            if(g_inFlight == 0)
                return;
        }
        std::this_thread::sleep_for(std::chrono::milliseconds(5));
    }
}
void compileEntry(Entry *e, std::string source)
{
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<char> cubin;
    std::string log;
    int rc = specializeCompile(source, cubin, log);
    std::lock_guard<std::mutex> g(cacheLock());
    e->compileMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    e->log = log;
    e->cubin.swap(cubin);
    e->state = rc == 0 ? 1 : 2;
    g_inFlight--;
}
} // namespace

const SpecKernels *specializeGet(const std::string &source, int device, int mode, std::string &err)
{
    Entry *e = nullptr;
    bool compileHere = false;
    if(mode != SPEC_PEEK)
    { // load NVRTC on this thread, then (once) register the exit handler: see waitForCompiles
        std::lock_guard<std::mutex> g(g_lock);
        static bool registered = false;
        if(loadNvrtc() && !registered)
        {
            registered = true;
            atexit(waitForCompiles);
        }
    }
    {
        std::lock_guard<std::mutex> g(cacheLock());
        auto it = cache().find(source);
        if(it != cache().end())
            e = it->second;
        else if(mode != SPEC_PEEK)
        {
            // Bounded: a process that renders an endless stream of distinct scenes (an animation, one scene per frame)
            // neither piles up cubins and loaded modules nor stacks compiles.  Past kMaxEntries programs new ones stay
            // interpreted (entries are never evicted: a scene may be launching their kernels); a background compile is
            // started only while no other one is running (every compile uses all host cores: --split-compile=0).
            const size_t kMaxEntries = 64;
            if(cache().size() >= kMaxEntries)
            {
                err = "specialised-kernel cache full (64 scene programs in this process)";
                return nullptr;
            }
            if(mode == SPEC_START && g_inFlight > 0)
                return nullptr; // try again at the next pass / call
            e = new Entry();
            cache()[source] = e;
            g_inFlight++;
            if(mode == SPEC_START)
                std::thread(compileEntry, e, source).detach();
            else
                compileHere = true;
        }
    }
    if(!e)
        return nullptr;
    if(compileHere)
        compileEntry(e, source);
    if(mode == SPEC_WAIT)
        for(;;)
        { // somebody else's compile: wait for it
            {
                std::lock_guard<std::mutex> g(cacheLock());
                if(e->state != 0)
                    break;
            }
            std::this_thread::sleep_for(std::chrono::milliseconds(5));
        }
    std::lock_guard<std::mutex> g(cacheLock());
    if(e->state == 0)
        return nullptr; // still compiling: the caller keeps the interpreter for now
    if(e->state == 2)
    {
        err = e->log;
        return nullptr;
    }
    auto it = e->loaded.find(device);
    if(it != e->loaded.end())
    {
        if(!it->second)
            err = "loading the specialised kernels failed before on this device";
        return it->second;
    }
    // first use on this device: load the cubin into the current (primary) context
    SpecKernels *k = nullptr;
    void *module = nullptr;
    if(!loadDriver())
    {
        err = g_libs.why;
        e->loaded[device] = nullptr;
        return nullptr;
    }
    int rc = g_libs.cuModuleLoadData(&module, e->cubin.data());
    if(rc != 0)
        err = "cuModuleLoadData: " + cuError(rc);
    else
    {
        k = new SpecKernels();
        k->module = module;
        struct
        {
            void **fn;
            const char *name;
        } want[] = {{&k->tracePrimary, "ptbSpecTracePrimary"}, {&k->traceQueue, "ptbSpecTraceQueue"}, {&k->intersect, "ptbSpecIntersect"}, {&k->intersectO0, "ptbSpecIntersectO0"}};
        for(auto &x : want)
            if(g_libs.cuModuleGetFunction(x.fn, module, x.name) != 0)
                *x.fn = nullptr; // the other program holds this one
        if(!k->tracePrimary && !k->intersect)
        {
            err = "the compiled module holds none of the expected kernels";
            delete k;
            k = nullptr;
        }
        if(k)
        {
            g_libs.cuFuncGetAttribute(&k->regsTrace, 4 /* CU_FUNC_ATTRIBUTE_NUM_REGS */, k->tracePrimary ? k->tracePrimary : k->intersect);
            k->cubinBytes = e->cubin.size();
            k->compileMs = e->compileMs;
        }
    }
    e->loaded[device] = k; // failures are remembered too: no retry storm
    return k;
}

int specOccupancy(void *fn, int threads, size_t smemBytes)
{
    int n = 0;
    if(smemBytes > 48 * 1024 && g_libs.cuFuncSetAttribute(fn, 8 /* CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES */, (int)smemBytes) != 0)
        return 0;
    if(g_libs.cuOccupancy(&n, fn, threads, smemBytes) != 0)
        return 0;
    return n;
}

int specLaunch(void *fn, int blocks, int threads, size_t smemBytes, cudaStream_t stream, void **args, std::string &err)
{
    int rc = g_libs.cuLaunchKernel(fn, (unsigned)blocks, 1, 1, (unsigned)threads, 1, 1, (unsigned)smemBytes, (void *)stream, args, nullptr);
    if(rc != 0)
    {
        err = "cuLaunchKernel: " + cuError(rc);
        return -1;
    }
    return 0;
}

} // namespace ptb
