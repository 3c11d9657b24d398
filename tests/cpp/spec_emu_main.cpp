// spec_emu_main.cpp — TEST: the register-resident evaluator GENERATED for one scene (ptb_specialize.cu, second
// generation) against the interpreter (evalFast, ptb_csg_fast.cuh), both compiled as HOST code through
// csrc/ptb_host_emu.h and run one ray at a time.  They must agree bit for bit: root span list (length, every t,
// every tag) and the first boundary with t >= eps.
//   -DSPEC_LIST_INC="<file>"   the generated specList  (ray-batch module, functions only)
//   -DSPEC_TRACE_INC="<file>"  the generated specTrace (renderer module, functions only)
//   argv: program.hex rays.bin(float32 n x 6: o, d) cap origin0(0|1)
// prints: rays, overflowed rays (generated code gave up: the kernel reruns those through the interpreter), mismatches
#define PTB_HOST_EMU 1
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "ptb200.h"
#include "ptb_csg_reg.cuh"

namespace ptb
{
inline void specStageTableList(uint32_t tab); // (the list module's copy of specStageTable, renamed by the test)
}
#include SPEC_LIST_INC
#include SPEC_TRACE_INC

// where the two modules' constant tables are staged in the emulated shared memory (pools and stacks lie below)
static const uint32_t kTabTrace = 96u << 10, kTabList = 160u << 10;

using namespace ptb;

template <bool O0>
static int run(const std::vector<uint4> &prog, const std::vector<float> &rays, int cap, int maxStack)
{
    const size_t n = rays.size() / 6;
    size_t overflowed = 0, bad = 0, hits = 0, multi = 0;
    std::vector<float> ovfArea(3 * 1024);
    for(size_t i = 0; i < n; i++)
    {
        const V3 o = mk(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]), d = mk(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
        // interpreter: pool at shared offset 0, start stack behind it (layout of setupFast<GP = true, 1>)
        FastCtx cx;
        cx.prog = 0;
        cx.poolStride = (uint32_t)cap * (12u * PTB_FAST_THREADS);
        cx.stkStride = (uint32_t)(maxStack + 1) * (4u * PTB_FAST_THREADS);
        cx.pool.base = 0, cx.pool.cap = cap, cx.pool.ovf = ovfArea.data(), cx.pool.ovfStride = 1;
        cx.ovfRay = 0;
        cx.stk = cx.pool.base + cx.poolStride;
        V3 oo[1] = {o}, dd[1] = {d};
        int tops[1];
        evalFast<true, O0, 1>(cx, prog.data(), oo, dd, tops);
        std::vector<SpanRec> want;
        for(int k = 0; k < tops[0]; k++)
            want.push_back(cx.pool.load(k));
        float wt = -1.f;
        uint32_t wtag = 0u;
        const int wkind = firstHit(cx.pool, tops[0], wt, wtag);
        hits += wkind != 0, multi += tops[0] >= 2;
        // generated code, list flavour
        REval ev;
        ev.poolBase = 0, ev.cap = cap, ev.top = 0, ev.ovf = false, ev.tab = kTabList;
        const int got = specList<O0>(ev, o, d);
        bool ovf = ev.ovf, ok = true;
        if(!ovf)
        {
            ok = got == tops[0];
            for(int k = 0; ok && k < got; k++)
            {
                const SpanRec s = rpoolLoad(0, k);
                ok = __float_as_uint(s.t0) == __float_as_uint(want[k].t0) && __float_as_uint(s.t1) == __float_as_uint(want[k].t1) && s.tags == want[k].tags;
            }
            if(got >= 2 && ev.top != got)
                ok = false; // the root must be the only list left on the pool stack
        }
        // generated code, first-hit flavour
        ev.poolBase = 0, ev.cap = cap, ev.top = 0, ev.ovf = false, ev.tab = kTabTrace;
        float t = -1.f;
        uint32_t tag = 0u;
        const int kind = specTrace<O0>(ev, o, d, t, tag);
        if(!ev.ovf)
            ok = ok && kind == wkind && __float_as_uint(t) == __float_as_uint(wt) && tag == wtag;
        ovf = ovf || ev.ovf;
        overflowed += ovf;
        if(!ok)
        {
            if(bad < 5)
                printf("mismatch ray %zu: interpreter n=%d kind=%d t=%g tag=%u; generated n=%d kind=%d t=%g tag=%u\n", i, tops[0], wkind, wt, wtag, got, kind, t, tag);
            bad++;
        }
    }
    printf("rays %zu hits %zu multi_span %zu overflowed %zu mismatches %zu\n", n, hits, multi, overflowed, bad);
    return bad ? 1 : 0;
}

int main(int argc, char **argv)
{
    if(argc < 6)
        return 2;
    std::vector<uint4> prog;
    FILE *f = fopen(argv[1], "r");
    uint4 w;
    while(f && fscanf(f, "%x %x %x %x", &w.x, &w.y, &w.z, &w.w) == 4)
        prog.push_back(w);
    if(f)
        fclose(f);
    std::vector<float> rays;
    f = fopen(argv[2], "rb");
    float buf[6];
    while(f && fread(buf, 4, 6, f) == 6)
        rays.insert(rays.end(), buf, buf + 6);
    if(f)
        fclose(f);
    for(threadIdx.x = 0; threadIdx.x < PTB_FAST_THREADS; threadIdx.x++) // the CTA-wide staging loops of the two modules
        ptb::specStageTable(kTabTrace), ptb::specStageTableList(kTabList);
    threadIdx.x = 0;
    const int cap = atoi(argv[3]), maxStack = atoi(argv[5]);
    return atoi(argv[4]) ? run<true>(prog, rays, cap, maxStack) : run<false>(prog, rays, cap, maxStack);
}
