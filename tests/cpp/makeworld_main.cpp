// makeworld_main.cpp — compiles the reference's OWN scene code, src/test.cpp:52-145 (unionArray, makeLens,
// makeSkySphericalCoordinates, makeWorld ...), unchanged against this repository's host scene API
// (path-trace_b200/host/ptb200_scene.hpp) — the claim of INTEGRATION.md "Option A".  The line range is extracted
// by the test at run time into a temporary directory (never into the repository) and handed in as
// -DMAKEWORLD_INC="<path>".  The program calls makeWorld() in the current directory (which must hold test2.hdr
// and Serpentine_Valley_3k.hdr, src/test.cpp:121,123), flattens the Object graph and prints the flattened scene
// as text for tests/test_makeworld.py to compare with ptb200/scenes.py::config1.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <cassert>
#include <iostream>
#include <string>

#include "ptb200_render.hpp"

using namespace std;
using namespace PathTrace;

#include MAKEWORLD_INC

static void words(const char *name, const void *p, size_t bytes)
{
    printf("%s", name);
    const unsigned char *b = (const unsigned char *)p;
    for(size_t i = 0; i + 4 <= bytes; i += 4)
    {
        unsigned v;
        memcpy(&v, b + i, 4);
        printf(" %08x", v);
    }
    printf("\n");
}

int main()
{
    try
    {
        Object *world = makeWorld();
        ptb200::Flattener f;
        world->flatten(f);
        PtbSceneDesc d = f.desc();
        printf("counts %d %d %d %d %d %d %d %d %d\n", d.n_prims, d.n_xforms, d.n_chains, d.n_chain_xf, d.n_ops, d.n_materials, d.n_texops, d.n_texdata, d.n_images);
        words("prims", d.prims, sizeof(PtbPrim) * d.n_prims);
        words("xforms", d.xforms, sizeof(PtbXform) * d.n_xforms);
        words("chains", d.chains, sizeof(PtbChain) * d.n_chains);
        words("chain_xf", d.chain_xf, 4 * (size_t)d.n_chain_xf);
        words("ops", d.ops, 4 * (size_t)d.n_ops);
        words("materials", d.materials, sizeof(PtbMaterial) * d.n_materials);
        words("texops", d.texops, sizeof(PtbTexOp) * d.n_texops);
        words("texdata", d.texdata, 4 * (size_t)d.n_texdata);
        for(int i = 0; i < d.n_images; i++)
        {
            double sum = 0;
            for(size_t k = 0; k < (size_t)4 * d.images[i].width * d.images[i].height; k++)
                sum += d.images[i].rgba[k];
            printf("image %d %d %.17g\n", d.images[i].width, d.images[i].height, sum);
        }
        delete world;
        return 0;
    }
    catch(std::exception &e)
    {
        printf("error %s\n", e.what());
        return 1;
    }
}
