// multi_gpu_main.cpp — a C++-only caller of the multi-GPU entry points (no Python, no torch): builds
// BASELINE config 2 (diffuse sphere on a plane under a constant sky) with the host scene API, renders it on ONE
// device (ptb200_render) and sample-sharded on N devices (ptb200_comm_init_all + ptb200_scene_create_multi +
// ptb200_render_multi: NCCL reduce-scatter inside the library), and compares the two frames.  The N-device frame
// is the same estimator over the same global sample set, so they agree up to float summation order.
//   usage: multi_gpu_main [n_devices=2] [width=192] [height=108] [spp=64]
// exit code 0 = frames agree, 1 = they differ, 2 = fewer than n devices (the test skips), 3 = error.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "ptb200_render.hpp"

using namespace PathTrace;

static Object *unionArray(std::vector<Object *> &objs, size_t start, size_t end)
{ // balanced Union tree, the shape of the reference's unionArray (src/test.cpp:52-64)
    if(end - start == 1)
        return objs[start];
    size_t mid = (start + end) / 2;
    return new Union(unionArray(objs, start, mid), unionArray(objs, mid, end));
}

int main(int argc, char **argv)
{
    const int n = argc > 1 ? atoi(argv[1]) : 2, W = argc > 2 ? atoi(argv[2]) : 192, H = argc > 3 ? atoi(argv[3]) : 108, spp = argc > 4 ? atoi(argv[4]) : 64;
    if(ptb200_device_count() < n)
    {
        printf("multi_gpu_main: %d device(s), need %d\n", ptb200_device_count(), n);
        return 2;
    }
    try
    {
        static Material white(new ColorTexture(Color(0.8f)), new ColorTexture(Color(1)));
        static Material sky(new ColorTexture(Color(0)), new ColorTexture(Color(0)), new ColorTexture(Color(1)));
        std::vector<Object *> objs;
        objs.push_back(new Sphere(Vector3D(0, 0, -4), 1.0f, &white));
        objs.push_back(new Plane(Vector3D(0, 1, 0), Vector3D(0, -1, 0), &white));
        const float ext = 200;
        const Vector3D dirs[6] = {Vector3D(0, 0, -1), Vector3D(0, 0, 1), Vector3D(0, -1, 0), Vector3D(0, 1, 0), Vector3D(1, 0, 0), Vector3D(-1, 0, 0)};
        for(int i = 0; i < 6; i++)
            objs.push_back(new Plane(dirs[i], ext, &sky)); // the six emissive far planes, src/test.cpp:134-140
        Object *world = unionArray(objs, 0, objs.size());
        PtbRenderParams params = ptb200::defaultParams(W, H, spp);
        std::vector<float> one((size_t)W * H * 4, -1.f), many((size_t)W * H * 4, -2.f);
        PtbStats st1;
        {
            ptb200::DeviceScene scene(world);
            ptb200::renderFrame(scene, params, one.data(), &st1);
        }
        std::vector<int> devices;
        for(int i = 0; i < n; i++)
            devices.push_back(i);
        std::vector<PtbStats> stN(n);
        {
            ptb200::MultiDeviceScene scenes(world, devices);
            scenes.render(params, many.data(), stN.data());
        }
        delete world;
        unsigned long long paths = 0, rays = 0;
        for(int i = 0; i < n; i++)
            paths += stN[i].paths, rays += stN[i].rays;
        double worst = 0, mean = 0;
        for(size_t i = 0; i < one.size(); i++)
        {
            double d = fabs((double)one[i] - many[i]) / (fabs((double)one[i]) + 1e-3);
            worst = d > worst ? d : worst;
            mean += one[i];
        }
        mean /= (double)one.size();
        printf("multi_gpu_main: %d devices, %dx%d x %d spp: 1-device paths %llu rays %llu, %d-device paths %llu rays %llu, frame mean %.5f, worst relative difference %.3g\n", n, W, H,
               spp, (unsigned long long)st1.paths, (unsigned long long)st1.rays, n, paths, rays, mean, worst);
        if(paths != st1.paths || rays != st1.rays)
        {
            printf("multi_gpu_main: the shards do not add up to the one-device sample set\n");
            return 1;
        }
        return worst <= 1e-4 ? 0 : 1;
    }
    catch(std::exception &e)
    {
        printf("multi_gpu_main: %s\n", e.what());
        return 3;
    }
}
