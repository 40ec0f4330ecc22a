// scenes_api.cpp -- C entry points of oracle/libsrtscenes.so: the seeded scene generators
// compiled ALONE, by the host compiler, from simpleraytracer_b200/csrc/scenes.cpp -- no CUDA, no other product
// code.  Test infrastructure: it lets the CPU arm of bench.py (`--impl reference`) and the oracle tests obtain
// the benchmark's input arrays without loading libModelRunner.so, so the checker's input does not come from the
// thing being checked.  tests/test_host_logic.py pins that both builds of the generator emit identical arrays.
#include "scene.h"

extern "C" {

size_t orc_gen_cornell_box(float* vertices, float* albedo, srt_camera* camera, srt_light* light,
                           uint32_t width, uint32_t height)
{
    return ML::GenCornellBox(vertices, albedo, camera, light, width, height);
}

int orc_gen_mesh(uint32_t seed, size_t n_triangles, float* vertices, float* albedo, srt_camera* camera,
                 srt_light* light, uint32_t width, uint32_t height)
{
    try
    {
        ML::GenMesh(seed, n_triangles, vertices, albedo, camera, light, width, height);
        return 0;
    }
    catch (...)
    {
        return 1;
    }
}

} // extern "C"
