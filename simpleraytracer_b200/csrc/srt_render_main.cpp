// srt_render -- command-line client of the render loop, the producer side of the
// raw-image wire format that the reference's test_app consumes
// (test_app.cpp:137-192, model_runner/README.md:173): headerless dense HWC,
// little-endian fp32, width*height*3 values, on stdout or into a file.
//
//   srt_render -s cornell|mesh [-n triangles] [-seed N] -w W -h H [-spp N]
//              [-shade albedo|lambert|shadow] [-accel bvh|brute] [-o file] [-gpus N | -devices a,b,...]
//   srt_render -s mesh -n 100000 -w 800 -h 600 -spp 16 | test_app -m denoiser.pb -w 800 -h 600 > out.bin
//
// -gpus N (or an explicit -devices list) renders the one frame on N GPUs of this process through
// srtCreateSceneGroup / srtRenderGroup (interleaved row blocks, peer stores into GPU 0's frame, device-side
// completion flags); the output is bit-identical to the single-GPU frame.
//
// Uses only the C ABI (model_runner.h + srt_render.h), like test_app does; the
// frame is rendered into an ml_image and read through mlMapImage.  Exit status
// and option style follow test_app (errors on stderr, status 255).
#include "model_runner.h"
#include "srt_render.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace {

[[noreturn]] void Usage()
{
    std::fprintf(stderr,
                 "usage: srt_render -s cornell|mesh [-n triangles] [-seed N] -w width -h height [-spp N]\n"
                 "                  [-shade albedo|lambert|shadow] [-accel bvh|brute] [-o output_file]\n"
                 "                  [-gpus N | -devices a,b,...]\n"
                 "writes width*height*3 fp32 values (HWC, no header) to output_file or stdout\n");
    std::exit(255);
}

std::string ContextError(ml_context ctx)
{
    char buf[1024];
    return mlGetContextError(ctx, buf, sizeof(buf));
}

} // namespace

int main(int argc, char** argv)
{
    std::string scene_name = "cornell", out_path, shade = "shadow", accel = "bvh";
    size_t n_tris = 100000;
    uint32_t seed = 1, width = 0, height = 0, spp = 1;
    std::vector<int> devices;
    for (int i = 1; i < argc; i++)
    {
        std::string a = argv[i];
        auto next = [&]() -> char const* {
            if (i + 1 >= argc) Usage();
            return argv[++i];
        };
        if (a == "-s") scene_name = next();
        else if (a == "-n") n_tris = std::strtoull(next(), nullptr, 10);
        else if (a == "-seed") seed = static_cast<uint32_t>(std::strtoul(next(), nullptr, 10));
        else if (a == "-w") width = static_cast<uint32_t>(std::strtoul(next(), nullptr, 10));
        else if (a == "-h") height = static_cast<uint32_t>(std::strtoul(next(), nullptr, 10));
        else if (a == "-spp") spp = static_cast<uint32_t>(std::strtoul(next(), nullptr, 10));
        else if (a == "-shade") shade = next();
        else if (a == "-accel") accel = next();
        else if (a == "-o") out_path = next();
        else if (a == "-gpus")
        {
            int const n = std::atoi(next());
            if (n < 1) Usage();
            devices.clear();
            for (int d = 0; d < n; d++) devices.push_back(d);
        }
        else if (a == "-devices")
        {
            devices.clear();
            for (char const* c = next(); *c;)
            {
                devices.push_back(std::atoi(c));
                while (*c && *c != ',') c++;
                if (*c == ',') c++;
            }
            if (devices.empty()) Usage();
        }
        else Usage();
    }
    if (width == 0 || height == 0 || spp == 0) Usage();

    ml_context ctx = nullptr;
    srt_scene scene = nullptr;
    srt_scene_group group = nullptr;
    ml_image image = nullptr;
    int status = 0;
    try
    {
        bool const multi = devices.size() > 1;
        if (!multi)
        {
            ctx = devices.empty() ? mlCreateContext() : nullptr;
            if (!devices.empty())
            {
                // one explicit device: the reference API has no device argument, MLB200_DEVICE selects it
                setenv("MLB200_DEVICE", std::to_string(devices[0]).c_str(), 1);
                ctx = mlCreateContext();
            }
            if (ctx == nullptr) throw std::runtime_error("mlCreateContext failed (no CUDA device?)");
        }

        std::vector<float> vertices, albedo;
        srt_camera cam{};
        srt_light light{};
        if (scene_name == "cornell")
        {
            n_tris = srtGenCornellBox(nullptr, nullptr, nullptr, nullptr, width, height);
            vertices.resize(n_tris * 9);
            albedo.resize(n_tris * 3);
            srtGenCornellBox(vertices.data(), albedo.data(), &cam, &light, width, height);
        }
        else if (scene_name == "mesh")
        {
            vertices.resize(n_tris * 9);
            albedo.resize(n_tris * 3);
            if (srtGenMesh(seed, n_tris, vertices.data(), albedo.data(), &cam, &light, width, height) != ML_OK)
                throw std::runtime_error("srtGenMesh failed");
        }
        else Usage();

        srt_render_params p{};
        p.width = width; p.height = height; p.y0 = 0; p.y1 = height; p.buf_y0 = 0;
        p.spp = spp; p.jitter = spp > 1 ? 1u : 0u; p.seed = seed;
        p.shade = shade == "albedo" ? SRT_SHADE_ALBEDO : shade == "lambert" ? SRT_SHADE_LAMBERT : SRT_SHADE_SHADOW;
        p.accel = accel == "bvh" ? SRT_ACCEL_BVH : SRT_ACCEL_BRUTE;
        p.t_min = 1e-4f; p.shadow_eps = 1e-3f;
        p.background[0] = 0.5f; p.background[1] = 0.7f; p.background[2] = 0.9f;
        srt_render_stats st{};
        std::vector<float> host_frame;
        size_t bytes = 0;
        void* data = nullptr;
        if (multi)
        {
            char err[1024];
            group = srtCreateSceneGroup(devices.data(), static_cast<uint32_t>(devices.size()), n_tris, vertices.data(),
                                        albedo.data(), accel == "bvh" ? 1 : 0, err, sizeof(err));
            if (group == nullptr) throw std::runtime_error(std::string("srtCreateSceneGroup: ") + err);
            host_frame.resize(static_cast<size_t>(width) * height * 3);
            if (srtRenderGroup(group, &cam, &light, &p, host_frame.data(), &st) != ML_OK)
                throw std::runtime_error(std::string("srtRenderGroup: ") + srtGetSceneGroupError(group, err, sizeof(err)));
            data = host_frame.data();
            bytes = host_frame.size() * sizeof(float);
        }
        else
        {
            scene = srtCreateScene(ctx, n_tris, vertices.data(), albedo.data(), accel == "bvh" ? 1 : 0);
            if (scene == nullptr) throw std::runtime_error("srtCreateScene: " + ContextError(ctx));
            ml_image_info info{ML_FLOAT32, width, height, 3};
            image = mlCreateImage(ctx, &info);
            if (image == nullptr) throw std::runtime_error("mlCreateImage: " + ContextError(ctx));
            if (srtRenderToImage(scene, &cam, &light, &p, image, &st) != ML_OK)
            {
                char buf[1024];
                throw std::runtime_error(std::string("srtRenderToImage: ") + srtGetSceneError(scene, buf, sizeof(buf)));
            }
            data = mlMapImage(image, &bytes);
            if (data == nullptr) throw std::runtime_error("mlMapImage failed");
        }
        FILE* f = out_path.empty() ? stdout : std::fopen(out_path.c_str(), "wb");
        if (f == nullptr) throw std::runtime_error("cannot open " + out_path);
        size_t const written = std::fwrite(data, 1, bytes, f);
        if (f != stdout) std::fclose(f); else std::fflush(stdout);
        if (!multi) mlUnmapImage(image, data);
        if (written != bytes) throw std::runtime_error("short write");
        std::fprintf(stderr, "srt_render: %zu GPU(s), %ux%u, %zu triangles, %llu primary + %llu shadow rays, kernel %.3f ms (%.1f Mrays/s)\n",
                     multi ? devices.size() : size_t(1), width, height, n_tris, static_cast<unsigned long long>(st.primary_rays),
                     static_cast<unsigned long long>(st.shadow_rays), st.kernel_ms,
                     st.kernel_ms > 0 ? (st.primary_rays + st.shadow_rays) / (st.kernel_ms * 1e3) : 0.0);
    }
    catch (std::exception const& e)
    {
        std::fprintf(stderr, "%s\n", e.what());
        status = 255;
    }
    mlReleaseImage(image);
    if (scene != nullptr) srtReleaseScene(scene);
    if (group != nullptr) srtReleaseSceneGroup(group);
    mlReleaseContext(ctx);
    return status;
}
