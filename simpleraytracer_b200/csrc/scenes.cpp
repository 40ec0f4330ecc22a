// scenes.cpp -- seeded synthetic scene generators (SURVEY.md row R6;
// BUILD-DEFINED: the reference ships no scene or mesh I/O).  They only produce
// INPUT arrays (9 floats per triangle + 3 albedo floats); the same arrays are
// handed to the CUDA path and to the CPU oracle by the tests and the bench.
#include "scene.h"

#include <cmath>

namespace ML {
namespace {

struct Rng
{
    uint32_t state;
    explicit Rng(uint32_t seed) : state(seed * 0x9E3779B9u + 0x7F4A7C15u) {}
    uint32_t Next()
    {
        state += 0x9E3779B9u;
        uint32_t h = state;
        h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
        return h;
    }
    float Uniform() { return static_cast<float>(Next() >> 8) * (1.0f / 16777216.0f); }
    float Range(float lo, float hi) { return lo + (hi - lo) * Uniform(); }
};

struct TriSink
{
    float* vertices;
    float* albedo;
    size_t count = 0;
    size_t capacity;
    bool Full() const { return count >= capacity; }
    void Push(float const* a, float const* b, float const* c, float const* rgb)
    {
        if (Full()) return;
        if (vertices)
        {
            float* v = vertices + 9 * count;
            for (int k = 0; k < 3; k++) { v[k] = a[k]; v[3 + k] = b[k]; v[6 + k] = c[k]; }
        }
        if (albedo)
        {
            float* m = albedo + 3 * count;
            m[0] = rgb[0]; m[1] = rgb[1]; m[2] = rgb[2];
        }
        count++;
    }
    void Quad(float const* a, float const* b, float const* c, float const* d, float const* rgb)
    {
        Push(a, b, c, rgb);
        Push(a, c, d, rgb);
    }
};

void LookAt(srt_camera* cam, double const eye[3], double const target[3], double fov_y_deg,
            uint32_t width, uint32_t height)
{
    if (!cam) return;
    double f[3] = {target[0] - eye[0], target[1] - eye[1], target[2] - eye[2]};
    double fl = std::sqrt(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
    for (double& v : f) v /= fl;
    double up[3] = {0, 1, 0};
    double r[3] = {f[1] * up[2] - f[2] * up[1], f[2] * up[0] - f[0] * up[2], f[0] * up[1] - f[1] * up[0]};
    double rl = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
    for (double& v : r) v /= rl;
    double u[3] = {r[1] * f[2] - r[2] * f[1], r[2] * f[0] - r[0] * f[2], r[0] * f[1] - r[1] * f[0]};
    for (int k = 0; k < 3; k++)
    {
        cam->origin[k] = static_cast<float>(eye[k]);
        cam->forward[k] = static_cast<float>(f[k]);
        cam->right[k] = static_cast<float>(r[k]);
        cam->up[k] = static_cast<float>(u[k]);
    }
    double half_h = std::tan(fov_y_deg * 0.5 * 3.14159265358979323846 / 180.0);
    double aspect = height ? static_cast<double>(width) / static_cast<double>(height) : 1.0;
    cam->half_h = static_cast<float>(half_h);
    cam->half_w = static_cast<float>(half_h * aspect);
}

float TerrainHeight(float x, float z)
{
    return 4.0f * std::sin(0.13f * x) * std::cos(0.11f * z)
         + 1.5f * std::sin(0.37f * x + 0.5f * z)
         + 0.6f * std::sin(0.9f * z - 0.3f * x);
}

} // namespace

size_t GenCornellBox(float* vertices, float* albedo, srt_camera* camera, srt_light* light,
                     uint32_t width, uint32_t height)
{
    TriSink sink{vertices, albedo, 0, 12};
    float const white[3] = {0.73f, 0.73f, 0.73f};
    float const red[3] = {0.65f, 0.05f, 0.05f};
    float const green[3] = {0.12f, 0.45f, 0.15f};
    float const lamp[3] = {1.0f, 1.0f, 0.9f};

    float const p000[3] = {-1, -1, -1}, p100[3] = {1, -1, -1}, p010[3] = {-1, 1, -1}, p110[3] = {1, 1, -1};
    float const p001[3] = {-1, -1, 1}, p101[3] = {1, -1, 1}, p011[3] = {-1, 1, 1}, p111[3] = {1, 1, 1};

    sink.Quad(p000, p100, p101, p001, white); // floor   y = -1
    sink.Quad(p010, p011, p111, p110, white); // ceiling y = +1
    sink.Quad(p000, p010, p110, p100, white); // back    z = -1
    sink.Quad(p000, p001, p011, p010, red);   // left    x = -1
    sink.Quad(p100, p110, p111, p101, green); // right   x = +1
    float const l0[3] = {-0.25f, 0.995f, -0.25f}, l1[3] = {0.25f, 0.995f, -0.25f};
    float const l2[3] = {0.25f, 0.995f, 0.25f}, l3[3] = {-0.25f, 0.995f, 0.25f};
    sink.Quad(l0, l1, l2, l3, lamp);          // lamp quad just below the ceiling

    double const eye[3] = {0.0, 0.0, 3.4};
    double const target[3] = {0.0, 0.0, 0.0};
    LookAt(camera, eye, target, 45.0, width, height);
    if (light)
    {
        light->position[0] = 0.0f; light->position[1] = 0.9f; light->position[2] = 0.0f;
        light->intensity[0] = 1.6f; light->intensity[1] = 1.6f; light->intensity[2] = 1.4f;
    }
    return sink.count;
}

void GenMesh(uint32_t seed, size_t n_triangles, float* vertices, float* albedo,
             srt_camera* camera, srt_light* light, uint32_t width, uint32_t height)
{
    TriSink sink{vertices, albedo, 0, n_triangles};
    Rng rng(seed);

    // --- 1. displaced-grid terrain: 2*G*G triangles, ~60 % of the budget ----
    size_t grid = static_cast<size_t>(std::sqrt(0.3 * static_cast<double>(n_triangles)));
    if (n_triangles < 64) grid = 0;
    float const half = 50.0f;
    if (grid > 0)
    {
        float const step = 2.0f * half / static_cast<float>(grid);
        std::vector<float> h((grid + 1) * (grid + 1));
        for (size_t j = 0; j <= grid; j++)
            for (size_t i = 0; i <= grid; i++)
            {
                float x = -half + step * static_cast<float>(i);
                float z = -half + step * static_cast<float>(j);
                h[j * (grid + 1) + i] = TerrainHeight(x, z) + 0.15f * (rng.Uniform() - 0.5f) * step;
            }
        for (size_t j = 0; j < grid; j++)
            for (size_t i = 0; i < grid; i++)
            {
                float x0 = -half + step * static_cast<float>(i), x1 = x0 + step;
                float z0 = -half + step * static_cast<float>(j), z1 = z0 + step;
                float a[3] = {x0, h[j * (grid + 1) + i], z0};
                float b[3] = {x1, h[j * (grid + 1) + i + 1], z0};
                float c[3] = {x1, h[(j + 1) * (grid + 1) + i + 1], z1};
                float d[3] = {x0, h[(j + 1) * (grid + 1) + i], z1};
                float t = 0.5f + 0.08f * a[1];
                t = t < 0.f ? 0.f : (t > 1.f ? 1.f : t);
                float tint = 0.9f + 0.1f * rng.Uniform();
                float rgb[3] = {(0.25f + 0.45f * t) * tint, (0.45f + 0.25f * t) * tint, (0.20f + 0.30f * t) * tint};
                sink.Push(a, c, b, rgb);
                sink.Push(a, d, c, rgb);
            }
    }

    // --- 2. sphere soup: 24 lat-long spheres, ~35 % of the budget ------------
    size_t const n_spheres = n_triangles >= 2000 ? 24 : 0;
    if (n_spheres)
    {
        double per = 0.35 * static_cast<double>(n_triangles) / static_cast<double>(n_spheres);
        int stacks = static_cast<int>((1.0 + std::sqrt(1.0 + per)) * 0.5);
        if (stacks < 3) stacks = 3;
        int slices = 2 * stacks;
        for (size_t s = 0; s < n_spheres; s++)
        {
            float cx = rng.Range(-40.f, 40.f), cz = rng.Range(-40.f, 40.f);
            float r = rng.Range(1.5f, 4.0f);
            float cy = TerrainHeight(cx, cz) + r + rng.Range(0.2f, 6.0f);
            float rgb[3] = {rng.Range(0.3f, 0.9f), rng.Range(0.3f, 0.9f), rng.Range(0.3f, 0.9f)};
            auto point = [&](int st, int sl, float* out) {
                double theta = 3.14159265358979323846 * st / stacks;
                double phi = 2.0 * 3.14159265358979323846 * (sl % slices) / slices;
                out[0] = cx + r * static_cast<float>(std::sin(theta) * std::cos(phi));
                out[1] = cy + r * static_cast<float>(std::cos(theta));
                out[2] = cz + r * static_cast<float>(std::sin(theta) * std::sin(phi));
            };
            for (int st = 0; st < stacks; st++)
                for (int sl = 0; sl < slices; sl++)
                {
                    float a[3], b[3], c[3], d[3];
                    point(st, sl, a); point(st, sl + 1, b);
                    point(st + 1, sl + 1, c); point(st + 1, sl, d);
                    if (st != 0) sink.Push(a, b, c, rgb);          // skip the degenerate pole half
                    if (st != stacks - 1) sink.Push(a, c, d, rgb);
                }
        }
    }

    // --- 3. confetti: small random triangles until the count is exact ---------
    while (!sink.Full())
    {
        float x = rng.Range(-45.f, 45.f), z = rng.Range(-45.f, 45.f);
        float y = TerrainHeight(x, z) + rng.Range(0.5f, 12.0f);
        float a[3] = {x, y, z};
        float b[3] = {x + rng.Range(-0.5f, 0.5f), y + rng.Range(-0.5f, 0.5f), z + rng.Range(-0.5f, 0.5f)};
        float c[3] = {x + rng.Range(-0.5f, 0.5f), y + rng.Range(-0.5f, 0.5f), z + rng.Range(-0.5f, 0.5f)};
        float rgb[3] = {rng.Range(0.2f, 1.0f), rng.Range(0.2f, 1.0f), rng.Range(0.2f, 1.0f)};
        sink.Push(a, b, c, rgb);
    }

    double const eye[3] = {0.0, 30.0, 66.0};
    double const target[3] = {0.0, -2.0, 0.0};
    LookAt(camera, eye, target, 42.0, width, height);
    if (light)
    {
        light->position[0] = 30.0f; light->position[1] = 60.0f; light->position[2] = 25.0f;
        light->intensity[0] = 5200.0f; light->intensity[1] = 5000.0f; light->intensity[2] = 4600.0f;
    }
}

} // namespace ML
