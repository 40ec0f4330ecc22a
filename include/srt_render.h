/*
 * srt_render.h -- C ABI of the B200-native render loop (BUILD-DEFINED).
 *
 * The reference tree (achernigin1987/SimpleRayTracer) ships no renderer: its
 * only code is model_runner/ (SURVEY.md section 0).  BASELINE.json still puts the
 * per-pixel render loop on the hot path, so this header defines the boundary a
 * host application would bind for it: plain pointers and sizes, no C++ or
 * torch types.  The frame is written into the same dense HWC fp32 layout the
 * reference uses for images (model_runner.h:139, image.cpp:43), so a rendered
 * frame can be consumed by mlInfer() without a host copy (srtRenderToImage).
 *
 * The arithmetic every function below performs is fixed by docs/RENDER_SPEC.md;
 * oracle/ holds the independent CPU restatement used by the tests.
 */
#ifndef SRT_RENDER_H_B200
#define SRT_RENDER_H_B200

#include <stddef.h>
#include <stdint.h>
#include "model_runner.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct srt_scene_t* srt_scene;

/* Pinhole camera.  forward/right/up are unit vectors; half_w = tan(fov_x/2),
 * half_h = tan(fov_y/2), evaluated once on the host (never on the device). */
struct srt_camera
{
    float origin[3];
    float forward[3];
    float right[3];
    float up[3];
    float half_w;
    float half_h;
};

/* One point light; radiant intensity per channel. */
struct srt_light
{
    float position[3];
    float intensity[3];
};

enum srt_shade_mode
{
    SRT_SHADE_ALBEDO = 0,  /* closest hit only: rgb = albedo of the hit triangle    */
    SRT_SHADE_LAMBERT = 1, /* Lambert + point light, no visibility test             */
    SRT_SHADE_SHADOW = 2   /* Lambert + point light + one any-hit shadow ray        */
};

enum srt_accel_mode
{
    SRT_ACCEL_BRUTE = 0, /* every ray tests every triangle (FP32-pipe roofline)   */
    SRT_ACCEL_BVH = 1    /* accelerated: exact candidate grids built on the device (screen tiles for primary
                            rays, a perspective grid around the light for shadow rays; docs/RENDER_SPEC.md
                            section 8) in front of a stack traversal of the scene BVH.  Same image as
                            SRT_ACCEL_BRUTE, bit for bit.                                                  */
};

struct srt_render_params
{
    uint32_t width;      /* full frame size, pixels                               */
    uint32_t height;
    uint32_t y0;         /* rows [y0,y1) are rendered (row-band sharding)         */
    uint32_t y1;
    uint32_t buf_y0;     /* frame row that the output pointers' first row holds   */
    uint32_t row_block_stride; /* 0 or 1: every row of the band.  N > 1: only the  */
    uint32_t row_block_phase;  /* 8-row blocks b = y/8 with b % N == phase are      */
                               /* rendered (interleaved sharding over N GPUs for    */
                               /* load balance); needs y0 % 8 == 0                  */
    uint32_t spp;        /* samples per pixel, accumulated in index order         */
    uint32_t jitter;     /* 0: pixel centre; 1: hashed sub-pixel jitter           */
    uint32_t seed;
    uint32_t shade;      /* srt_shade_mode                                        */
    uint32_t accel;      /* srt_accel_mode                                        */
    uint32_t collect_stats; /* 1: also count box/triangle tests (slower variant)  */
    float t_min;         /* primary rays accept t > t_min                         */
    float shadow_eps;    /* shadow-ray origin offset along the facing normal      */
    float background[3];
};

struct srt_render_stats
{
    uint64_t primary_rays;
    uint64_t shadow_rays;
    uint64_t box_tests;   /* only with collect_stats                             */
    uint64_t tri_tests;   /* only with collect_stats                             */
    float kernel_ms;      /* CUDA-event time of the call's device work: grid build + render kernel(s) */
    float grid_build_ms;  /* of which: per-frame candidate-grid build (srtRenderDevice* only, else 0) */
    float light_grid_build_ms; /* build time of the light grid now in use (built once per light position,
                                  outside later frames: 0 when no light grid exists)                   */
};

struct srt_scene_info
{
    uint64_t n_triangles;
    uint64_t n_bvh_nodes;
    uint64_t device_bytes;
    float bvh_build_ms;   /* host wall time                                      */
    float upload_ms;      /* host->device copy, CUDA events                      */
    float bounds_min[3];
    float bounds_max[3];
};

/* ---- scene ------------------------------------------------------------ */

/* vertices: 9 floats per triangle (v0,v1,v2); albedo: 3 floats per triangle
 * (NULL: 0.8 grey).  build_bvh = 0 skips the BVH (brute force only).
 * Failure: returns NULL and leaves a message on the context
 * (same convention as mlCreateImage/mlCreateModel). */
ML_API_ENTRY srt_scene srtCreateScene(ml_context context, size_t n_triangles,
                                      float const* vertices, float const* albedo,
                                      int build_bvh);
ML_API_ENTRY ml_status srtGetSceneInfo(srt_scene scene, srt_scene_info* info);
ML_API_ENTRY char* srtGetSceneError(srt_scene scene, char* buffer, size_t buffer_size);
ML_API_ENTRY void srtReleaseScene(srt_scene scene);

/* Copies the flattened BVH to the caller for inspection/tests.  nodes: 16
 * floats per node (see docs/RENDER_SPEC.md), order: n_triangles original ids
 * in leaf order.  Either pointer may be NULL.  Returns the node count. */
ML_API_ENTRY size_t srtGetSceneBvh(srt_scene scene, float* nodes, size_t max_nodes,
                                   int32_t* order, size_t max_order);

/* ---- synthetic scenes (host-side generators, seeded, deterministic) ---- */

/* Cornell box: 5 walls + 2 inner blocks... returns the triangle count; call
 * with NULL arrays to query it.  Fills a matching camera and light. */
ML_API_ENTRY size_t srtGenCornellBox(float* vertices, float* albedo,
                                     srt_camera* camera, srt_light* light,
                                     uint32_t width, uint32_t height);
/* Displaced-grid terrain plus a sphere soup, exactly n_triangles triangles. */
ML_API_ENTRY ml_status srtGenMesh(uint32_t seed, size_t n_triangles,
                                  float* vertices, float* albedo,
                                  srt_camera* camera, srt_light* light,
                                  uint32_t width, uint32_t height);

/* ---- render ----------------------------------------------------------- */

/* Host-buffer entry point: renders rows [y0,y1) and copies them into the
 * caller's buffers (any of rgb/hit_id/hit_t may be NULL).  Synchronous.
 * rgb: H*W*3 floats (HWC); hit_id: H*W int32 (-1 = miss); hit_t: H*W floats
 * (+inf = miss).  With spp > 1, hit_id/hit_t are those of sample 0.
 * The band is rendered, copied and scattered in overlapped chunks; buffers that
 * are page-locked as a whole (cudaHostAlloc / cudaHostRegister) are filled
 * straight from the device, pageable ones through pinned staging. */
ML_API_ENTRY ml_status srtRender(srt_scene scene, srt_camera const* camera,
                                 srt_light const* light, srt_render_params const* params,
                                 float* rgb, int32_t* hit_id, float* hit_t,
                                 srt_render_stats* stats);

/* Device-buffer entry point: pointers are device addresses on the scene's
 * GPU (they may be peer-mapped memory of another GPU), stream is a
 * cudaStream_t (NULL: the context's stream).  Asynchronous; ray counters are
 * fetched with srtGetRenderStats(), which synchronises. */
ML_API_ENTRY ml_status srtRenderDevice(srt_scene scene, srt_camera const* camera,
                                       srt_light const* light, srt_render_params const* params,
                                       void* d_rgb, void* d_hit_id, void* d_hit_t,
                                       void* stream);
ML_API_ENTRY ml_status srtGetRenderStats(srt_scene scene, srt_render_stats* stats);

/* Renders straight into an ml_image (FLOAT32, 3 channels, width x height):
 * the frame stays in HBM and becomes the image's authoritative copy, so a
 * following mlInfer(model, image, out) reads it without a host round trip. */
ML_API_ENTRY ml_status srtRenderToImage(srt_scene scene, srt_camera const* camera,
                                        srt_light const* light, srt_render_params const* params,
                                        ml_image image, srt_render_stats* stats);

/* ---- multi-GPU frames ---------------------------------------------------
 * One frame is sharded over N GPUs in interleaved 8-row blocks (params.row_block_stride = N,
 * row_block_phase = rank; scene replicated).  Every GPU's render kernel stores its pixels straight into
 * ONE frame buffer (the memory of GPU 0, peer-mapped: cudaDeviceEnablePeerAccess inside one process,
 * mlxIpcOpen across processes) and its last block publishes a completion flag next to it; GPU 0 waits
 * for the flags on the device.  No gather pass, no host round trip, no collective.
 *
 * Flag array (uint32, zero-initialised, in GPU 0's memory): flags[chunk * n_ranks + rank].  A frame uses a
 * new `value` (an epoch that only grows); waiting is "flag - value >= 0" in wrap-safe arithmetic. */

/* Rows [y0, y1) cut into at most requested_chunks chunks of whole (8 * stride)-row groups: the layout both
 * srtRenderDeviceSignal and srtGatherFrameHost use.  Chunk k covers rows [y0 + k*chunk_rows, ...). */
ML_API_ENTRY ml_status srtChunkLayout(srt_render_params const* params, uint32_t requested_chunks,
                                      uint32_t* chunk_rows, uint32_t* n_chunks);

/* srtRenderDevice + completion flags: the band is rendered as the chunks of srtChunkLayout(params, n_chunks)
 * and chunk k's last block stores `value` to ((uint32_t*)d_flags)[k * chunk_stride] with system-scope
 * release ordering after all of the chunk's pixels (d_flags may be peer memory).  Asynchronous. */
ML_API_ENTRY ml_status srtRenderDeviceSignal(srt_scene scene, srt_camera const* camera,
                                             srt_light const* light, srt_render_params const* params,
                                             void* d_rgb, void* d_hit_id, void* d_hit_t, void* stream,
                                             void* d_flags, uint32_t chunk_stride, uint32_t n_chunks,
                                             uint32_t value);

/* Enqueues on `stream` (NULL: the context's stream) a device-side wait until the `count` flags
 * d_flags[k * stride] have reached `value`; work enqueued behind it sees the signalling kernels' pixels.
 * The wait gives up after timeout_ms (0: 2000) so a dead peer cannot hang the GPU; timeouts are counted
 * (srtFlagWaitTimeouts, cumulative per context, valid after the stream has been synchronised). */
ML_API_ENTRY ml_status srtStreamWaitFlags(ml_context context, void* stream, void const* d_flags,
                                          uint32_t count, uint32_t stride, uint32_t value, uint32_t timeout_ms);
/* Enqueues a release store of `value` to *d_flag (e.g. "frame consumed", the reverse direction). */
ML_API_ENTRY ml_status srtStreamWriteFlag(ml_context context, void* stream, void* d_flag, uint32_t value);
ML_API_ENTRY uint32_t srtFlagWaitTimeouts(ml_context context);

/* GPU 0's half of a host-buffer frame: for every chunk of srtChunkLayout(params, n_chunks) waits (on the
 * device) for the n_ranks flags of that chunk and copies the chunk's rows of d_rgb (full frame, W*H*3
 * floats on the context's GPU) into rgb_host while later chunks are still being rendered.  Page-locked
 * destinations are written straight from the device, pageable ones through pinned staging.  Synchronous;
 * ML_FAIL (message on the context) if a wait timed out. */
ML_API_ENTRY ml_status srtGatherFrameHost(ml_context context, srt_render_params const* params,
                                          void const* d_rgb, void const* d_flags, uint32_t n_ranks,
                                          uint32_t n_chunks, uint32_t value, float* rgb_host,
                                          uint32_t timeout_ms);

/* One process, N GPUs (the host side stays C++; no torch, no NCCL): a scene group holds a context and a
 * scene replica per device (BVH built once), GPU devices[0] owns the frame and the flags, every other
 * device gets peer access to it, and one persistent host thread per device issues that device's work.
 * error/error_size receive the reason when creation fails (NULL return). */
typedef struct srt_scene_group_t* srt_scene_group;
ML_API_ENTRY srt_scene_group srtCreateSceneGroup(int const* devices, uint32_t n_devices, size_t n_triangles,
                                                 float const* vertices, float const* albedo, int build_bvh,
                                                 char* error, size_t error_size);
/* Renders the full frame on all devices of the group.  rgb_host != NULL: H*W*3 floats receive the frame
 * (chunked: the copy of chunk k overlaps the rendering of chunk k+1); NULL: the frame stays in GPU 0's
 * memory (srtSceneGroupFrame).  params.y0/y1/buf_y0/row_block_* are set by the call (whole frame).
 * Synchronous.  stats: rays summed over the devices; kernel_ms = GPU 0's timeline from its first launch to
 * "all flags seen" (the frame time on the device). */
ML_API_ENTRY ml_status srtRenderGroup(srt_scene_group group, srt_camera const* camera, srt_light const* light,
                                      srt_render_params const* params, float* rgb_host, srt_render_stats* stats);
ML_API_ENTRY void* srtSceneGroupFrame(srt_scene_group group);
ML_API_ENTRY uint32_t srtSceneGroupSize(srt_scene_group group);
ML_API_ENTRY char* srtGetSceneGroupError(srt_scene_group group, char* buffer, size_t buffer_size);
ML_API_ENTRY void srtReleaseSceneGroup(srt_scene_group group);

/* ---- render without allocation -------------------------------------------
 * srtRenderDeviceNoAlloc has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  It is srtRenderDevice -- same arguments, same pixels, bit for bit -- with all of
 * the call's scratch memory (ray counters, the screen grid's tile lists and per-triangle records) in ONE static
 * device array of the library (about 4.3 MB that every GPU loading the library holds until exit) instead of
 * buffers the scene allocates on demand.  It calls no cudaMalloc / cudaFree / allocating routine and creates
 * no events; inputs (the resident scene) and outputs (d_rgb / d_hit_id / d_hit_t, device pointers, any may be
 * NULL) are the caller's.  Limits: the scene holds at most SRT_NOALLOC_MAX_TRIANGLES triangles and the frame at
 * most SRT_NOALLOC_MAX_TILES 8x8-pixel tiles (e.g. 1024 x 1024); beyond either the call fails with ML_FAIL and a
 * message on the scene.  Shadow rays walk the BVH (no light grid: its scratch scales with the scene).
 * Asynchronous on `stream` (NULL: the context's stream); no ray statistics (srtGetRenderStats does not see this
 * call).  Every call on a GPU shares the one array, which keeps the previous call's contents (what must start at
 * zero is cleared by the call itself): calls that OVERLAP on the GPU -- other streams, other scenes, other
 * threads -- race and give undefined pixels; serialise them.  The first call on a device resolves the array's
 * address (cudaGetSymbolAddress): make one call before starting a graph capture. */
#define SRT_NOALLOC_MAX_TRIANGLES 32768
#define SRT_NOALLOC_MAX_TILES 16384
ML_API_ENTRY ml_status srtRenderDeviceNoAlloc(srt_scene scene, srt_camera const* camera,
                                              srt_light const* light, srt_render_params const* params,
                                              void* d_rgb, void* d_hit_id, void* d_hit_t, void* stream);

/* Independent-FFMA microbenchmark over all SMs: the FP32 roofline denominator
 * (MEASURED_PEAKS.json carries no FP32 figure).  Returns TFLOP/s (FMA = 2). */
ML_API_ENTRY double srtMeasureFp32Peak(ml_context context, int repeats);

#ifdef __cplusplus
} /* extern "C" */
#endif

#endif /* SRT_RENDER_H_B200 */
