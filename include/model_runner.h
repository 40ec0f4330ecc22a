/*
 * model_runner.h -- C ABI of the B200-native model runner (drop-in boundary).
 *
 * ABI-compatible with the reference header model_runner/model_runner.h
 * (handles :65,:70,:75; ml_status :82-86; ml_data_type :91-95; ml_image_info
 * :100-106; ml_model_params :53-60; the 13 entry points :115-:258).  A client
 * compiled against the reference header (e.g. the reference's test_app.cpp)
 * links against libModelRunner.so built from this tree without changes.
 *
 * What differs is behind the boundary: images live in pinned host memory with
 * a device-resident twin in HBM, the model is executed by hand-written sm_100a
 * kernels, and every call is synchronous as seen by the caller.
 *
 * Call protocol (same as the reference, model_runner.h:6-24):
 *   ctx = mlCreateContext(); model = mlCreateModel(ctx,&params);
 *   mlGetModelInfo -> fill width/height -> mlSetModelInputInfo -> mlGetModelInfo
 *   mlCreateImage x2; mlMapImage/write/mlUnmapImage; mlInfer; mlMapImage/read.
 *   Release images, model and context (any order; the context owns nothing).
 */
#ifndef MODEL_RUNNER_H_B200
#define MODEL_RUNNER_H_B200

#include <stddef.h>

#if defined(_WIN32)
#  ifdef RADEONPROML_BUILD
#    define ML_API_ENTRY __declspec(dllexport)
#  else
#    define ML_API_ENTRY __declspec(dllimport)
#  endif
#elif defined(__GNUC__)
#  ifdef RADEONPROML_BUILD
#    define ML_API_ENTRY __attribute__((visibility("default")))
#  else
#    define ML_API_ENTRY
#  endif
#else
#  define ML_API_ENTRY
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* Only these three leading fields are ever read by the library, so callers
 * built against the longer struct in ml.h remain compatible. Zero unused
 * fields. (reference: model_runner.h:53-60, model.cpp:80,101,107) */
struct ml_model_params
{
    char const* model_path;  /* frozen GraphDef, binary protobuf            */
    char const* input_node;  /* NULL: first node of the graph               */
    char const* output_node; /* NULL: last node of the graph                */
};

typedef struct ml_context_t* ml_context;
typedef struct ml_model_t*   ml_model;
typedef struct ml_image_t*   ml_image;

#define ML_INVALID_HANDLE NULL

enum ml_status
{
    ML_OK,
    ML_FAIL
};

enum ml_data_type
{
    ML_FLOAT32,
    ML_FLOAT16,
};

/* Dense row-major (height, width, channels) image; 0 = dimension unknown. */
struct ml_image_info
{
    ml_data_type dtype;
    size_t width;
    size_t height;
    size_t channels;
};

/* ---- context ---------------------------------------------------------- */
ML_API_ENTRY ml_context mlCreateContext();
ML_API_ENTRY char* mlGetContextError(ml_context context, char* buffer, size_t buffer_size);
ML_API_ENTRY void mlReleaseContext(ml_context context);

/* ---- image ------------------------------------------------------------ */
ML_API_ENTRY ml_image mlCreateImage(ml_context context, ml_image_info const* info);
ML_API_ENTRY ml_status mlGetImageInfo(ml_image image, ml_image_info* info);
ML_API_ENTRY void* mlMapImage(ml_image image, size_t* size);
ML_API_ENTRY ml_status mlUnmapImage(ml_image image, void* data);
ML_API_ENTRY void mlReleaseImage(ml_image image);

/* ---- model ------------------------------------------------------------ */
ML_API_ENTRY ml_model mlCreateModel(ml_context context, ml_model_params const* params);
ML_API_ENTRY char* mlGetModelError(ml_model model, char* buffer, size_t buffer_size);
ML_API_ENTRY ml_status mlGetModelInfo(ml_model model,
                                      ml_image_info* input_info,
                                      ml_image_info* output_info);
ML_API_ENTRY ml_status mlSetModelInputInfo(ml_model model, ml_image_info const* info);
ML_API_ENTRY ml_status mlInfer(ml_model model, ml_image input, ml_image output);
ML_API_ENTRY void mlReleaseModel(ml_model model);

#ifdef __cplusplus
} /* extern "C" */
#endif

#endif /* MODEL_RUNNER_H_B200 */
