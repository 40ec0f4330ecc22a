/*
 * model_runner_ext.h -- BUILD-DEFINED extensions next to the reference's C API
 * (none of these exist in the reference; the 14 ml* entry points of
 * model_runner.h are unchanged and sufficient for reference-style callers).
 */
#ifndef MODEL_RUNNER_EXT_H_B200
#define MODEL_RUNNER_EXT_H_B200

#include <stddef.h>
#include "model_runner.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Parses a frozen GraphDef on the HOST only (no CUDA call) and writes a text
 * description of what mlCreateModel would build: detected input/output nodes
 * (reference model.cpp:94-120), dtype and dimensions (model.cpp:23-48) and the
 * fused execution plan.  On failure the text is the message mlCreateModel
 * would leave on the context.  Returns ML_OK / ML_FAIL; `buffer` always
 * receives a NUL-terminated string (truncated to buffer_size-1). */
ML_API_ENTRY ml_status mlxInspectModel(char const* model_path, char const* input_node,
                                       char const* output_node, char* buffer, size_t buffer_size);

/* Device pointer of an image's HBM twin, made current first (uploads the host
 * copy if that is the valid one).  for_write != 0 declares that the caller will
 * overwrite the whole image on `stream` and makes the device copy authoritative.
 * Lets an external producer/consumer (e.g. a torch tensor) share the buffer. */
ML_API_ENTRY void* mlxImageDevicePointer(ml_image image, int for_write, void* stream);

/* mlCreateContext() on an explicit CUDA device ordinal (the reference API has
 * no device argument; mlCreateContext uses MLB200_DEVICE or the CUDA runtime's
 * current device).  One process per GPU passes its LOCAL_RANK here: the
 * library links its own CUDA runtime, so another runtime's cudaSetDevice (for
 * example torch's) is not visible to it. */
ML_API_ENTRY ml_context mlxCreateContextOnDevice(int device);

/* Peer mapping of an image's HBM buffer into another PROCESS of the same node
 * (one process per GPU): the owner exports a 64-byte handle, every other rank
 * opens it and receives a device pointer that its kernels can store to over
 * NVLink/NVSwitch.  This is how the multi-GPU render writes each rank's rows
 * straight into rank 0's frame (no gather pass).  mlxIpcExportImage also makes
 * the device copy authoritative (peers are about to overwrite it).
 * mlxIpcOpen returns NULL on failure (message on the context). */
#define MLX_IPC_HANDLE_BYTES 64
ML_API_ENTRY ml_status mlxIpcExportImage(ml_image image, void* handle_out);
ML_API_ENTRY void* mlxIpcOpen(ml_context context, void const* handle);
ML_API_ENTRY void mlxIpcClose(ml_context context, void* device_pointer);

/* Raw device memory on the context's GPU (zero-filled) that can be shared with other processes: the
 * completion flags of a multi-GPU frame (srt_render.h) live in such a buffer on GPU 0.  mlxIpcExportPointer
 * writes the 64-byte handle of a buffer obtained from mlxDeviceAlloc; peers map it with mlxIpcOpen. */
ML_API_ENTRY void* mlxDeviceAlloc(ml_context context, size_t bytes);
ML_API_ENTRY void mlxDeviceFree(ml_context context, void* device_pointer);
ML_API_ENTRY ml_status mlxIpcExportPointer(ml_context context, void* device_pointer, void* handle_out);

/* Arithmetic of the model's 3x3 / stride 1 / SAME convolution layers with 16, 32 or 64k input
 * channels (a thinner first layer, e.g. RGB, is zero-padded to 16) whose filter fits in shared
 * memory; every other layer always runs in fp32 on the CUDA cores.
 *   MLX_PRECISION_BF16X3 (default; MLB200_PRECISION=bf16x3): tensor cores (tcgen05 + TMEM) with
 *     fp32-grade results -- every fp32 operand x is split into two bf16 halves (x = hi + lo to 16
 *     mantissa bits) and each product is issued as hi*hi + hi*lo + lo*hi into the fp32 accumulator
 *     (the dropped lo*lo term is 2^-16 relative).  Stated bound: 1e-4 of the output range against
 *     the fp32 oracle (measured ~1e-5), inside the 1e-3 bar of fp32 inference.  Non-finite inputs
 *     give NaN where fp32 arithmetic would give +-inf.
 *   MLX_PRECISION_FP32 (MLB200_PRECISION=fp32): plain fp32 FMAs on the CUDA cores (~1e-6).
 *   MLX_PRECISION_BF16 (MLB200_PRECISION=bf16): bf16 operands, fp32 accumulation on the tensor
 *     cores; activations between such layers are stored as bf16.  Stated bound 3e-2 of the range.
 * Image I/O stays ML_FLOAT32 / ML_FLOAT16 as declared by the graph.  May be called before or after
 * mlSetModelInputInfo.  mlxModelTensorCoreLayers reports how many layers run on the tensor cores. */
#define MLX_PRECISION_FP32 0
#define MLX_PRECISION_BF16 1
#define MLX_PRECISION_BF16X3 2
ML_API_ENTRY ml_status mlxSetModelPrecision(ml_model model, int precision);
ML_API_ENTRY size_t mlxModelTensorCoreLayers(ml_model model);

/* `count` independent mlInfer() calls as one pipelined batch (the reference API is batch-1 only,
 * model.cpp:208-213): the host->device copy of image i+1 and the device->host copy of result i-1
 * run on copy streams while the kernels of image i execute.  Per-image checks and error messages
 * are those of mlInfer, prefixed "Image <i>: ".  Synchronous like mlInfer: on ML_OK every output
 * is complete. */
ML_API_ENTRY ml_status mlxInferBatch(ml_model model, ml_image const* inputs, ml_image const* outputs, size_t count);

/* One mlInfer() with every fused op of the plan timed on the device (CUDA events, op by op).
 * op_ms / op_flops / op_bytes (each may be NULL) receive, for the first max_ops ops in execution
 * order (the order mlxInspectModel lists them), the duration in ms and the ALGORITHMIC flops and
 * bytes of the op; *n_ops receives the number of ops.  Measurement aid for roofline reports. */
ML_API_ENTRY ml_status mlxProfileModel(ml_model model, ml_image input, ml_image output, float* op_ms,
                                       double* op_flops, double* op_bytes, size_t max_ops, size_t* n_ops);

/* Device time of the last successful mlxInferBatch() in ms, CUDA events inside the library: from the first
 * image's upload being enqueued on the copy-in stream to the last image's read-back (or, when no host copy
 * is due, its last kernel) completing.  Measurement aid for bench.py. */
ML_API_ENTRY float mlxModelLastBatchMs(ml_model model);

/* Number of kernel launches one mlInfer() of this model enqueues. */
ML_API_ENTRY size_t mlxModelLaunchCount(ml_model model);

#ifdef __cplusplus
}
#endif

#endif /* MODEL_RUNNER_EXT_H_B200 */
