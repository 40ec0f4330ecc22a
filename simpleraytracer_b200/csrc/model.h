// model.h -- ML::Model: frozen-graph model (reference model.h:18-42) executed
// by hand-written CUDA kernels instead of a TensorFlow session.
#pragma once

#include "common.h"
#include "graphdef.h"
#include "nn_kernels.h"

#include <map>
#include <vector>

namespace ML {

class Image;

class Model
{
public:
    static ml_model MakeHandle(Model* m) { return reinterpret_cast<ml_model>(m); }
    static Model* FromHandle(ml_model m) { return reinterpret_cast<Model*>(m); }

    Model(DeviceStatePtr device, ml_model_params const* params);
    ~Model();
    Model(Model const&) = delete;
    Model& operator=(Model const&) = delete;

    ml_status GetInfo(ml_image_info* input_info, ml_image_info* output_info);
    ml_status SetInputInfo(ml_image_info const* info);
    ml_status Infer(ml_image input, ml_image output);
    // n independent inferences, pipelined: H2D of image i+1 and D2H of result i-1 overlap the
    // kernels of image i (three streams).  Same checks and messages as Infer, per image.
    ml_status InferBatch(ml_image const* inputs, ml_image const* outputs, size_t n);
    char* GetError(char* buffer, size_t buffer_size) const;
    ml_status Profile(ml_image input, ml_image output, float* op_ms, double* op_flops, double* op_bytes,
                      size_t max_ops, size_t* n_ops);

    // Arithmetic of the 3x3 / stride 1 / SAME layers conv_tc_kernels.cu supports: 0 = fp32 on the CUDA
    // cores, 1 = bf16 on the tensor cores (tcgen05, fp32 accumulation), 2 = bf16x3 (default): the
    // tensor cores with every fp32 operand split into two bf16 halves and three products per term --
    // fp32-grade results (~1e-5 of the range).  Initial value from MLB200_PRECISION=fp32|bf16|bf16x3;
    // mlxSetModelPrecision changes it later.
    ml_status SetPrecision(int precision);
    int Precision() const { return m_precision; }
    size_t TensorCoreLayers() const;

    // Number of kernel launches one Infer() enqueues (bench bookkeeping).
    size_t LaunchesPerInfer() const;
    float LastBatchMs() const { return m_last_batch_ms; }
    // Text dump of the detected nodes, shapes and fused plan (mlxInspectModel).
    std::string Describe() const;

private:
    // ---- execution plan ---------------------------------------------------
    // Pool .. Binary: the ops around the convolutions of an encoder / decoder graph (U-Net style denoisers)
    enum class OpKind { Conv, BiasAct, Add, Concat, Copy, Pool, Resize, DepthToSpace, Pad, Binary };

    struct TensorInfo
    {
        int C = 0;             // channels (static)
        int H = 0, W = 0;      // set by shape inference
        int buffer = -1;       // index into m_buffers; -1 = external (input / output image)
        int last_use = -1;     // index of the last op reading it
        bool bf16 = false;     // stored as bf16 NHWC (between tensor-core layers); else fp32
        bool split = false;    // stored as bf16 [hi: C | lo: C] per pixel (between bf16x3 layers; 4 bytes per element)
    };

    struct Op
    {
        OpKind kind;
        std::vector<int> in;   // tensor ids
        int out = -1;          // tensor id
        // Conv
        int KH = 0, KW = 0, stride = 1, dilation = 1;
        bool same_padding = true;
        int weights = -1;      // index into m_consts (HWIO filter)
        int bias = -1;         // index into m_consts or -1
        nn::Act act = nn::Act::None;
        float alpha = 0.2f;
        int scale = -1;        // BiasAct: per-channel multiplier applied BEFORE the bias (Mul by a constant, FusedBatchNorm)
        // Pool: KH = KW = window, stride, same_padding; pool_max.  Resize: bilinear / align_corners / half_pixel and the
        // output size, either absolute (size_h, size_w) or size_ref's (H, W) times (mul_h, mul_w) -- the
        // Shape -> StridedSlice -> Mul pattern Keras UpSampling2D exports.  DepthToSpace: block.  Pad: pads + mode.
        bool pool_max = false;
        bool bilinear = false, align_corners = false, half_pixel = false;
        int size_h = 0, size_w = 0, size_ref = -1, mul_h = 1, mul_w = 1;
        int block = 1;
        int pad_t = 0, pad_b = 0, pad_l = 0, pad_r = 0, pad_mode = 0;   // 0 CONSTANT, 1 REFLECT, 2 SYMMETRIC
        int binary = 0;        // Binary: 1 = a - b, 2 = a * b (a + b is OpKind::Add)
        int residual = -1;     // Conv: tensor added after the activation (a following Add folded in); also in `in`
        std::string add_name;  // name of the folded Add node, for its error message
        std::string name;      // for error messages
        // bf16 tensor-core execution (decided by AssignPrecision, bound by AllocateBuffers)
        bool tc = false;
        int packed = -1;       // index into m_packed (filter repacked [tap][co][ci] bf16)
        nn::ConvTc* tc_op = nullptr;
        bool pad_in = false;   // fp32 input with < 16 channels: zero-padded to 16 bf16 channels first
        void* pad_buffer = nullptr;
        // bf16x3 (precision 2): fp32 tensors either side; the input is split into [hi | lo] bf16 halves
        // in m_split_buffer right before the layer, the filter is packed as hi/lo halves
        bool split = false;
        int packed_split = -1; // index into m_packed
    };

    struct DeviceConst
    {
        std::vector<float> host;
        float* dev = nullptr;
    };

    struct Value // what a graph node evaluates to while the plan is built
    {
        bool is_const = false;
        int tensor = -1;                     // activation tensor id
        std::shared_ptr<pb::Tensor> constant;
        // an integer vector that depends on the input size: shape_ref's [N, H, W, C] (Shape), its [H, W] slice
        // (StridedSlice [1:3]) and that slice times constants (Mul) -- only what a resize's `size` input needs
        int shape_ref = -1;
        bool shape_hw = false;
        int shape_mul_h = 1, shape_mul_w = 1;
    };

    void BuildPlan(pb::Graph const& graph, int input_idx, int output_idx);
    Value Evaluate(pb::Graph const& graph, int node_idx, std::vector<Value>& memo,
                   std::vector<char>& state, int input_idx);
    void FusePlan();
    void InferShapes(size_t H, size_t W, size_t C); // throws on mismatch
    void AssignPrecision();
    void AllocateBuffers();
    void ReleaseBuffers();
    int AddConst(std::vector<float> values);
    int NewTensor(int channels);
    void Run(float const* d_in, float* d_out, cudaStream_t stream);
    void RunOp(Op const& op, float const* d_in, float* d_out, cudaStream_t stream);
    bool CheckInferArgs(ml_image input, ml_image output); // sets m_error, false on failure
    nn::ConvDesc MakeConvDesc(Op const& op) const;

    DeviceStatePtr m_device;
    std::string m_input_node;
    std::string m_output_node;
    ml_image_info m_input_info{};
    ml_image_info m_output_info{};
    bool m_ready = false;        // shapes inferred and buffers allocated
    bool m_loaded = false;       // plan built (the reference's "session != nullptr")

    std::vector<TensorInfo> m_tensors;
    std::vector<Op> m_ops;
    std::vector<DeviceConst> m_consts;
    std::map<std::string, int> m_name_index;
    int m_input_tensor = -1;
    int m_output_tensor = -1;

    int m_precision = 2;
    std::vector<cudaEvent_t> m_events; // pool for InferBatch
    cudaEvent_t m_batch_ev0 = nullptr, m_batch_ev1 = nullptr; // device span of the last InferBatch (timing enabled)
    float m_last_batch_ms = 0.f;
    std::vector<void*> m_packed;  // device copies of repacked bf16 filters
    std::vector<void*> m_buffers;
    std::vector<size_t> m_buffer_bytes;
    void* m_split_buffer = nullptr; // bf16 [H][W][hi|lo] input of the current bf16x3 layer (ops run one at a time)
    void* m_alias_in = nullptr;  // copy of the input for in-place inference through tensor-core layers (see Run)
    float* m_in_f32 = nullptr;   // staging when images are FLOAT16
    float* m_out_f32 = nullptr;
    std::string m_error;
};

} // namespace ML
