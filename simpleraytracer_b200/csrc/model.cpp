// model.cpp -- model object and its six C entry points.
//
// Contract (what callers can observe) follows the reference:
//   constructor       model.cpp:73-139   path checks, node auto-detection (first /
//                                        last node), dtype + last-3-dims discovery
//   GetInfo           model.cpp:141-159
//   SetInputInfo      model.cpp:161-235  dtype / known-dimension checks, no-op when
//                                        unchanged, output shape known afterwards
//   Infer             model.cpp:237-294  handle + output-dimension checks
//   InferToCache      model.cpp:301-347  unspecified-dimension + byte-size checks
//   C wrappers        model.cpp:352-395
// The arithmetic (reference: one tensorflow::Session::Run, model.cpp:339) is an
// execution plan of hand-written CUDA kernels over device-resident NHWC tensors;
// the input is read from and the result written to the images' device buffers.
#include "model.h"

#include <cmath>

#include "dtype.h"
#include "image.h"
#include "model_runner_ext.h"

#include <algorithm>
#include <map>

namespace ML {
namespace {

// Largest stride / dilation a graph may ask for (TensorFlow accepts any positive value; image networks use 1..4).
// The bound keeps the int arithmetic of shape inference and padding far from overflow for any frame that fits in memory.
constexpr int64_t kMaxStrideOrDilation = 64;
// Largest tensor side and element count a plan may reach: the kernels index elements with 32-bit arithmetic, and an
// untrusted graph must not be able to wrap the plan's int shape arithmetic (a resize by 2^31 - 1, say).
constexpr int64_t kMaxTensorSide = int64_t(1) << 20;
constexpr uint64_t kMaxTensorElements = uint64_t(1) << 31;

// dtype from attr "dtype" or "T"; H,W,C = last three dims of _output_shapes[0]
// with -1 -> 0 (reference model.cpp:23-48).  Divergence: a node without
// _output_shapes falls back to a Placeholder's "shape" attr, else all-unknown,
// instead of aborting inside protobuf's Map::at.
void FillImageInfo(pb::Node const& node, ml_image_info& info)
{
    pb::Attr const* dtype = node.Find("dtype");
    if (dtype == nullptr || !dtype->has_type) dtype = node.Find("T");
    if (dtype == nullptr || !dtype->has_type)
        throw std::runtime_error("Unable to detect node dtype: " + node.name);
    info.dtype = DataTypeFromTF(dtype->type);

    pb::Shape const* shape = nullptr;
    if (pb::Attr const* os = node.Find("_output_shapes"))
        if (!os->list_shape.empty()) shape = &os->list_shape.front();
    if (shape == nullptr)
        if (pb::Attr const* sh = node.Find("shape"))
            if (sh->has_shape) shape = &sh->shape;

    auto dim = [shape](int from_end) -> size_t {
        if (shape == nullptr) return 0;
        int n = static_cast<int>(shape->dims.size());
        int i = n - from_end;
        if (i < 0) return 0;
        int64_t v = shape->dims[static_cast<size_t>(i)];
        return v > 0 ? static_cast<size_t>(v) : 0;
    };
    info.height = dim(3);
    info.width = dim(2);
    info.channels = dim(1);
}

std::string CleanInputName(std::string const& raw)
{
    std::string name = raw;
    size_t colon = name.rfind(':');
    if (colon != std::string::npos)
    {
        if (name.substr(colon + 1) != "0")
            throw std::runtime_error("Unimplemented: output " + raw + " (only output :0 is supported)");
        name.erase(colon);
    }
    return name;
}

[[noreturn]] void Unimplemented(pb::Node const& node, std::string const& why)
{
    throw std::runtime_error("Error creating graph: Unimplemented: " + why + " (node '" + node.name
                             + "', op '" + node.op + "')");
}

} // namespace

// ---------------------------------------------------------------------------
// construction
// ---------------------------------------------------------------------------
Model::Model(DeviceStatePtr device, ml_model_params const* params) : m_device(std::move(device))
{
    if (params == nullptr) throw std::runtime_error("Bad parameters argument");
    if (params->model_path == nullptr) throw std::runtime_error("Bad model_path model parameter value");

    pb::Graph graph = pb::ReadGraphDefFile(params->model_path);

    int input_idx = 0;
    int output_idx = static_cast<int>(graph.nodes.size()) - 1;
    for (int i = 0; i < static_cast<int>(graph.nodes.size()); i++)
    {
        if (params->input_node != nullptr && graph.nodes[i].name == params->input_node) input_idx = i;
        if (params->output_node != nullptr && graph.nodes[i].name == params->output_node) output_idx = i;
    }
    FillImageInfo(graph.nodes[input_idx], m_input_info);
    FillImageInfo(graph.nodes[output_idx], m_output_info);
    m_input_node = graph.nodes[input_idx].name;
    m_output_node = graph.nodes[output_idx].name;

    BuildPlan(graph, input_idx, output_idx);
    FusePlan();
    if (char const* e = std::getenv("MLB200_PRECISION"))
    {
        std::string const v(e);
        if (v == "fp32" || v == "FP32") m_precision = 0;
        else if (v == "bf16" || v == "BF16") m_precision = 1;
        else if (v == "bf16x3" || v == "BF16X3") m_precision = 2;
    }

    // Upload constants once; they stay resident in HBM for the model's life.
    // (m_device == nullptr: host-only inspection, mlxInspectModel.)
    if (m_device) m_device->MakeCurrent();
    try
    {
        for (auto& c : m_consts)
            if (m_device)
        {
            ML_CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&c.dev), std::max<size_t>(c.host.size(), 1) * sizeof(float)));
            ML_CUDA_CHECK(cudaMemcpyAsync(c.dev, c.host.data(), c.host.size() * sizeof(float),
                                          cudaMemcpyHostToDevice, m_device->stream));
        }
        if (m_device) ML_CUDA_CHECK(cudaStreamSynchronize(m_device->stream));
    }
    catch (...)
    {
        for (auto& c : m_consts) cudaFree(c.dev);
        throw;
    }
    m_loaded = true;

    // Static channel propagation gives the output channel count even when the
    // output node carries no _output_shapes.
    try
    {
        InferShapes(m_input_info.height, m_input_info.width, m_input_info.channels);
        TensorInfo const& out = m_tensors[m_output_tensor];
        if (m_output_info.channels == 0) m_output_info.channels = static_cast<size_t>(out.C);
        if (m_input_info.height && m_input_info.width && m_input_info.channels)
        {
            // Fully specified graph: ready to run without mlSetModelInputInfo
            // (the reference would dereference an empty input map here).
            m_output_info.height = static_cast<size_t>(out.H);
            m_output_info.width = static_cast<size_t>(out.W);
            m_output_info.channels = static_cast<size_t>(out.C);
            if (m_device)
            {
                AllocateBuffers();
                m_ready = true;
            }
        }
    }
    catch (std::exception const& e)
    {
        // AllocateBuffers may have got part of the way: give everything back before the object disappears
        if (m_device)
        {
            ReleaseBuffers();
            for (void* p : m_packed) cudaFree(p);
            m_packed.clear();
        }
        for (auto& c : m_consts) cudaFree(c.dev);
        throw std::runtime_error(std::string("Error creating graph: Invalid argument: ") + e.what());
    }
}

Model::~Model()
{
    if (!m_device) return;
    cudaSetDevice(m_device->device);
    cudaStreamSynchronize(m_device->stream);
    ReleaseBuffers();
    for (void* p : m_packed) cudaFree(p);
    for (cudaEvent_t e : m_events) cudaEventDestroy(e);
    if (m_batch_ev0) cudaEventDestroy(m_batch_ev0);
    if (m_batch_ev1) cudaEventDestroy(m_batch_ev1);
    for (auto& c : m_consts) cudaFree(c.dev);
}

int Model::AddConst(std::vector<float> values)
{
    m_consts.push_back(DeviceConst{std::move(values), nullptr});
    return static_cast<int>(m_consts.size()) - 1;
}

int Model::NewTensor(int channels)
{
    TensorInfo t;
    t.C = channels;
    m_tensors.push_back(t);
    return static_cast<int>(m_tensors.size()) - 1;
}

void Model::BuildPlan(pb::Graph const& graph, int input_idx, int output_idx)
{
    std::vector<Value> memo(graph.nodes.size());
    std::vector<char> state(graph.nodes.size(), 0);
    m_name_index.clear();
    for (int i = 0; i < static_cast<int>(graph.nodes.size()); i++) m_name_index[graph.nodes[i].name] = i;
    Value out = Evaluate(graph, output_idx, memo, state, input_idx);
    if (out.is_const) Unimplemented(graph.nodes[output_idx], "output node is a constant");
    if (m_input_tensor < 0)
        throw std::runtime_error("Error creating graph: Invalid argument: output node '"
                                 + graph.nodes[output_idx].name + "' does not depend on input node '"
                                 + graph.nodes[input_idx].name + "'");
    m_output_tensor = out.tensor;
    if (m_output_tensor == m_input_tensor)
    {
        Op op;
        op.kind = OpKind::Copy;
        op.in = {m_input_tensor};
        op.out = NewTensor(m_tensors[m_input_tensor].C);
        op.name = graph.nodes[output_idx].name;
        m_ops.push_back(op);
        m_output_tensor = op.out;
    }
}

Model::Value Model::Evaluate(pb::Graph const& graph, int idx, std::vector<Value>& memo,
                             std::vector<char>& state, int input_idx)
{
    if (state[idx] == 2) return memo[idx];
    pb::Node const& node = graph.nodes[idx];
    if (state[idx] == 1) Unimplemented(node, "graph cycle");
    state[idx] = 1;

    auto input_value = [&](size_t k) -> Value {
        size_t seen = 0;
        for (auto const& raw : node.inputs)
        {
            if (!raw.empty() && raw[0] == '^') continue; // control dependency
            if (seen++ == k)
            {
                std::string name = CleanInputName(raw);
                auto it = m_name_index.find(name);
                if (it == m_name_index.end()) Unimplemented(node, "unknown input '" + raw + "'");
                return Evaluate(graph, it->second, memo, state, input_idx);
            }
        }
        Unimplemented(node, "missing input #" + std::to_string(k));
    };
    auto data_input_count = [&]() {
        size_t n = 0;
        for (auto const& raw : node.inputs)
            if (raw.empty() || raw[0] != '^') n++;
        return n;
    };
    auto activation = [&](size_t k) -> int {
        Value v = input_value(k);
        if (v.is_const) Unimplemented(node, "constant activation input #" + std::to_string(k));
        return v.tensor;
    };
    auto parse_act = [&](std::string const& op, nn::Act& act, float& alpha) -> bool {
        alpha = 0.2f;
        if (op == "Relu") act = nn::Act::Relu;
        else if (op == "Relu6") act = nn::Act::Relu6;
        else if (op == "Sigmoid") act = nn::Act::Sigmoid;
        else if (op == "Tanh") act = nn::Act::Tanh;
        else if (op == "LeakyRelu")
        {
            act = nn::Act::LeakyRelu;
            if (pb::Attr const* a = node.Find("alpha"))
                if (a->has_f) alpha = a->f;
        }
        else return false;
        return true;
    };

    Value result;
    nn::Act act;
    float alpha;
    if (idx == input_idx)
    {
        if (m_input_tensor < 0) m_input_tensor = NewTensor(static_cast<int>(m_input_info.channels));
        result.tensor = m_input_tensor;
    }
    else if (node.op == "Placeholder")
    {
        Unimplemented(node, "more than one graph input");
    }
    else if (node.op == "Const")
    {
        pb::Attr const* v = node.Find("value");
        if (v == nullptr || !v->tensor) Unimplemented(node, "Const without value");
        result.is_const = true;
        result.constant = v->tensor;
    }
    else if (node.op == "Identity" || node.op == "StopGradient" || node.op == "PlaceholderWithDefault")
    {
        result = input_value(0);
    }
    else if (node.op == "Conv2D")
    {
        int in = activation(0);
        Value f = input_value(1);
        if (!f.is_const || f.constant->shape.dims.size() != 4 || f.constant->values.empty())
            Unimplemented(node, "Conv2D filter must be a 4-D float constant");
        if (pb::Attr const* df = node.Find("data_format"))
            if (df->has_s && df->s != "NHWC") Unimplemented(node, "data_format " + df->s);
        Op op;
        op.kind = OpKind::Conv;
        op.name = node.name;
        op.in = {in};
        // every dim must fit an int and the values must be exactly KH*KW*Cin*Cout (untrusted file)
        for (int64_t d : f.constant->shape.dims)
            if (d < 1 || d > (int64_t(1) << 20)) Unimplemented(node, "Conv2D filter dimension " + std::to_string(d));
        if (f.constant->values.size() != f.constant->ElementCount())
            Unimplemented(node, "Conv2D filter holds " + std::to_string(f.constant->values.size()) + " values for "
                                    + std::to_string(f.constant->ElementCount()) + " elements");
        op.KH = static_cast<int>(f.constant->shape.dims[0]);
        op.KW = static_cast<int>(f.constant->shape.dims[1]);
        int cin = static_cast<int>(f.constant->shape.dims[2]);
        int cout = static_cast<int>(f.constant->shape.dims[3]);
        if (pb::Attr const* s = node.Find("strides"))
            if (s->list_i.size() == 4)
            {
                // bounded above as well: shape inference does int arithmetic on (H + stride - 1) and (K - 1) * dilation
                if (s->list_i[0] != 1 || s->list_i[3] != 1 || s->list_i[1] != s->list_i[2] || s->list_i[1] < 1
                    || s->list_i[1] > kMaxStrideOrDilation)
                    Unimplemented(node, "strides");
                op.stride = static_cast<int>(s->list_i[1]);
            }
        if (pb::Attr const* dl = node.Find("dilations"))
            if (dl->list_i.size() == 4)
            {
                if (dl->list_i[0] != 1 || dl->list_i[3] != 1 || dl->list_i[1] != dl->list_i[2] || dl->list_i[1] < 1
                    || dl->list_i[1] > kMaxStrideOrDilation)
                    Unimplemented(node, "dilations");
                op.dilation = static_cast<int>(dl->list_i[1]);
            }
        pb::Attr const* pad = node.Find("padding");
        if (pad == nullptr || !pad->has_s || (pad->s != "SAME" && pad->s != "VALID"))
            Unimplemented(node, "padding");
        op.same_padding = pad->s == "SAME";
        int known = m_tensors[in].C;
        if (known != 0 && known != cin)
            throw std::runtime_error("Error creating graph: Invalid argument: Conv2D '" + node.name
                                     + "' expects " + std::to_string(cin) + " input channels, got "
                                     + std::to_string(known));
        op.weights = AddConst(f.constant->values);
        op.out = NewTensor(cout);
        m_ops.push_back(op);
        result.tensor = op.out;
    }
    else if (node.op == "BiasAdd" || node.op == "Add" || node.op == "AddV2")
    {
        if (data_input_count() != 2) Unimplemented(node, "expected two inputs");
        Value a = input_value(0), b = input_value(1);
        if (a.is_const && b.is_const) Unimplemented(node, "constant folding");
        if (a.is_const) std::swap(a, b);
        if (b.is_const)
        {
            size_t n = b.constant->values.size();
            int C = m_tensors[a.tensor].C;
            if (n == 0 || (n != 1 && C != 0 && static_cast<int>(n) != C))
                Unimplemented(node, "bias of " + std::to_string(n) + " values on " + std::to_string(C) + " channels");
            Op op;
            op.kind = OpKind::BiasAct;
            op.name = node.name;
            op.in = {a.tensor};
            std::vector<float> bias = b.constant->values;
            if (n == 1 && C > 1) bias.assign(static_cast<size_t>(C), bias[0]);
            op.bias = AddConst(std::move(bias));
            op.out = NewTensor(C != 0 ? C : static_cast<int>(n));
            m_ops.push_back(op);
            result.tensor = op.out;
        }
        else
        {
            if (node.op == "BiasAdd") Unimplemented(node, "BiasAdd with a non-constant bias");
            Op op;
            op.kind = OpKind::Add;
            op.name = node.name;
            op.in = {a.tensor, b.tensor};
            op.out = NewTensor(m_tensors[a.tensor].C);
            m_ops.push_back(op);
            result.tensor = op.out;
        }
    }
    else if (parse_act(node.op, act, alpha))
    {
        Op op;
        op.kind = OpKind::BiasAct;
        op.name = node.name;
        op.in = {activation(0)};
        op.act = act;
        op.alpha = alpha;
        op.out = NewTensor(m_tensors[op.in[0]].C);
        m_ops.push_back(op);
        result.tensor = op.out;
    }
    else if (node.op == "ConcatV2")
    {
        size_t n = data_input_count();
        if (n < 2) Unimplemented(node, "ConcatV2 without inputs");
        Value axis = input_value(n - 1);
        if (!axis.is_const || axis.constant->ints.size() != 1) Unimplemented(node, "ConcatV2 axis");
        int64_t ax = axis.constant->ints[0];
        if (ax != 3 && ax != -1) Unimplemented(node, "ConcatV2 on axis " + std::to_string(ax));
        Op op;
        op.kind = OpKind::Concat;
        op.name = node.name;
        int total = 0;
        for (size_t k = 0; k + 1 < n; k++)
        {
            int t = activation(k);
            if (m_tensors[t].C == 0) Unimplemented(node, "ConcatV2 of tensors with unknown channel count");
            op.in.push_back(t);
            total += m_tensors[t].C;
        }
        op.out = NewTensor(total);
        m_ops.push_back(op);
        result.tensor = op.out;
    }
    else if (node.op == "MaxPool" || node.op == "AvgPool")
    {
        if (pb::Attr const* df = node.Find("data_format"))
            if (df->has_s && df->s != "NHWC") Unimplemented(node, "data_format " + df->s);
        pb::Attr const* ks = node.Find("ksize");
        pb::Attr const* st = node.Find("strides");
        pb::Attr const* pad = node.Find("padding");
        if (ks == nullptr || ks->list_i.size() != 4 || ks->list_i[0] != 1 || ks->list_i[3] != 1
            || ks->list_i[1] != ks->list_i[2] || ks->list_i[1] < 1 || ks->list_i[1] > 16)
            Unimplemented(node, "ksize (square windows over H and W only)");
        if (st == nullptr || st->list_i.size() != 4 || st->list_i[0] != 1 || st->list_i[3] != 1
            || st->list_i[1] != st->list_i[2] || st->list_i[1] < 1 || st->list_i[1] > kMaxStrideOrDilation)
            Unimplemented(node, "strides");
        if (pad == nullptr || !pad->has_s || (pad->s != "SAME" && pad->s != "VALID")) Unimplemented(node, "padding");
        Op op;
        op.kind = OpKind::Pool;
        op.name = node.name;
        op.in = {activation(0)};
        op.KH = op.KW = static_cast<int>(ks->list_i[1]);
        op.stride = static_cast<int>(st->list_i[1]);
        op.same_padding = pad->s == "SAME";
        op.pool_max = node.op == "MaxPool";
        op.out = NewTensor(m_tensors[op.in[0]].C);
        m_ops.push_back(op);
        result.tensor = op.out;
    }
    else if (node.op == "ResizeNearestNeighbor" || node.op == "ResizeBilinear")
    {
        Op op;
        op.kind = OpKind::Resize;
        op.name = node.name;
        op.in = {activation(0)};
        op.bilinear = node.op == "ResizeBilinear";
        if (pb::Attr const* a = node.Find("align_corners")) op.align_corners = a->has_b && a->b;
        if (pb::Attr const* a = node.Find("half_pixel_centers")) op.half_pixel = a->has_b && a->b;
        if (op.align_corners && op.half_pixel) Unimplemented(node, "align_corners together with half_pixel_centers");
        Value size = input_value(1);
        if (size.is_const)
        {
            if (size.constant->ints.size() != 2 || size.constant->ints[0] < 1 || size.constant->ints[1] < 1)
                Unimplemented(node, "size must hold two positive integers");
            if (size.constant->ints[0] > kMaxTensorSide || size.constant->ints[1] > kMaxTensorSide)
                Unimplemented(node, "size beyond " + std::to_string(kMaxTensorSide) + " rows or columns");
            op.size_h = static_cast<int>(size.constant->ints[0]);
            op.size_w = static_cast<int>(size.constant->ints[1]);
        }
        else if (size.shape_ref >= 0 && size.shape_hw)
        {
            op.size_ref = size.shape_ref;
            op.mul_h = size.shape_mul_h;
            op.mul_w = size.shape_mul_w;
        }
        else Unimplemented(node, "size must be a constant or Shape(x)[1:3] * constant");
        op.out = NewTensor(m_tensors[op.in[0]].C);
        m_ops.push_back(op);
        result.tensor = op.out;
    }
    else if (node.op == "Shape")
    {
        result.shape_ref = activation(0);
    }
    else if (node.op == "StridedSlice" || node.op == "Mul" || node.op == "Sub" || node.op == "RealDiv")
    {
        Value a = input_value(0);
        if (node.op == "StridedSlice")
        {
            // only Shape(x)[1:3]: the (H, W) of a tensor
            if (a.shape_ref < 0 || a.shape_hw) Unimplemented(node, "StridedSlice of anything but a Shape");
            Value b = input_value(1), e = input_value(2);
            bool ok = b.is_const && e.is_const && b.constant->ints.size() == 1 && e.constant->ints.size() == 1
                      && b.constant->ints[0] == 1 && e.constant->ints[0] == 3;
            if (data_input_count() > 3)
            {
                Value st = input_value(3);
                ok = ok && st.is_const && st.constant->ints.size() == 1 && st.constant->ints[0] == 1;
            }
            for (char const* mask : {"begin_mask", "end_mask", "ellipsis_mask", "new_axis_mask", "shrink_axis_mask"})
                if (pb::Attr const* m = node.Find(mask)) ok = ok && (!m->has_i || m->i == 0);
            if (!ok) Unimplemented(node, "StridedSlice other than Shape(x)[1:3]");
            result = a;
            result.shape_hw = true;
        }
        else
        {
            Value b = input_value(1);
            if (node.op == "Mul" && ((a.shape_ref >= 0) != (b.shape_ref >= 0)))
            {
                // Shape(x)[1:3] * [sy, sx]
                Value const& sh = a.shape_ref >= 0 ? a : b;
                Value const& k = a.shape_ref >= 0 ? b : a;
                if (!sh.shape_hw || !k.is_const || (k.constant->ints.size() != 2 && k.constant->ints.size() != 1)
                    || k.constant->ints[0] < 1 || k.constant->ints.back() < 1)
                    Unimplemented(node, "shape arithmetic other than Shape(x)[1:3] * positive constants");
                int64_t const mh = static_cast<int64_t>(sh.shape_mul_h) * k.constant->ints[0];
                int64_t const mw = static_cast<int64_t>(sh.shape_mul_w) * k.constant->ints.back();
                if (k.constant->ints[0] > kMaxTensorSide || k.constant->ints.back() > kMaxTensorSide || mh > kMaxTensorSide
                    || mw > kMaxTensorSide)
                    Unimplemented(node, "shape factor beyond " + std::to_string(kMaxTensorSide));
                result = sh;
                result.shape_mul_h = static_cast<int>(mh);
                result.shape_mul_w = static_cast<int>(mw);
            }
            else if (a.shape_ref >= 0 || b.shape_ref >= 0) Unimplemented(node, "shape arithmetic");
            else if (a.is_const && b.is_const) Unimplemented(node, "constant folding");
            else if (a.is_const || b.is_const)
            {
                // tensor (op) per-channel or scalar constant: out = in * scale + bias
                bool const const_first = a.is_const;
                Value const& x = const_first ? b : a;
                Value const& k = const_first ? a : b;
                size_t const n = k.constant->values.size();
                int const C = m_tensors[x.tensor].C;
                if (n == 0 || (n != 1 && C != 0 && static_cast<int>(n) != C))
                    Unimplemented(node, "constant of " + std::to_string(n) + " values on " + std::to_string(C) + " channels");
                size_t const len = (n == 1 && C > 1) ? static_cast<size_t>(C) : n;
                std::vector<float> scale(len, 1.0f), bias(len, 0.0f);
                bool use_scale = false, use_bias = false;
                for (size_t c = 0; c < len; c++)
                {
                    float const v = k.constant->values[n == 1 ? 0 : c];
                    if (node.op == "Mul") { scale[c] = v; use_scale = true; }
                    else if (node.op == "Sub")
                    {
                        if (const_first) { scale[c] = -1.0f; bias[c] = v; use_scale = true; }   // k - x
                        else bias[c] = -v;                                                      // x - k
                        use_bias = true;
                    }
                    else  // RealDiv
                    {
                        if (const_first) Unimplemented(node, "constant / tensor");
                        scale[c] = 1.0f / v;      // x / k as x * (1 / k): within 1 ulp of the division
                        use_scale = true;
                    }
                }
                Op op;
                op.kind = OpKind::BiasAct;
                op.name = node.name;
                op.in = {x.tensor};
                if (use_scale) op.scale = AddConst(std::move(scale));
                if (use_bias) op.bias = AddConst(std::move(bias));
                op.out = NewTensor(C != 0 ? C : static_cast<int>(len));
                m_ops.push_back(op);
                result.tensor = op.out;
            }
            else
            {
                if (node.op == "RealDiv") Unimplemented(node, "tensor / tensor");
                Op op;
                op.kind = OpKind::Binary;
                op.name = node.name;
                op.binary = node.op == "Sub" ? 1 : 2;
                op.in = {a.tensor, b.tensor};
                op.out = NewTensor(m_tensors[a.tensor].C);
                m_ops.push_back(op);
                result.tensor = op.out;
            }
        }
    }
    else if (node.op == "DepthToSpace")
    {
        if (pb::Attr const* df = node.Find("data_format"))
            if (df->has_s && df->s != "NHWC") Unimplemented(node, "data_format " + df->s);
        pb::Attr const* bs = node.Find("block_size");
        if (bs == nullptr || !bs->has_i || bs->i < 2 || bs->i > 8) Unimplemented(node, "block_size");
        Op op;
        op.kind = OpKind::DepthToSpace;
        op.name = node.name;
        op.in = {activation(0)};
        op.block = static_cast<int>(bs->i);
        int const C = m_tensors[op.in[0]].C;
        if (C == 0 || C % (op.block * op.block) != 0)
            Unimplemented(node, std::to_string(C) + " channels with block_size " + std::to_string(op.block));
        op.out = NewTensor(C / (op.block * op.block));
        m_ops.push_back(op);
        result.tensor = op.out;
    }
    else if (node.op == "Pad" || node.op == "PadV2" || node.op == "MirrorPad")
    {
        Op op;
        op.kind = OpKind::Pad;
        op.name = node.name;
        op.in = {activation(0)};
        Value pads = input_value(1);
        if (!pads.is_const || pads.constant->ints.size() != 8) Unimplemented(node, "paddings must be a constant [4, 2]");
        auto const& pv = pads.constant->ints;
        if (pv[0] != 0 || pv[1] != 0 || pv[6] != 0 || pv[7] != 0) Unimplemented(node, "padding of the batch or channel axis");
        for (int k = 2; k < 6; k++)
            if (pv[k] < 0 || pv[k] > 4096) Unimplemented(node, "padding amount");
        op.pad_t = static_cast<int>(pv[2]); op.pad_b = static_cast<int>(pv[3]);
        op.pad_l = static_cast<int>(pv[4]); op.pad_r = static_cast<int>(pv[5]);
        if (node.op == "MirrorPad")
        {
            pb::Attr const* m = node.Find("mode");
            if (m == nullptr || !m->has_s || (m->s != "REFLECT" && m->s != "SYMMETRIC")) Unimplemented(node, "mode");
            op.pad_mode = m->s == "REFLECT" ? 1 : 2;
        }
        else if (node.op == "PadV2")
        {
            Value cv = input_value(2);
            if (!cv.is_const || cv.constant->values.size() != 1 || cv.constant->values[0] != 0.0f)
                Unimplemented(node, "PadV2 with a non-zero constant");
        }
        op.out = NewTensor(m_tensors[op.in[0]].C);
        m_ops.push_back(op);
        result.tensor = op.out;
    }
    else if (node.op == "FusedBatchNorm" || node.op == "FusedBatchNormV2" || node.op == "FusedBatchNormV3")
    {
        // inference form: y = (x - mean) * scale / sqrt(variance + epsilon) + offset, as one per-channel multiply-add
        if (pb::Attr const* df = node.Find("data_format"))
            if (df->has_s && df->s != "NHWC") Unimplemented(node, "data_format " + df->s);
        if (pb::Attr const* tr = node.Find("is_training"))
            if (tr->has_b && tr->b) Unimplemented(node, "is_training = true");
        int const x = activation(0);
        Value sc = input_value(1), of = input_value(2), mean = input_value(3), var = input_value(4);
        if (!sc.is_const || !of.is_const || !mean.is_const || !var.is_const) Unimplemented(node, "non-constant statistics");
        size_t const n = sc.constant->values.size();
        if (n == 0 || of.constant->values.size() != n || mean.constant->values.size() != n || var.constant->values.size() != n
            || (m_tensors[x].C != 0 && static_cast<size_t>(m_tensors[x].C) != n))
            Unimplemented(node, "statistics do not match the channel count");
        float eps = 1e-3f;
        if (pb::Attr const* e = node.Find("epsilon")) if (e->has_f) eps = e->f;
        std::vector<float> scale(n), bias(n);
        for (size_t c = 0; c < n; c++)
        {
            double const inv = static_cast<double>(sc.constant->values[c])
                               / std::sqrt(static_cast<double>(var.constant->values[c]) + static_cast<double>(eps));
            scale[c] = static_cast<float>(inv);
            bias[c] = static_cast<float>(static_cast<double>(of.constant->values[c]) - static_cast<double>(mean.constant->values[c]) * inv);
        }
        Op op;
        op.kind = OpKind::BiasAct;
        op.name = node.name;
        op.in = {x};
        op.scale = AddConst(std::move(scale));
        op.bias = AddConst(std::move(bias));
        op.out = NewTensor(static_cast<int>(n));
        m_ops.push_back(op);
        result.tensor = op.out;
    }
    else if (node.op == "Squeeze" || node.op == "ExpandDims" || node.op == "Reshape")
    {
        // N = 1 and dense NHWC: dropping / adding the batch axis or restating [1, H, W, C] moves no data.
        if (node.op == "Reshape")
        {
            Value shp = input_value(1);
            int const C = m_tensors[activation(0)].C;
            bool ok = shp.is_const && (shp.constant->ints.size() == 4 || shp.constant->ints.size() == 3);
            if (ok)
            {
                auto const& d = shp.constant->ints;
                if (d.size() == 4 && d[0] != 1 && d[0] != -1) ok = false;
                if (d.back() != -1 && C != 0 && d.back() != C) ok = false;    // the channel count must stay
            }
            if (!ok) Unimplemented(node, "Reshape other than to [1, H, W, C] / [H, W, C] with unchanged channels");
        }
        else if (node.op == "Squeeze")
        {
            if (pb::Attr const* sd = node.Find("squeeze_dims"))
                for (int64_t d : sd->list_i)
                    if (d != 0 && d != -4) Unimplemented(node, "squeeze of an axis other than the batch");
        }
        result = input_value(0);
        if (result.is_const) Unimplemented(node, "constant folding");
    }
    else
    {
        Unimplemented(node, "unsupported op");
    }

    state[idx] = 2;
    memo[idx] = result;
    return result;
}

// Folds BiasAdd and activations into the producing op when nothing else reads
// the intermediate tensor: Conv2D+BiasAdd+Relu becomes one kernel launch.
void Model::FusePlan()
{
    bool changed = true;
    while (changed)
    {
        changed = false;
        std::vector<int> uses(m_tensors.size(), 0);
        std::vector<int> producer(m_tensors.size(), -1);
        for (size_t i = 0; i < m_ops.size(); i++)
        {
            for (int t : m_ops[i].in) uses[t]++;
            producer[m_ops[i].out] = static_cast<int>(i);
        }
        uses[m_output_tensor]++;
        for (size_t i = 0; i < m_ops.size() && !changed; i++)
        {
            Op& op = m_ops[i];
            if (op.kind != OpKind::BiasAct) continue;
            int src = op.in[0];
            int p = producer[src];
            if (p < 0 || uses[src] != 1) continue;
            Op& prod = m_ops[p];
            if (prod.act != nn::Act::None) continue; // activation already applied: order matters
            if (prod.residual >= 0) continue;        // the residual is added after the activation
            bool const has_bias = op.bias >= 0;
            if (op.scale >= 0)
            {
                // out = prod(...) * s + b: a convolution takes s into its filter (per output channel) and bias, a
                // multiply-add composes; anything else keeps the op (FusedBatchNorm after Conv2D costs nothing)
                auto const& sc = m_consts[op.scale].host;
                if (prod.kind == OpKind::Conv)
                {
                    auto& w = m_consts[prod.weights].host;
                    size_t const cout = sc.size();
                    if (cout == 0 || w.size() % cout != 0 || static_cast<int>(cout) != m_tensors[prod.out].C) continue;
                    for (size_t k = 0; k < w.size(); k++) w[k] *= sc[k % cout];       // HWIO: output channel fastest
                    if (prod.bias >= 0)
                        for (size_t k = 0; k < cout; k++) m_consts[prod.bias].host[k] *= sc[k];
                }
                else if (prod.kind == OpKind::BiasAct)
                {
                    if (prod.bias >= 0 && m_consts[prod.bias].host.size() != sc.size()) continue;
                    if (prod.scale >= 0 && m_consts[prod.scale].host.size() != sc.size()) continue;
                    if (prod.bias >= 0)
                        for (size_t k = 0; k < sc.size(); k++) m_consts[prod.bias].host[k] *= sc[k];
                    if (prod.scale >= 0)
                        for (size_t k = 0; k < sc.size(); k++) m_consts[prod.scale].host[k] *= sc[k];
                    else prod.scale = op.scale;
                }
                else continue;
            }
            if (prod.kind == OpKind::Conv || prod.kind == OpKind::BiasAct)
            {
                if (has_bias && prod.bias >= 0)
                {
                    // bias + bias: add the vectors
                    auto& a = m_consts[prod.bias].host;
                    auto const& b = m_consts[op.bias].host;
                    if (a.size() != b.size()) continue;
                    for (size_t k = 0; k < a.size(); k++) a[k] += b[k];
                }
                else if (has_bias) prod.bias = op.bias;
            }
            else if (prod.kind == OpKind::Add)
            {
                if (has_bias) continue;
            }
            else continue;
            prod.act = op.act;
            prod.alpha = op.alpha;
            prod.out = op.out;
            m_ops.erase(m_ops.begin() + static_cast<long>(i));
            changed = true;
        }
        // Add(conv(...), r) -> the convolution's epilogue adds r (a global residual costs no pass of its
        // own).  Only when nothing else reads the conv result, r exists before the conv runs, and no
        // activation sits on the Add (the fused order is act(conv + bias) + r, exactly the graph's).
        for (size_t i = 0; i < m_ops.size() && !changed; i++)
        {
            Op& op = m_ops[i];
            if (op.kind != OpKind::Add || op.in.size() != 2 || op.in[0] == op.in[1]) continue;
            if (op.act != nn::Act::None) continue;
            for (int k = 0; k < 2 && !changed; k++)
            {
                int const conv_t = op.in[k], res_t = op.in[1 - k];
                int const p = producer[conv_t];
                if (p < 0 || uses[conv_t] != 1) continue;
                Op& prod = m_ops[p];
                if (prod.kind != OpKind::Conv || prod.residual >= 0) continue;
                if (res_t != m_input_tensor && (producer[res_t] < 0 || producer[res_t] > p)) continue;
                prod.residual = res_t;
                prod.in.push_back(res_t);
                prod.add_name = op.name;
                prod.out = op.out;
                m_ops.erase(m_ops.begin() + static_cast<long>(i));
                changed = true;
            }
        }
    }
}

// ---------------------------------------------------------------------------
// shapes and memory
// ---------------------------------------------------------------------------
void Model::InferShapes(size_t H, size_t W, size_t C)
{
    for (auto& t : m_tensors) { t.H = 0; t.W = 0; t.last_use = -1; }
    TensorInfo& in = m_tensors[m_input_tensor];
    if (C != 0)
    {
        if (in.C != 0 && in.C != static_cast<int>(C))
            throw std::runtime_error("input has " + std::to_string(C) + " channels, the graph expects "
                                     + std::to_string(in.C));
        in.C = static_cast<int>(C);
    }
    auto too_large = [](int64_t h, int64_t w, int64_t c) {
        return h < 0 || w < 0 || h > kMaxTensorSide || w > kMaxTensorSide
               || static_cast<uint64_t>(h) * static_cast<uint64_t>(w) * static_cast<uint64_t>(std::max<int64_t>(c, 1)) >= kMaxTensorElements;
    };
    if (H > static_cast<size_t>(kMaxTensorSide) || W > static_cast<size_t>(kMaxTensorSide)
        || too_large(static_cast<int64_t>(H), static_cast<int64_t>(W), in.C))
        throw std::runtime_error("input of " + std::to_string(W) + " x " + std::to_string(H) + " x " + std::to_string(in.C)
                                 + " exceeds the largest supported tensor (2^31 elements, 2^20 rows or columns)");
    in.H = static_cast<int>(H);
    in.W = static_cast<int>(W);

    for (size_t i = 0; i < m_ops.size(); i++)
    {
        Op& op = m_ops[i];
        for (int t : op.in) m_tensors[t].last_use = static_cast<int>(i);
        TensorInfo const& a = m_tensors[op.in[0]];
        TensorInfo& o = m_tensors[op.out];
        switch (op.kind)
        {
            case OpKind::Conv:
            {
                size_t wsize = m_consts[op.weights].host.size();
                int cin = static_cast<int>(wsize / (static_cast<size_t>(op.KH) * op.KW * o.C));
                if (a.C != 0 && a.C != cin)
                    throw std::runtime_error("Conv2D '" + op.name + "' expects " + std::to_string(cin)
                                             + " input channels, got " + std::to_string(a.C));
                auto out_dim = [&](int n, int k) {
                    if (n == 0) return 0;
                    int eff = (k - 1) * op.dilation + 1;
                    if (op.same_padding) return (n + op.stride - 1) / op.stride;
                    return n >= eff ? (n - eff) / op.stride + 1 : 0;
                };
                o.H = out_dim(a.H, op.KH);
                o.W = out_dim(a.W, op.KW);
                if ((a.H && !o.H) || (a.W && !o.W))
                    throw std::runtime_error("Conv2D '" + op.name + "' output is empty for a "
                                             + std::to_string(a.W) + " x " + std::to_string(a.H) + " input");
                if (op.residual >= 0)
                {
                    TensorInfo const& r = m_tensors[op.residual];
                    if (r.C != o.C || r.H != o.H || r.W != o.W)
                        throw std::runtime_error("'" + op.add_name + "' operands differ in shape");
                }
                break;
            }
            case OpKind::BiasAct:
            case OpKind::Copy:
                if (o.C == 0) o.C = a.C;
                if (op.bias >= 0 && a.C != 0 && static_cast<int>(m_consts[op.bias].host.size()) != a.C)
                    throw std::runtime_error("'" + op.name + "' bias size does not match "
                                             + std::to_string(a.C) + " channels");
                o.H = a.H; o.W = a.W;
                break;
            case OpKind::Add:
            {
                TensorInfo const& b = m_tensors[op.in[1]];
                if (a.C != b.C || a.H != b.H || a.W != b.W)
                    throw std::runtime_error("'" + op.name + "' operands differ in shape");
                if (o.C == 0) o.C = a.C;
                o.H = a.H; o.W = a.W;
                break;
            }
            case OpKind::Concat:
                for (int t : op.in)
                    if (m_tensors[t].H != a.H || m_tensors[t].W != a.W)
                        throw std::runtime_error("'" + op.name + "' operands differ in spatial size");
                o.H = a.H; o.W = a.W;
                break;
            case OpKind::Pool:
            {
                auto out_dim = [&](int n) {
                    if (n == 0) return 0;
                    if (op.same_padding) return (n + op.stride - 1) / op.stride;
                    return n >= op.KH ? (n - op.KH) / op.stride + 1 : 0;
                };
                if (o.C == 0) o.C = a.C;
                o.H = out_dim(a.H); o.W = out_dim(a.W);
                if ((a.H && !o.H) || (a.W && !o.W))
                    throw std::runtime_error("'" + op.name + "' output is empty for a " + std::to_string(a.W) + " x "
                                             + std::to_string(a.H) + " input");
                break;
            }
            case OpKind::Resize:
                if (o.C == 0) o.C = a.C;
                if (op.size_ref >= 0)
                {
                    int64_t const h = static_cast<int64_t>(m_tensors[op.size_ref].H) * op.mul_h;
                    int64_t const w = static_cast<int64_t>(m_tensors[op.size_ref].W) * op.mul_w;
                    if (too_large(h, w, o.C))
                        throw std::runtime_error("'" + op.name + "' output of " + std::to_string(w) + " x " + std::to_string(h)
                                                 + " exceeds the largest supported tensor");
                    o.H = static_cast<int>(h);
                    o.W = static_cast<int>(w);
                }
                else if (a.H != 0 && a.W != 0) { o.H = op.size_h; o.W = op.size_w; }
                break;
            case OpKind::DepthToSpace:
                if (too_large(static_cast<int64_t>(a.H) * op.block, static_cast<int64_t>(a.W) * op.block, o.C))
                    throw std::runtime_error("'" + op.name + "' output exceeds the largest supported tensor");
                o.H = a.H * op.block; o.W = a.W * op.block;
                break;
            case OpKind::Pad:
                if (o.C == 0) o.C = a.C;
                if (a.H != 0 && a.W != 0)
                {
                    if (op.pad_mode != 0)
                    {
                        int const lim_h = op.pad_mode == 1 ? a.H - 1 : a.H, lim_w = op.pad_mode == 1 ? a.W - 1 : a.W;
                        if (op.pad_t > lim_h || op.pad_b > lim_h || op.pad_l > lim_w || op.pad_r > lim_w)
                            throw std::runtime_error("'" + op.name + "' mirrors more rows or columns than the input has");
                    }
                    if (too_large(static_cast<int64_t>(a.H) + op.pad_t + op.pad_b, static_cast<int64_t>(a.W) + op.pad_l + op.pad_r, o.C))
                        throw std::runtime_error("'" + op.name + "' output exceeds the largest supported tensor");
                    o.H = a.H + op.pad_t + op.pad_b; o.W = a.W + op.pad_l + op.pad_r;
                }
                break;
            case OpKind::Binary:
            {
                TensorInfo const& b = m_tensors[op.in[1]];
                if (a.C != b.C || a.H != b.H || a.W != b.W)
                    throw std::runtime_error("'" + op.name + "' operands differ in shape");
                if (o.C == 0) o.C = a.C;
                o.H = a.H; o.W = a.W;
                break;
            }
        }
        if (too_large(o.H, o.W, o.C))
            throw std::runtime_error("'" + op.name + "' output of " + std::to_string(o.W) + " x " + std::to_string(o.H) + " x "
                                     + std::to_string(o.C) + " exceeds the largest supported tensor (2^31 elements)");
    }
    AssignPrecision();
}

// Decides which Conv ops run on the tensor cores and which tensors are stored as
// bf16.  A tensor is bf16 iff a Conv produces it, every reader is a tensor-core
// Conv and it is not the model's input/output; a Conv runs on the tensor cores
// iff its shape is supported and its input tensor is bf16.  Fixpoint because
// the two conditions feed each other.  The first layer of a chain (fp32 input,
// e.g. 3 channels) stays on CUDA cores and writes bf16 directly.
void Model::AssignPrecision()
{
    for (auto& t : m_tensors) t.bf16 = t.split = false;
    std::vector<char> tc(m_ops.size(), 0);
    auto plain3x3 = [](Op const& op) {
        return op.KH == 3 && op.KW == 3 && op.stride == 1 && op.dilation == 1;
    };
    for (size_t i = 0; i < m_ops.size(); i++)
    {
        Op& op = m_ops[i];
        op.tc = false;
        op.pad_in = false;
        op.split = false;
        if (m_precision == 2 && op.kind == OpKind::Conv)
        {
            // fp32-grade on the tensor cores: tensors stay fp32, every supported layer splits its own input
            int const c = m_tensors[op.in[0]].C;
            int const cp = c < 16 ? 16 : c;
            if (c > 0 && nn::ConvTcSupported(op.KH, op.KW, op.stride, op.dilation, op.same_padding, cp,
                                             m_tensors[op.out].C, true))
                tc[i] = 1, op.split = true;
            continue;
        }
        if (m_precision != 1 || op.kind != OpKind::Conv) continue;
        int const cin = m_tensors[op.in[0]].C;
        // a thin first layer (e.g. RGB) runs on the tensor cores too, on a 16-channel zero-padded bf16 copy
        op.pad_in = cin > 0 && cin < 16;
        tc[i] = nn::ConvTcSupported(op.KH, op.KW, op.stride, op.dilation, op.same_padding,
                                    op.pad_in ? 16 : cin, m_tensors[op.out].C);
    }
    bool changed = m_precision == 1;
    std::vector<char> bf(m_tensors.size(), 0);
    while (changed)
    {
        changed = false;
        std::fill(bf.begin(), bf.end(), 0);
        for (size_t i = 0; i < m_ops.size(); i++)
        {
            Op const& op = m_ops[i];
            if (op.kind != OpKind::Conv || !(tc[i] || plain3x3(op))) continue;
            if (op.out == m_output_tensor || op.out == m_input_tensor) continue;
            if (m_tensors[op.out].C % 16 != 0) continue;
            if (op.residual >= 0) continue;          // conv + residual writes fp32
            bool all_tc = false;
            for (size_t j = 0; j < m_ops.size(); j++)
                for (int t : m_ops[j].in)
                    if (t == op.out) { all_tc = true; }
            for (size_t j = 0; j < m_ops.size() && all_tc; j++)
                for (int t : m_ops[j].in)
                    if (t == op.out && (!tc[j] || m_ops[j].residual == t)) all_tc = false;   // a residual is read as fp32
            bf[op.out] = all_tc;
        }
        for (size_t i = 0; i < m_ops.size(); i++)
            if (tc[i] && !m_ops[i].pad_in && !bf[m_ops[i].in[0]]) { tc[i] = 0; changed = true; }
    }
    for (size_t i = 0; i < m_ops.size(); i++) m_ops[i].tc = tc[i] != 0;
    for (size_t t = 0; t < m_tensors.size(); t++) m_tensors[t].bf16 = bf[t] != 0;
    // bf16x3: a tensor between two split layers is written by the producer's epilogue directly as
    // [hi | lo] bf16 halves (no fp32 round trip, no split pass) iff every reader is a split layer
    // taking it as its convolution input.
    for (Op const& op : m_ops)
    {
        if (!(op.tc && op.split) || op.residual >= 0) continue;
        if (op.out == m_output_tensor || op.out == m_input_tensor) continue;
        if (m_tensors[op.out].C % 16 != 0) continue;
        bool any = false, all = true;
        for (Op const& r : m_ops)
            for (size_t k = 0; k < r.in.size(); k++)
                if (r.in[k] == op.out)
                {
                    any = true;
                    if (!(r.tc && r.split && r.kind == OpKind::Conv && k == 0 && r.residual != op.out)) all = false;
                }
        m_tensors[op.out].split = any && all;
    }
}

size_t Model::TensorCoreLayers() const
{
    size_t n = 0;
    for (auto const& op : m_ops) n += op.tc ? 1 : 0;
    return n;
}

ml_status Model::SetPrecision(int precision)
{
    m_error.clear();
    if (precision != 0 && precision != 1 && precision != 2)
    {
        m_error = "Unsupported precision: " + std::to_string(precision);
        return ML_FAIL;
    }
    if (precision == m_precision) return ML_OK;
    m_precision = precision;
    try
    {
        AssignPrecision();
        if (m_ready) AllocateBuffers();
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        m_ready = false;
        m_error = std::string("Inference error: ") + e.what();
        return ML_FAIL;
    }
}

void Model::ReleaseBuffers()
{
    for (Op& op : m_ops)
    {
        nn::ConvTcDestroy(op.tc_op);
        op.tc_op = nullptr;
        cudaFree(op.pad_buffer);
        op.pad_buffer = nullptr;
    }
    for (void* p : m_buffers) cudaFree(p);
    m_buffers.clear();
    m_buffer_bytes.clear();
    cudaFree(m_split_buffer);
    m_split_buffer = nullptr;
    cudaFree(m_alias_in);
    m_alias_in = nullptr;
    cudaFree(m_in_f32);
    cudaFree(m_out_f32);
    m_in_f32 = m_out_f32 = nullptr;
}

// Liveness-based buffer reuse: an intermediate's buffer returns to the free
// list right after the last op that reads it.
void Model::AllocateBuffers()
{
    m_device->MakeCurrent();
    ReleaseBuffers();
    std::vector<int> free_list;
    for (auto& t : m_tensors) t.buffer = -1;
    for (size_t i = 0; i < m_ops.size(); i++)
    {
        Op const& op = m_ops[i];
        if (op.out != m_output_tensor)
        {
            TensorInfo& o = m_tensors[op.out];
            size_t need = static_cast<size_t>(o.H) * o.W * o.C * (o.bf16 ? 2 : sizeof(float));
            int best = -1;
            for (size_t k = 0; k < free_list.size(); k++)
            {
                int b = free_list[k];
                if (m_buffer_bytes[b] >= need && (best < 0 || m_buffer_bytes[b] < m_buffer_bytes[free_list[best]]))
                    best = static_cast<int>(k);
            }
            if (best >= 0)
            {
                o.buffer = free_list[best];
                free_list.erase(free_list.begin() + best);
            }
            else
            {
                void* p = nullptr;
                ML_CUDA_CHECK(cudaMalloc(&p, std::max<size_t>(need, 16)));
                m_buffers.push_back(p);
                m_buffer_bytes.push_back(need);
                o.buffer = static_cast<int>(m_buffers.size()) - 1;
            }
        }
        for (int t : op.in)
            if (m_tensors[t].last_use == static_cast<int>(i) && m_tensors[t].buffer >= 0)
            {
                bool dup = std::find(free_list.begin(), free_list.end(), m_tensors[t].buffer) != free_list.end();
                if (!dup) free_list.push_back(m_tensors[t].buffer);
            }
    }
    // Tensor-core layers: repack the filter once, bind the TMA tensor maps to the (now fixed) buffers.
    size_t split_bytes = 0;
    for (Op const& op : m_ops)
        if (op.tc && op.split && !m_tensors[op.in[0]].split)
        {
            TensorInfo const& a = m_tensors[op.in[0]];
            split_bytes = std::max(split_bytes, static_cast<size_t>(a.H) * a.W * 2 * (a.C < 16 ? 16 : a.C) * 2);
        }
    if (split_bytes > 0) ML_CUDA_CHECK(cudaMalloc(&m_split_buffer, split_bytes));
    for (Op& op : m_ops)
    {
        if (!op.tc) continue;
        TensorInfo const& a = m_tensors[op.in[0]];
        TensorInfo const& o = m_tensors[op.out];
        if (op.split)
        {
            int const cp = a.C < 16 ? 16 : a.C;
            if (op.packed_split < 0)
            {
                std::vector<uint16_t> packed = nn::ConvTcPackWeightsSplit(m_consts[op.weights].host.data(), a.C, o.C, cp);
                void* d = nullptr;
                ML_CUDA_CHECK(cudaMalloc(&d, packed.size() * 2));
                m_packed.push_back(d);
                op.packed_split = static_cast<int>(m_packed.size()) - 1;
                ML_CUDA_CHECK(cudaMemcpy(d, packed.data(), packed.size() * 2, cudaMemcpyHostToDevice));
            }
            op.tc_op = nn::ConvTcCreate(a.split ? m_buffers[a.buffer] : m_split_buffer, m_packed[op.packed_split],
                                        o.split ? m_buffers[o.buffer] : nullptr, a.H, a.W, cp, o.C, false, op.act, op.alpha,
                                        m_device->prop.multiProcessorCount, true, o.split);
            continue;
        }
        int const cin = op.pad_in ? 16 : a.C;
        if (op.pad_in)
            ML_CUDA_CHECK(cudaMalloc(&op.pad_buffer, static_cast<size_t>(a.H) * a.W * 16 * 2));
        if (op.packed < 0)
        {
            std::vector<uint16_t> packed = nn::ConvTcPackWeights(m_consts[op.weights].host.data(), a.C, o.C, cin);
            void* d = nullptr;
            ML_CUDA_CHECK(cudaMalloc(&d, packed.size() * 2));
            m_packed.push_back(d);
            op.packed = static_cast<int>(m_packed.size()) - 1;
            ML_CUDA_CHECK(cudaMemcpy(d, packed.data(), packed.size() * 2, cudaMemcpyHostToDevice));
        }
        op.tc_op = nn::ConvTcCreate(op.pad_in ? op.pad_buffer : m_buffers[a.buffer], m_packed[op.packed],
                                    o.bf16 ? m_buffers[o.buffer] : nullptr, a.H, a.W, cin, o.C, o.bf16,
                                    op.act, op.alpha, m_device->prop.multiProcessorCount);
    }
    if (m_input_info.dtype == ML_FLOAT16)
    {
        TensorInfo const& in = m_tensors[m_input_tensor];
        TensorInfo const& out = m_tensors[m_output_tensor];
        ML_CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&m_in_f32), static_cast<size_t>(in.H) * in.W * in.C * 4));
        ML_CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&m_out_f32), static_cast<size_t>(out.H) * out.W * out.C * 4));
    }
}

std::string Model::Describe() const
{
    static char const* const kinds[] = {"Conv", "BiasAct", "Add", "Concat", "Copy", "Pool", "Resize", "DepthToSpace", "Pad",
                                        "Binary"};
    static char const* const acts[] = {"none", "relu", "leaky_relu", "relu6", "sigmoid", "tanh"};
    std::ostringstream os;
    os << "input_node=" << m_input_node << " output_node=" << m_output_node << "\n";
    os << "input dtype=" << m_input_info.dtype << " w=" << m_input_info.width << " h=" << m_input_info.height
       << " c=" << m_input_info.channels << "\n";
    os << "output dtype=" << m_output_info.dtype << " w=" << m_output_info.width << " h=" << m_output_info.height
       << " c=" << m_output_info.channels << "\n";
    os << "ops=" << m_ops.size() << " launches=" << LaunchesPerInfer() << "\n";
    for (Op const& op : m_ops)
    {
        os << kinds[static_cast<int>(op.kind)] << " name=" << op.name << " in=";
        for (size_t k = 0; k < op.in.size(); k++) os << (k ? "," : "") << "t" << op.in[k];
        os << " out=t" << op.out << " cout=" << m_tensors[op.out].C;
        if (op.kind == OpKind::Conv)
        {
            double sum = 0;
            for (float v : m_consts[op.weights].host) sum += v;
            os << " k=" << op.KH << "x" << op.KW << " stride=" << op.stride << " dilation=" << op.dilation
               << " padding=" << (op.same_padding ? "SAME" : "VALID") << " wsum=" << sum;
        }
        if (op.kind == OpKind::Pool)
            os << (op.pool_max ? " max" : " avg") << " k=" << op.KH << " stride=" << op.stride << " padding="
               << (op.same_padding ? "SAME" : "VALID");
        if (op.kind == OpKind::Resize)
        {
            os << (op.bilinear ? " bilinear" : " nearest") << " align_corners=" << op.align_corners
               << " half_pixel_centers=" << op.half_pixel;
            if (op.size_ref >= 0) os << " size=t" << op.size_ref << ".hw*(" << op.mul_h << "," << op.mul_w << ")";
            else os << " size=" << op.size_h << "x" << op.size_w;
        }
        if (op.kind == OpKind::DepthToSpace) os << " block=" << op.block;
        if (op.kind == OpKind::Pad)
            os << " pads=" << op.pad_t << "," << op.pad_b << "," << op.pad_l << "," << op.pad_r << " mode="
               << (op.pad_mode == 0 ? "CONSTANT" : op.pad_mode == 1 ? "REFLECT" : "SYMMETRIC");
        if (op.kind == OpKind::Binary) os << (op.binary == 1 ? " sub" : " mul");
        if (op.scale >= 0) os << " scale=yes";
        if (op.residual >= 0) os << " residual=t" << op.residual;
        if (op.tc) os << (op.split ? " engine=tcgen05_bf16x3" : " engine=tcgen05_bf16");
        if (m_tensors[op.out].bf16) os << " out=bf16";
        if (m_tensors[op.out].split) os << " out=bf16_hi_lo";
        os << " bias=" << (op.bias >= 0 ? "yes" : "no") << " act=" << acts[static_cast<int>(op.act)];
        if (op.act == nn::Act::LeakyRelu) os << " alpha=" << op.alpha;
        os << "\n";
    }
    return os.str();
}

size_t Model::LaunchesPerInfer() const
{
    size_t n = 0;
    for (auto const& op : m_ops)
        n += op.kind == OpKind::Concat ? op.in.size()
                                       : (1 + (op.tc && op.pad_in ? 1 : 0) + (op.tc && op.residual >= 0 ? 1 : 0)
                                          + (op.tc && op.split && !m_tensors[op.in[0]].split ? 1 : 0));
    if (m_input_info.dtype == ML_FLOAT16) n += 2;
    return n;
}

// ---------------------------------------------------------------------------
// API methods
// ---------------------------------------------------------------------------
ml_status Model::GetInfo(ml_image_info* input_info, ml_image_info* output_info)
{
    if (!m_loaded) return ML_FAIL;
    if (input_info != nullptr) *input_info = m_input_info;
    if (output_info != nullptr) *output_info = m_output_info;
    return ML_OK;
}

ml_status Model::SetInputInfo(ml_image_info const* info)
{
    m_error.clear();
    if (info == nullptr)
    {
        m_error = "Bad info parameter";
        return ML_FAIL;
    }
    if (m_input_info.dtype != info->dtype)
    {
        std::ostringstream os;
        os << "Overriding data type " << m_input_info.dtype << " with " << info->dtype;
        m_error = os.str();
        return ML_FAIL;
    }
    bool ok = ForEachDim([&](auto dim, char const* name) {
        if (m_input_info.*dim != 0 && info->*dim != m_input_info.*dim)
        {
            std::ostringstream os;
            os << "Overriding " << name << " dimension " << m_input_info.*dim << " with " << info->*dim;
            m_error = os.str();
            return false;
        }
        return true;
    });
    if (!ok) return ML_FAIL;

    bool same = ForEachDim([&](auto dim, char const*) { return m_input_info.*dim == info->*dim; });
    if (same) return ML_OK; // nothing changed

    // The reference allocates the input tensor and runs a dummy inference on
    // zeros to learn the output shape (model.cpp:204-229); here the shape is
    // computed analytically and the device arena is (re)allocated.
    ml_image_info const previous = m_input_info;
    try
    {
        ok = ForEachDim([&](auto dim, char const* name) {
            if (info->*dim == 0)
            {
                m_error = std::string("Input image ") + name + " dimension is not specified";
                return false;
            }
            return true;
        });
        if (!ok) return ML_FAIL;
        m_input_info = *info;
        m_ready = false;
        InferShapes(info->height, info->width, info->channels);
        AllocateBuffers();
        TensorInfo const& out = m_tensors[m_output_tensor];
        m_output_info.height = static_cast<size_t>(out.H);
        m_output_info.width = static_cast<size_t>(out.W);
        m_output_info.channels = static_cast<size_t>(out.C);
        m_ready = true;
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        m_input_info = previous;
        m_error = std::string("Inference error: Invalid argument: ") + e.what();
        return ML_FAIL;
    }
}

void Model::Run(float const* d_in, float* d_out, cudaStream_t stream)
{
    // In-place inference (input image == output image): a tensor-core layer adds its folded residual
    // in a pass AFTER the convolution has written the output, so a residual that is the model input
    // must not share storage with the output.  (The fp32 kernels add it in their epilogue, where each
    // element is read before the same thread overwrites it.)
    if (static_cast<void const*>(d_in) == static_cast<void const*>(d_out))
    {
        // Likewise any op that writes the output while reading the input NON-elementwise (a convolution's main
        // input, a concat, pool, resize, depth-to-space or pad: neighbouring pixels or another layout): the
        // reference is safe here because TensorFlow copies the input into its own tensor (model.cpp:336).
        bool clash = false;
        for (Op const& op : m_ops)
        {
            if (op.out != m_output_tensor) continue;
            clash |= op.tc && op.residual == m_input_tensor;
            bool const elementwise = op.kind == OpKind::BiasAct || op.kind == OpKind::Add || op.kind == OpKind::Binary
                                     || op.kind == OpKind::Copy;
            for (size_t k = 0; k < op.in.size(); k++)
                if (op.in[k] == m_input_tensor && !elementwise
                    && !(op.kind == OpKind::Conv && static_cast<int>(op.in[k]) == op.residual && k > 0))
                    clash = true;
        }
        if (clash)
        {
            TensorInfo const& t = m_tensors[m_input_tensor];
            size_t const bytes = static_cast<size_t>(t.H) * t.W * t.C * sizeof(float);
            if (m_alias_in == nullptr) ML_CUDA_CHECK(cudaMalloc(&m_alias_in, std::max<size_t>(bytes, 16)));
            ML_CUDA_CHECK(cudaMemcpyAsync(m_alias_in, d_in, bytes, cudaMemcpyDeviceToDevice, stream));
            d_in = static_cast<float const*>(m_alias_in);
        }
    }
    for (Op const& op : m_ops) RunOp(op, d_in, d_out, stream);
}

void Model::RunOp(Op const& op, float const* d_in, float* d_out, cudaStream_t stream)
{
    auto ptr = [&](int t) -> float* {
        if (t == m_input_tensor) return const_cast<float*>(d_in);
        if (t == m_output_tensor) return d_out;
        return static_cast<float*>(m_buffers[m_tensors[t].buffer]); // bf16 tensors: same storage, 2-byte elements
    };
    {
        TensorInfo const& a = m_tensors[op.in[0]];
        TensorInfo const& o = m_tensors[op.out];
        size_t const n_out = static_cast<size_t>(o.H) * o.W * o.C;
        switch (op.kind)
        {
            case OpKind::Conv:
            {
                nn::ConvDesc d{};
                d.H = a.H; d.W = a.W; d.Cin = a.C;
                d.OH = o.H; d.OW = o.W; d.Cout = o.C;
                d.KH = op.KH; d.KW = op.KW;
                d.stride = op.stride; d.dilation = op.dilation;
                if (op.same_padding)
                {
                    // TensorFlow SAME: total = max((out-1)*stride + eff_k - in, 0), low half first
                    int effh = (op.KH - 1) * op.dilation + 1, effw = (op.KW - 1) * op.dilation + 1;
                    int ph = std::max((o.H - 1) * op.stride + effh - a.H, 0);
                    int pw = std::max((o.W - 1) * op.stride + effw - a.W, 0);
                    d.pad_top = ph / 2;
                    d.pad_left = pw / 2;
                }
                d.act = op.act;
                d.alpha = op.alpha;
                float const* bias = op.bias >= 0 ? m_consts[op.bias].dev : nullptr;
                d.residual = op.residual >= 0 ? ptr(op.residual) : nullptr;
                if (op.tc)
                {
                    if (op.split)
                    {
                        if (!a.split)
                            nn::SplitToBf16HiLo(ptr(op.in[0]), m_split_buffer, static_cast<size_t>(a.H) * a.W, a.C,
                                                a.C < 16 ? 16 : a.C, stream);
                    }
                    else if (op.pad_in)
                        nn::PadToBf16x16(ptr(op.in[0]), op.pad_buffer, static_cast<size_t>(a.H) * a.W, a.C, stream);
                    nn::ConvTcRun(op.tc_op, bias, ptr(op.out), stream);
                    // The folded Add runs as its own in-place pass here: inside the tensor-core epilogue the
                    // residual's global loads sit in the serial per-tile drain (measured on the 64->3 layer
                    // at 1080p: 0.135 ms fused, even prefetched, against 0.091 + 0.022 ms as two launches).
                    if (d.residual != nullptr)
                        nn::AddAct(ptr(op.out), d.residual, ptr(op.out), n_out, nn::Act::None, 0.f, stream);
                }
                else if (o.bf16)
                    nn::Conv2dFp32ToBf16(ptr(op.in[0]), m_consts[op.weights].dev, bias, ptr(op.out), d, stream);
                else
                    nn::Conv2dFp32(ptr(op.in[0]), m_consts[op.weights].dev, bias, ptr(op.out), d, stream);
                break;
            }
            case OpKind::BiasAct:
                if (op.scale >= 0)
                    nn::ScaleBiasAct(ptr(op.in[0]), m_consts[op.scale].dev, op.bias >= 0 ? m_consts[op.bias].dev : nullptr,
                                     ptr(op.out), n_out, o.C, op.act, op.alpha, stream);
                else
                    nn::BiasAct(ptr(op.in[0]), op.bias >= 0 ? m_consts[op.bias].dev : nullptr, ptr(op.out),
                                n_out, o.C, op.act, op.alpha, stream);
                break;
            case OpKind::Binary:
                nn::BinaryAct(ptr(op.in[0]), ptr(op.in[1]), ptr(op.out), n_out, op.binary, op.act, op.alpha, stream);
                break;
            case OpKind::Pool:
            {
                nn::PoolDesc d{};
                d.H = a.H; d.W = a.W; d.C = a.C; d.OH = o.H; d.OW = o.W; d.K = op.KH; d.stride = op.stride;
                d.is_max = op.pool_max ? 1 : 0;
                if (op.same_padding)      // TensorFlow SAME: total = max((out-1)*stride + k - in, 0), low half first
                {
                    d.pad_top = std::max((o.H - 1) * op.stride + op.KH - a.H, 0) / 2;
                    d.pad_left = std::max((o.W - 1) * op.stride + op.KH - a.W, 0) / 2;
                }
                nn::Pool2d(ptr(op.in[0]), ptr(op.out), d, stream);
                break;
            }
            case OpKind::Resize:
            {
                nn::ResizeDesc d{};
                d.H = a.H; d.W = a.W; d.C = a.C; d.OH = o.H; d.OW = o.W;
                d.bilinear = op.bilinear; d.align_corners = op.align_corners; d.half_pixel = op.half_pixel;
                auto scale = [&](int in, int out) {
                    return (op.align_corners && out > 1) ? static_cast<float>(in - 1) / static_cast<float>(out - 1)
                                                         : static_cast<float>(in) / static_cast<float>(out);
                };
                d.scale_y = scale(a.H, o.H); d.scale_x = scale(a.W, o.W);
                nn::Resize(ptr(op.in[0]), ptr(op.out), d, stream);
                break;
            }
            case OpKind::DepthToSpace:
                nn::DepthToSpace(ptr(op.in[0]), ptr(op.out), a.H, a.W, a.C, op.block, stream);
                break;
            case OpKind::Pad:
                nn::Pad2d(ptr(op.in[0]), ptr(op.out), a.H, a.W, a.C, op.pad_t, op.pad_l, o.H, o.W, op.pad_mode, stream);
                break;
            case OpKind::Add:
                nn::AddAct(ptr(op.in[0]), ptr(op.in[1]), ptr(op.out), n_out, op.act, op.alpha, stream);
                break;
            case OpKind::Concat:
            {
                int c0 = 0;
                for (int t : op.in)
                {
                    nn::ConcatChannels(ptr(t), ptr(op.out), static_cast<size_t>(o.H) * o.W,
                                       m_tensors[t].C, o.C, c0, stream);
                    c0 += m_tensors[t].C;
                }
                break;
            }
            case OpKind::Copy:
                ML_CUDA_CHECK(cudaMemcpyAsync(ptr(op.out), ptr(op.in[0]), n_out * sizeof(float),
                                              cudaMemcpyDeviceToDevice, stream));
                break;
        }
    }
}

bool Model::CheckInferArgs(ml_image input, ml_image output)
{
    Image* in = Image::FromHandle(input);
    Image* out = Image::FromHandle(output);
    if (in == nullptr)
    {
        m_error = "Bad input image handle";
        return false;
    }
    if (out == nullptr)
    {
        m_error = "Bad output image handle";
        return false;
    }

    ml_image_info const& oi = out->Info();
    bool ok = ForEachDim([&](auto dim, char const* name) {
        if (oi.*dim != m_output_info.*dim)
        {
            std::ostringstream os;
            os << "Output image " << name << " dimension " << oi.*dim << " does not match "
               << m_output_info.*dim;
            m_error = os.str();
            return false;
        }
        return true;
    });
    if (!ok) return false;

    ok = ForEachDim([&](auto dim, char const* name) {
        if (m_input_info.*dim == 0)
        {
            m_error = std::string("Input image ") + name + " dimension is not specified";
            return false;
        }
        return true;
    });
    if (!ok) return false;

    size_t const item = DataTypeSize(m_input_info.dtype);
    size_t const in_bytes = m_input_info.width * m_input_info.height * m_input_info.channels * item;
    size_t const out_bytes = m_output_info.width * m_output_info.height * m_output_info.channels * item;
    if (in->Bytes() != in_bytes)
    {
        std::ostringstream os;
        os << "Internal error: input size does not match: " << in->Bytes() << " vs " << in_bytes;
        m_error = os.str();
        return false;
    }
    if (out->Bytes() != out_bytes)
    {
        std::ostringstream os;
        os << "Internal error: output size does not match: " << out->Bytes() << " vs " << out_bytes;
        m_error = os.str();
        return false;
    }
    return true;
}

ml_status Model::Infer(ml_image input, ml_image output)
{
    m_error.clear();
    if (!CheckInferArgs(input, output)) return ML_FAIL;
    Image* in = Image::FromHandle(input);
    Image* out = Image::FromHandle(output);
    size_t const item = DataTypeSize(m_input_info.dtype);
    size_t const in_bytes = in->Bytes();
    size_t const out_bytes = out->Bytes();

    try
    {
        if (!m_ready) throw std::runtime_error("model input shape has not been set");
        if (in->Device()->device != m_device->device || out->Device()->device != m_device->device)
            throw std::runtime_error("images and model live on different devices");
        m_device->MakeCurrent();
        cudaStream_t s = m_device->stream;
        void* d_in = in->DeviceRead(s);
        void* d_out = out->DeviceWrite();
        if (m_input_info.dtype == ML_FLOAT16)
        {
            nn::HalfToFloat(d_in, m_in_f32, in_bytes / item, s);
            Run(m_in_f32, m_out_f32, s);
            nn::FloatToHalf(m_out_f32, d_out, out_bytes / item, s);
        }
        else
        {
            Run(static_cast<float const*>(d_in), static_cast<float*>(d_out), s);
        }
        out->CommitDeviceWrite(s);
        ML_CUDA_CHECK(cudaStreamSynchronize(s));
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        m_error = std::string("Inference error: ") + e.what();
        return ML_FAIL;
    }
}

ml_status Model::InferBatch(ml_image const* inputs, ml_image const* outputs, size_t n)
{
    m_error.clear();
    if ((inputs == nullptr || outputs == nullptr) && n != 0)
    {
        m_error = "Bad image array argument";
        return ML_FAIL;
    }
    for (size_t i = 0; i < n; i++)
        if (!CheckInferArgs(inputs[i], outputs[i]))
        {
            m_error = "Image " + std::to_string(i) + ": " + m_error;
            return ML_FAIL;
        }
    // An image that is one item's output and ANOTHER item's input (or output) cannot be pipelined: the upload of the
    // later item and the read-back of the earlier one would touch the same buffers concurrently.  Run in order.
    bool aliased = false;
    for (size_t i = 0; i < n && !aliased; i++)
        for (size_t j = 0; j < n && !aliased; j++)
            aliased = i != j && (outputs[i] == inputs[j] || outputs[i] == outputs[j]);
    if (m_input_info.dtype == ML_FLOAT16 || aliased)
    {
        // (also: the fp16 <-> fp32 staging buffers are single) run one after the other
        for (size_t i = 0; i < n; i++)
            if (Infer(inputs[i], outputs[i]) != ML_OK) return ML_FAIL;
        return ML_OK;
    }
    try
    {
        if (!m_ready) throw std::runtime_error("model input shape has not been set");
        m_device->MakeCurrent();
        cudaStream_t const s_in = m_device->copy_in, s_run = m_device->stream, s_out = m_device->copy_out;
        while (m_events.size() < 2 * n)
        {
            cudaEvent_t e = nullptr;
            ML_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            m_events.push_back(e);
        }
        if (m_batch_ev0 == nullptr)
        {
            ML_CUDA_CHECK(cudaEventCreate(&m_batch_ev0));
            ML_CUDA_CHECK(cudaEventCreate(&m_batch_ev1));
        }
        // device span of the batch (mlxModelLastBatchMs): first upload enqueued -> last read-back done.
        // s_run / s_out only start behind events of s_in, so the s_in record is the earliest point.
        ML_CUDA_CHECK(cudaEventRecord(m_batch_ev0, s_in));
        for (size_t i = 0; i < n; i++)
        {
            Image* in = Image::FromHandle(inputs[i]);
            Image* out = Image::FromHandle(outputs[i]);
            if (in->Device()->device != m_device->device || out->Device()->device != m_device->device)
                throw std::runtime_error("images and model live on different devices");
            // upload (if the host copy is the valid one) on the copy-in stream, ahead of the kernels
            void* d_in = in->DeviceRead(s_in);
            ML_CUDA_CHECK(cudaEventRecord(m_events[2 * i], s_in));
            ML_CUDA_CHECK(cudaStreamWaitEvent(s_run, m_events[2 * i], 0));
            void* d_out = out->DeviceWrite();
            Run(static_cast<float const*>(d_in), static_cast<float*>(d_out), s_run);
            ML_CUDA_CHECK(cudaEventRecord(m_events[2 * i + 1], s_run));
            // read-back (if the caller holds the host pointer) on the copy-out stream, behind the kernels
            ML_CUDA_CHECK(cudaStreamWaitEvent(s_out, m_events[2 * i + 1], 0));
            out->CommitDeviceWrite(s_out);
        }
        ML_CUDA_CHECK(cudaEventRecord(m_batch_ev1, s_out));   // s_out waits for the last image's kernels above
        ML_CUDA_CHECK(cudaStreamSynchronize(s_in));
        ML_CUDA_CHECK(cudaStreamSynchronize(s_run));
        ML_CUDA_CHECK(cudaStreamSynchronize(s_out));
        if (n > 0) ML_CUDA_CHECK(cudaEventElapsedTime(&m_last_batch_ms, m_batch_ev0, m_batch_ev1));
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        cudaStreamSynchronize(m_device->copy_in);
        cudaStreamSynchronize(m_device->stream);
        cudaStreamSynchronize(m_device->copy_out);
        m_error = std::string("Inference error: ") + e.what();
        return ML_FAIL;
    }
}

// One inference with every op bracketed by CUDA events on the compute stream (op by op, so the
// numbers are kernel durations, not a pipeline).  flops/bytes are the ALGORITHMIC figures of
// conv 2*OH*OW*KH*KW*Cin*Cout and (in + out) elements in their stored width + filter.
ml_status Model::Profile(ml_image input, ml_image output, float* op_ms, double* op_flops, double* op_bytes,
                         size_t max_ops, size_t* n_ops)
{
    m_error.clear();
    if (n_ops != nullptr) *n_ops = m_ops.size();
    if (!CheckInferArgs(input, output)) return ML_FAIL;
    try
    {
        if (!m_ready) throw std::runtime_error("model input shape has not been set");
        if (m_input_info.dtype != ML_FLOAT32) throw std::runtime_error("profiling needs ML_FLOAT32 images");
        Image* in = Image::FromHandle(input);
        Image* out = Image::FromHandle(output);
        m_device->MakeCurrent();
        cudaStream_t s = m_device->stream;
        float const* d_in = static_cast<float const*>(in->DeviceRead(s));
        float* d_out = static_cast<float*>(out->DeviceWrite());
        ML_CUDA_CHECK(cudaStreamSynchronize(s));
        for (size_t i = 0; i < m_ops.size(); i++)
        {
            Op const& op = m_ops[i];
            ML_CUDA_CHECK(cudaEventRecord(m_device->ev_start, s));
            RunOp(op, d_in, d_out, s);
            ML_CUDA_CHECK(cudaEventRecord(m_device->ev_stop, s));
            ML_CUDA_CHECK(cudaStreamSynchronize(s));
            float ms = 0.f;
            ML_CUDA_CHECK(cudaEventElapsedTime(&ms, m_device->ev_start, m_device->ev_stop));
            if (i >= max_ops) continue;
            TensorInfo const& o = m_tensors[op.out];
            double const out_elems = static_cast<double>(o.H) * o.W * o.C;
            double flops = out_elems, bytes = 0;
            for (int t : op.in)
                bytes += static_cast<double>(m_tensors[t].H) * m_tensors[t].W * m_tensors[t].C * (m_tensors[t].bf16 ? 2 : 4);
            bytes += out_elems * (o.bf16 ? 2 : 4);
            if (op.kind == OpKind::Conv)
            {
                TensorInfo const& a = m_tensors[op.in[0]];
                flops = 2.0 * out_elems * op.KH * op.KW * a.C;
                bytes += static_cast<double>(op.KH) * op.KW * a.C * o.C * (op.tc ? 2 : 4);
            }
            if (op_ms != nullptr) op_ms[i] = ms;
            if (op_flops != nullptr) op_flops[i] = flops;
            if (op_bytes != nullptr) op_bytes[i] = bytes;
        }
        out->CommitDeviceWrite(s);
        ML_CUDA_CHECK(cudaStreamSynchronize(s));
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        m_error = std::string("Inference error: ") + e.what();
        return ML_FAIL;
    }
}

char* Model::GetError(char* buffer, size_t buffer_size) const
{
    return FillBuffer(buffer, buffer_size, m_error);
}

} // namespace ML

extern "C" {

ML_API_ENTRY char* mlGetModelError(ml_model model, char* buffer, size_t buffer_size)
{
    if (ML::Model::FromHandle(model) == nullptr)
        return ML::FillBuffer(buffer, buffer_size, "Bad model handle");
    return ML::Model::FromHandle(model)->GetError(buffer, buffer_size);
}

ML_API_ENTRY ml_status mlGetModelInfo(ml_model model, ml_image_info* input_info, ml_image_info* output_info)
{
    if (ML::Model::FromHandle(model) == nullptr) return ML_FAIL;
    return ML::Model::FromHandle(model)->GetInfo(input_info, output_info);
}

ML_API_ENTRY ml_status mlSetModelInputInfo(ml_model model, ml_image_info const* info)
{
    if (ML::Model::FromHandle(model) == nullptr) return ML_FAIL;
    try
    {
        return ML::Model::FromHandle(model)->SetInputInfo(info);
    }
    catch (...)
    {
        return ML_FAIL;
    }
}

ML_API_ENTRY ml_status mlInfer(ml_model model, ml_image input, ml_image output)
{
    if (ML::Model::FromHandle(model) == nullptr) return ML_FAIL;
    try
    {
        return ML::Model::FromHandle(model)->Infer(input, output);
    }
    catch (...)
    {
        return ML_FAIL;
    }
}

ML_API_ENTRY ml_status mlxInferBatch(ml_model model, ml_image const* inputs, ml_image const* outputs, size_t count)
{
    if (ML::Model::FromHandle(model) == nullptr) return ML_FAIL;
    try
    {
        return ML::Model::FromHandle(model)->InferBatch(inputs, outputs, count);
    }
    catch (...)
    {
        return ML_FAIL;
    }
}

ML_API_ENTRY ml_status mlxProfileModel(ml_model model, ml_image input, ml_image output, float* op_ms,
                                       double* op_flops, double* op_bytes, size_t max_ops, size_t* n_ops)
{
    if (ML::Model::FromHandle(model) == nullptr) return ML_FAIL;
    try
    {
        return ML::Model::FromHandle(model)->Profile(input, output, op_ms, op_flops, op_bytes, max_ops, n_ops);
    }
    catch (...)
    {
        return ML_FAIL;
    }
}

ML_API_ENTRY void mlReleaseModel(ml_model model)
{
    delete ML::Model::FromHandle(model);
}

ML_API_ENTRY ml_status mlxInspectModel(char const* model_path, char const* input_node,
                                       char const* output_node, char* buffer, size_t buffer_size)
{
    try
    {
        ml_model_params params{model_path, input_node, output_node};
        ML::Model model(nullptr, &params);
        ML::FillBuffer(buffer, buffer_size, model.Describe());
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        ML::FillBuffer(buffer, buffer_size, e.what());
        return ML_FAIL;
    }
}

ML_API_ENTRY ml_status mlxSetModelPrecision(ml_model model, int precision)
{
    ML::Model* m = ML::Model::FromHandle(model);
    if (m == nullptr) return ML_FAIL;
    try { return m->SetPrecision(precision); }
    catch (...) { return ML_FAIL; }
}

ML_API_ENTRY size_t mlxModelTensorCoreLayers(ml_model model)
{
    ML::Model* m = ML::Model::FromHandle(model);
    return m != nullptr ? m->TensorCoreLayers() : 0;
}

ML_API_ENTRY float mlxModelLastBatchMs(ml_model model)
{
    if (ML::Model::FromHandle(model) == nullptr) return 0.f;
    return ML::Model::FromHandle(model)->LastBatchMs();
}

ML_API_ENTRY size_t mlxModelLaunchCount(ml_model model)
{
    if (ML::Model::FromHandle(model) == nullptr) return 0;
    return ML::Model::FromHandle(model)->LaunchesPerInfer();
}

} // extern "C"
