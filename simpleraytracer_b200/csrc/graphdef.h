// graphdef.h -- hand-written reader for the subset of TensorFlow's frozen
// GraphDef (binary protobuf) that the model runner needs.
// Replaces tensorflow::ReadBinaryProto at reference model.cpp:87; no protobuf
// or TensorFlow library is involved.  Field numbers follow TensorFlow's public
// .proto files (graph.proto, node_def.proto, attr_value.proto, tensor.proto,
// tensor_shape.proto, types.proto).
#pragma once

#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace ML {
namespace pb {

struct Shape
{
    std::vector<int64_t> dims; // -1 = unknown
    bool unknown_rank = false;
};

struct Tensor
{
    int dtype = 0;             // types.proto DataType
    Shape shape;
    std::vector<float> values; // decoded to fp32 (DT_FLOAT, DT_HALF, DT_INT32, DT_INT64)
    std::vector<int64_t> ints; // same data for integer tensors (axis / size constants)
    size_t ElementCount() const;
};

struct Attr
{
    bool has_s = false, has_i = false, has_f = false, has_b = false, has_type = false,
         has_shape = false;
    std::string s;
    int64_t i = 0;
    float f = 0.f;
    bool b = false;
    int type = 0;
    Shape shape;
    std::shared_ptr<Tensor> tensor;
    // list(...)
    std::vector<int64_t> list_i;
    std::vector<float> list_f;
    std::vector<std::string> list_s;
    std::vector<Shape> list_shape;
    std::vector<int> list_type;
};

struct Node
{
    std::string name;
    std::string op;
    std::vector<std::string> inputs;
    std::map<std::string, Attr> attr;

    Attr const* Find(char const* key) const
    {
        auto it = attr.find(key);
        return it == attr.end() ? nullptr : &it->second;
    }
};

struct Graph
{
    std::vector<Node> nodes;
};

// Throws std::runtime_error on malformed input.
Graph ParseGraphDef(uint8_t const* data, size_t size);
// Reads the file and parses it; error text mirrors the reference's
// "Error reading graph definition: <path>: <status>" (model.cpp:90).
Graph ReadGraphDefFile(std::string const& path);

} // namespace pb
} // namespace ML
