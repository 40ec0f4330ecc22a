// graphdef.cpp -- protobuf wire-format reader for frozen TensorFlow graphs.
// See graphdef.h.  Wire types: 0 varint, 1 fixed64, 2 length-delimited,
// 5 fixed32 (groups 3/4 are rejected).
#include "graphdef.h"

#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace ML {
namespace pb {
namespace {

class Reader
{
public:
    Reader(uint8_t const* p, size_t n) : m_p(p), m_end(p + n) {}
    bool AtEnd() const { return m_p >= m_end; }

    uint64_t Varint()
    {
        uint64_t v = 0;
        for (int shift = 0; shift < 64; shift += 7)
        {
            if (m_p >= m_end) throw std::runtime_error("truncated varint");
            uint8_t b = *m_p++;
            v |= static_cast<uint64_t>(b & 0x7F) << shift;
            if (!(b & 0x80)) return v;
        }
        throw std::runtime_error("varint too long");
    }

    void Tag(uint32_t& field, uint32_t& wire)
    {
        uint64_t t = Varint();
        field = static_cast<uint32_t>(t >> 3);
        wire = static_cast<uint32_t>(t & 7);
        if (field == 0) throw std::runtime_error("invalid field number 0");
    }

    Reader Sub()
    {
        uint64_t n = Varint();
        if (n > static_cast<uint64_t>(m_end - m_p)) throw std::runtime_error("truncated length-delimited field");
        Reader r(m_p, static_cast<size_t>(n));
        m_p += n;
        return r;
    }

    std::string Bytes()
    {
        Reader r = Sub();
        return std::string(reinterpret_cast<char const*>(r.m_p), static_cast<size_t>(r.m_end - r.m_p));
    }

    uint32_t Fixed32()
    {
        if (m_end - m_p < 4) throw std::runtime_error("truncated fixed32");
        uint32_t v;
        std::memcpy(&v, m_p, 4);
        m_p += 4;
        return v;
    }

    uint64_t Fixed64()
    {
        if (m_end - m_p < 8) throw std::runtime_error("truncated fixed64");
        uint64_t v;
        std::memcpy(&v, m_p, 8);
        m_p += 8;
        return v;
    }

    float Float()
    {
        uint32_t u = Fixed32();
        float f;
        std::memcpy(&f, &u, 4);
        return f;
    }

    void Skip(uint32_t wire)
    {
        switch (wire)
        {
            case 0: Varint(); break;
            case 1: Fixed64(); break;
            case 2: Sub(); break;
            case 5: Fixed32(); break;
            default: throw std::runtime_error("unsupported wire type " + std::to_string(wire));
        }
    }


private:
    uint8_t const* m_p;
    uint8_t const* m_end;
};

float HalfToFloat(uint16_t h)
{
    uint32_t sign = static_cast<uint32_t>(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1Fu;
    uint32_t man = h & 0x3FFu;
    uint32_t bits;
    if (exp == 0)
    {
        if (man == 0) bits = sign;
        else
        {
            int e = -1;
            do { man <<= 1; e++; } while (!(man & 0x400u));
            bits = sign | static_cast<uint32_t>(127 - 15 - e) << 23 | (man & 0x3FFu) << 13;
        }
    }
    else if (exp == 31) bits = sign | 0x7F800000u | man << 13;
    else bits = sign | (exp + 127 - 15) << 23 | man << 13;
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
}

// repeated scalar fields arrive either packed (wire 2) or one by one.
template <class Push>
void RepeatedVarint(Reader& r, uint32_t wire, Push push)
{
    if (wire == 2)
    {
        Reader s = r.Sub();
        while (!s.AtEnd()) push(s.Varint());
    }
    else if (wire == 0) push(r.Varint());
    else r.Skip(wire);
}

template <class Push>
void RepeatedFloat(Reader& r, uint32_t wire, Push push)
{
    if (wire == 2)
    {
        Reader s = r.Sub();
        while (!s.AtEnd()) push(s.Float());
    }
    else if (wire == 5) push(r.Float());
    else r.Skip(wire);
}

Shape ParseShape(Reader r)
{
    Shape shape;
    while (!r.AtEnd())
    {
        uint32_t f, w;
        r.Tag(f, w);
        if (f == 2 && w == 2) // Dim
        {
            Reader d = r.Sub();
            int64_t size = 0;
            while (!d.AtEnd())
            {
                uint32_t df, dw;
                d.Tag(df, dw);
                if (df == 1 && dw == 0) size = static_cast<int64_t>(d.Varint());
                else d.Skip(dw);
            }
            shape.dims.push_back(size);
        }
        else if (f == 3 && w == 0) shape.unknown_rank = r.Varint() != 0;
        else r.Skip(w);
    }
    return shape;
}

std::shared_ptr<Tensor> ParseTensor(Reader r)
{
    auto t = std::make_shared<Tensor>();
    std::string content;
    std::vector<float> float_val;
    std::vector<int64_t> int_val;
    std::vector<uint16_t> half_val;
    while (!r.AtEnd())
    {
        uint32_t f, w;
        r.Tag(f, w);
        if (f == 1 && w == 0) t->dtype = static_cast<int>(r.Varint());
        else if (f == 2 && w == 2) t->shape = ParseShape(r.Sub());
        else if (f == 4 && w == 2) content = r.Bytes();
        else if (f == 5) RepeatedFloat(r, w, [&](float v) { float_val.push_back(v); });
        else if (f == 7 || f == 10) RepeatedVarint(r, w, [&](uint64_t v) { int_val.push_back(static_cast<int64_t>(v)); });
        else if (f == 13) RepeatedVarint(r, w, [&](uint64_t v) { half_val.push_back(static_cast<uint16_t>(v)); });
        else r.Skip(w);
    }

    // The dims come from an untrusted file: multiply with an overflow check and cap the element count, so a crafted
    // shape whose product wraps cannot pass the size checks below with a tiny values vector.
    constexpr size_t kMaxConstElements = size_t(1) << 31;
    size_t count = 1;
    for (int64_t d : t->shape.dims)
    {
        if (d < 0) throw std::runtime_error("constant tensor with unknown dimension");
        if (d != 0 && count > kMaxConstElements / static_cast<size_t>(d))
            throw std::runtime_error("constant tensor too large");
        count *= static_cast<size_t>(d);
    }

    auto fill_from = [&](auto const& src, auto convert) {
        if (src.empty()) return false;
        t->values.resize(count);
        for (size_t i = 0; i < count; i++)
            t->values[i] = convert(src[i < src.size() ? i : src.size() - 1]); // TF repeats the last value
        return true;
    };

    switch (t->dtype)
    {
        case 1: // DT_FLOAT
            if (!content.empty())
            {
                if (content.size() != count * 4) throw std::runtime_error("tensor_content size mismatch");
                t->values.resize(count);
                std::memcpy(t->values.data(), content.data(), content.size());
            }
            else if (!fill_from(float_val, [](float v) { return v; })) t->values.assign(count, 0.f);
            break;
        case 19: // DT_HALF
            if (!content.empty())
            {
                if (content.size() != count * 2) throw std::runtime_error("tensor_content size mismatch");
                t->values.resize(count);
                for (size_t i = 0; i < count; i++)
                {
                    uint16_t h;
                    std::memcpy(&h, content.data() + 2 * i, 2);
                    t->values[i] = HalfToFloat(h);
                }
            }
            else if (!fill_from(half_val, [](uint16_t v) { return HalfToFloat(v); })) t->values.assign(count, 0.f);
            break;
        case 3: // DT_INT32
        case 9: // DT_INT64
        {
            size_t const item = t->dtype == 3 ? 4 : 8;
            t->ints.resize(count);
            if (!content.empty())
            {
                if (content.size() != count * item) throw std::runtime_error("tensor_content size mismatch");
                for (size_t i = 0; i < count; i++)
                {
                    if (item == 4) { int32_t v; std::memcpy(&v, content.data() + 4 * i, 4); t->ints[i] = v; }
                    else { int64_t v; std::memcpy(&v, content.data() + 8 * i, 8); t->ints[i] = v; }
                }
            }
            else
            {
                for (size_t i = 0; i < count; i++)
                    t->ints[i] = int_val.empty() ? 0 : int_val[i < int_val.size() ? i : int_val.size() - 1];
            }
            // int32 int_val entries are sign-extended varints
            if (t->dtype == 3)
                for (auto& v : t->ints) v = static_cast<int32_t>(v);
            t->values.resize(count);
            for (size_t i = 0; i < count; i++) t->values[i] = static_cast<float>(t->ints[i]);
            break;
        }
        default:
            // Other dtypes are carried without data; the model rejects them if used.
            break;
    }
    return t;
}

void ParseList(Reader r, Attr& a)
{
    while (!r.AtEnd())
    {
        uint32_t f, w;
        r.Tag(f, w);
        if (f == 2 && w == 2) a.list_s.push_back(r.Bytes());
        else if (f == 3) RepeatedVarint(r, w, [&](uint64_t v) { a.list_i.push_back(static_cast<int64_t>(v)); });
        else if (f == 4) RepeatedFloat(r, w, [&](float v) { a.list_f.push_back(v); });
        else if (f == 6) RepeatedVarint(r, w, [&](uint64_t v) { a.list_type.push_back(static_cast<int>(v)); });
        else if (f == 7 && w == 2) a.list_shape.push_back(ParseShape(r.Sub()));
        else r.Skip(w);
    }
}

Attr ParseAttr(Reader r)
{
    Attr a;
    while (!r.AtEnd())
    {
        uint32_t f, w;
        r.Tag(f, w);
        if (f == 1 && w == 2) ParseList(r.Sub(), a);
        else if (f == 2 && w == 2) { a.s = r.Bytes(); a.has_s = true; }
        else if (f == 3 && w == 0) { a.i = static_cast<int64_t>(r.Varint()); a.has_i = true; }
        else if (f == 4 && w == 5) { a.f = r.Float(); a.has_f = true; }
        else if (f == 5 && w == 0) { a.b = r.Varint() != 0; a.has_b = true; }
        else if (f == 6 && w == 0) { a.type = static_cast<int>(r.Varint()); a.has_type = true; }
        else if (f == 7 && w == 2) { a.shape = ParseShape(r.Sub()); a.has_shape = true; }
        else if (f == 8 && w == 2) a.tensor = ParseTensor(r.Sub());
        else r.Skip(w);
    }
    return a;
}

Node ParseNode(Reader r)
{
    Node n;
    while (!r.AtEnd())
    {
        uint32_t f, w;
        r.Tag(f, w);
        if (f == 1 && w == 2) n.name = r.Bytes();
        else if (f == 2 && w == 2) n.op = r.Bytes();
        else if (f == 3 && w == 2) n.inputs.push_back(r.Bytes());
        else if (f == 5 && w == 2) // map<string, AttrValue> entry
        {
            Reader e = r.Sub();
            std::string key;
            Attr value;
            while (!e.AtEnd())
            {
                uint32_t ef, ew;
                e.Tag(ef, ew);
                if (ef == 1 && ew == 2) key = e.Bytes();
                else if (ef == 2 && ew == 2) value = ParseAttr(e.Sub());
                else e.Skip(ew);
            }
            n.attr[key] = std::move(value);
        }
        else r.Skip(w);
    }
    return n;
}

} // namespace

size_t Tensor::ElementCount() const
{
    size_t c = 1;
    for (int64_t d : shape.dims) c *= static_cast<size_t>(d > 0 ? d : 0);
    return c;
}

Graph ParseGraphDef(uint8_t const* data, size_t size)
{
    Graph g;
    Reader r(data, size);
    while (!r.AtEnd())
    {
        uint32_t f, w;
        r.Tag(f, w);
        if (f == 1 && w == 2) g.nodes.push_back(ParseNode(r.Sub()));
        else r.Skip(w); // versions, library, ...
    }
    return g;
}

Graph ReadGraphDefFile(std::string const& path)
{
    std::ifstream file(path, std::ios::binary);
    if (!file)
        throw std::runtime_error("Error reading graph definition: " + path + ": Not found: " + path
                                 + "; No such file or directory");
    std::ostringstream ss;
    ss << file.rdbuf();
    std::string bytes = ss.str();
    try
    {
        Graph g = ParseGraphDef(reinterpret_cast<uint8_t const*>(bytes.data()), bytes.size());
        if (g.nodes.empty()) throw std::runtime_error("graph has no nodes");
        return g;
    }
    catch (std::exception const& e)
    {
        throw std::runtime_error("Error reading graph definition: " + path
                                 + ": Data loss: Can't parse " + path + " as binary proto (" + e.what() + ")");
    }
}

} // namespace pb
} // namespace ML
