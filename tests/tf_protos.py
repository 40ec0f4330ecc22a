"""Builds the TensorFlow GraphDef message classes dynamically with the protobuf
runtime (no protoc, no TensorFlow): an INDEPENDENT encoder/decoder used to
cross-check simpleraytracer_b200/graphdef_writer.py and the C++ reader.
Field numbers are those of TensorFlow's public .proto files."""
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

F = descriptor_pb2.FieldDescriptorProto


def _field(msg, name, number, ftype, label=F.LABEL_OPTIONAL, type_name=None, packed=None):
    f = msg.field.add()
    f.name, f.number, f.type, f.label = name, number, ftype, label
    if type_name:
        f.type_name = type_name
    if packed is not None:
        f.options.packed = packed
    return f


def build():
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "tf_subset.proto"
    fd.package = "tensorflow"
    fd.syntax = "proto3"

    shape = fd.message_type.add(); shape.name = "TensorShapeProto"
    dim = shape.nested_type.add(); dim.name = "Dim"
    _field(dim, "size", 1, F.TYPE_INT64)
    _field(dim, "name", 2, F.TYPE_STRING)
    _field(shape, "dim", 2, F.TYPE_MESSAGE, F.LABEL_REPEATED, ".tensorflow.TensorShapeProto.Dim")
    _field(shape, "unknown_rank", 3, F.TYPE_BOOL)

    tensor = fd.message_type.add(); tensor.name = "TensorProto"
    _field(tensor, "dtype", 1, F.TYPE_INT32)
    _field(tensor, "tensor_shape", 2, F.TYPE_MESSAGE, type_name=".tensorflow.TensorShapeProto")
    _field(tensor, "version_number", 3, F.TYPE_INT32)
    _field(tensor, "tensor_content", 4, F.TYPE_BYTES)
    _field(tensor, "float_val", 5, F.TYPE_FLOAT, F.LABEL_REPEATED, packed=True)
    _field(tensor, "int_val", 7, F.TYPE_INT32, F.LABEL_REPEATED, packed=True)
    _field(tensor, "half_val", 13, F.TYPE_INT32, F.LABEL_REPEATED, packed=True)

    attr = fd.message_type.add(); attr.name = "AttrValue"
    lst = attr.nested_type.add(); lst.name = "ListValue"
    _field(lst, "s", 2, F.TYPE_BYTES, F.LABEL_REPEATED)
    _field(lst, "i", 3, F.TYPE_INT64, F.LABEL_REPEATED, packed=True)
    _field(lst, "f", 4, F.TYPE_FLOAT, F.LABEL_REPEATED, packed=True)
    _field(lst, "b", 5, F.TYPE_BOOL, F.LABEL_REPEATED, packed=True)
    _field(lst, "type", 6, F.TYPE_INT32, F.LABEL_REPEATED, packed=True)
    _field(lst, "shape", 7, F.TYPE_MESSAGE, F.LABEL_REPEATED, ".tensorflow.TensorShapeProto")
    attr.oneof_decl.add().name = "value"
    for name, num, t, tn in [("list", 1, F.TYPE_MESSAGE, ".tensorflow.AttrValue.ListValue"),
                             ("s", 2, F.TYPE_BYTES, None), ("i", 3, F.TYPE_INT64, None),
                             ("f", 4, F.TYPE_FLOAT, None), ("b", 5, F.TYPE_BOOL, None),
                             ("type", 6, F.TYPE_INT32, None),
                             ("shape", 7, F.TYPE_MESSAGE, ".tensorflow.TensorShapeProto"),
                             ("tensor", 8, F.TYPE_MESSAGE, ".tensorflow.TensorProto")]:
        f = _field(attr, name, num, t, type_name=tn)
        f.oneof_index = 0

    node = fd.message_type.add(); node.name = "NodeDef"
    entry = node.nested_type.add(); entry.name = "AttrEntry"
    entry.options.map_entry = True
    _field(entry, "key", 1, F.TYPE_STRING)
    _field(entry, "value", 2, F.TYPE_MESSAGE, type_name=".tensorflow.AttrValue")
    _field(node, "name", 1, F.TYPE_STRING)
    _field(node, "op", 2, F.TYPE_STRING)
    _field(node, "input", 3, F.TYPE_STRING, F.LABEL_REPEATED)
    _field(node, "device", 4, F.TYPE_STRING)
    _field(node, "attr", 5, F.TYPE_MESSAGE, F.LABEL_REPEATED, ".tensorflow.NodeDef.AttrEntry")

    versions = fd.message_type.add(); versions.name = "VersionDef"
    _field(versions, "producer", 1, F.TYPE_INT32)

    graph = fd.message_type.add(); graph.name = "GraphDef"
    _field(graph, "node", 1, F.TYPE_MESSAGE, F.LABEL_REPEATED, ".tensorflow.NodeDef")
    _field(graph, "version", 3, F.TYPE_INT32)
    _field(graph, "versions", 4, F.TYPE_MESSAGE, type_name=".tensorflow.VersionDef")

    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = getattr(message_factory, "GetMessageClass", None)
    if get is None:
        factory = message_factory.MessageFactory(pool)
        get = factory.GetPrototype
    return {n: get(pool.FindMessageTypeByName("tensorflow." + n))
            for n in ["GraphDef", "NodeDef", "AttrValue", "TensorProto", "TensorShapeProto"]}
