"""Wire schema of the reference's ``node.proto``, built programmatically.

``grpc_tools``/``protoc`` are not available offline, so instead of committing
generated stubs the file descriptor is assembled with ``descriptor_pb2`` at
import time.  Field numbers, types, labels, the proto3-optional synthetic
oneofs, the package (``node``) and the service/method names reproduce
p2pfl/communication/grpc/proto/node.proto:26-57, which makes the bytes on the
wire interchangeable with a reference node.

A private ``DescriptorPool`` is used so this module can live in the same
process as the reference's own ``node_pb2`` (both describe ``node.proto``).
"""

from __future__ import annotations

from google.protobuf import descriptor_pb2, descriptor_pool, empty_pb2, message_factory

_F = descriptor_pb2.FieldDescriptorProto

SERVICE = "node.NodeServices"


def _field(msg, name, number, ftype, label=_F.LABEL_OPTIONAL, proto3_optional=False):
    f = msg.field.add()
    f.name, f.number, f.type, f.label = name, number, ftype, label
    f.json_name = name
    if proto3_optional:
        od = msg.oneof_decl.add()
        od.name = "_" + name
        f.oneof_index = len(msg.oneof_decl) - 1
        f.proto3_optional = True
    return f


def _build_file() -> descriptor_pb2.FileDescriptorProto:
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "node.proto"
    fd.package = "node"
    fd.syntax = "proto3"
    fd.dependency.append("google/protobuf/empty.proto")

    m = fd.message_type.add()
    m.name = "Message"
    _field(m, "source", 1, _F.TYPE_STRING)
    _field(m, "ttl", 2, _F.TYPE_INT32)
    _field(m, "hash", 3, _F.TYPE_INT64)
    _field(m, "cmd", 4, _F.TYPE_STRING)
    _field(m, "args", 5, _F.TYPE_STRING, _F.LABEL_REPEATED)
    _field(m, "round", 6, _F.TYPE_INT32, proto3_optional=True)

    w = fd.message_type.add()
    w.name = "Weights"
    _field(w, "source", 1, _F.TYPE_STRING)
    _field(w, "round", 2, _F.TYPE_INT32)
    _field(w, "weights", 3, _F.TYPE_BYTES)
    _field(w, "contributors", 4, _F.TYPE_STRING, _F.LABEL_REPEATED)
    _field(w, "weight", 5, _F.TYPE_INT32)
    _field(w, "cmd", 6, _F.TYPE_STRING)

    h = fd.message_type.add()
    h.name = "HandShakeRequest"
    _field(h, "addr", 1, _F.TYPE_STRING)

    r = fd.message_type.add()
    r.name = "ResponseMessage"
    _field(r, "error", 1, _F.TYPE_STRING, proto3_optional=True)

    svc = fd.service.add()
    svc.name = "NodeServices"
    for name, req, resp in (
        ("handshake", ".node.HandShakeRequest", ".node.ResponseMessage"),
        ("disconnect", ".node.HandShakeRequest", ".google.protobuf.Empty"),
        ("send_message", ".node.Message", ".node.ResponseMessage"),
        ("send_weights", ".node.Weights", ".node.ResponseMessage"),
    ):
        meth = svc.method.add()
        meth.name, meth.input_type, meth.output_type = name, req, resp
    return fd


_pool = descriptor_pool.DescriptorPool()
_pool.AddSerializedFile(empty_pb2.DESCRIPTOR.serialized_pb)
_pool.Add(_build_file())


def _cls(name: str):
    return message_factory.GetMessageClass(_pool.FindMessageTypeByName(name))


PbMessage = _cls("node.Message")
PbWeights = _cls("node.Weights")
PbHandShakeRequest = _cls("node.HandShakeRequest")
PbResponseMessage = _cls("node.ResponseMessage")
PbEmpty = empty_pb2.Empty

# method -> (request class, response class)
METHODS = {
    "handshake": (PbHandShakeRequest, PbResponseMessage),
    "disconnect": (PbHandShakeRequest, PbEmpty),
    "send_message": (PbMessage, PbResponseMessage),
    "send_weights": (PbWeights, PbResponseMessage),
}


def method_path(name: str) -> str:
    return f"/{SERVICE}/{name}"
