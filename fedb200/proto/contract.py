"""Run-time construction of the ``federated`` protobuf contract.

The reference ships protoc output (reference: src/federated_pb2.py:17-107,
src/federated_pb2_grpc.py:8-92).  This image has no ``grpc_tools``, and generated code
is a poor thing to maintain anyway, so the contract is described ONCE as a small table
and turned into a ``FileDescriptorProto`` here.  The message classes live in a private
descriptor pool: importing this module next to the reference's ``federated_pb2`` (which
registers ``federated.proto`` in the default pool) can never collide.

The serialized descriptor is byte-identical to the reference's (tests/test_proto.py),
hence so is every message on the wire.
"""
from __future__ import annotations

import grpc
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

_F = descriptor_pb2.FieldDescriptorProto

PACKAGE = "federated"
SERVICE = "Trainer"

# message -> [(field, number, type)]; order is the reference's declaration order.
MESSAGES = (
    ("Request", ()),
    ("HeartBeatResponse", (("status", 1, _F.TYPE_INT32),)),
    ("TrainRequest", (("rank", 1, _F.TYPE_INT32), ("world", 2, _F.TYPE_INT32))),
    ("TrainReply", (("message", 1, _F.TYPE_STRING),)),
    ("SendModelRequest", (("model", 1, _F.TYPE_STRING),)),
    ("SendModelReply", (("reply", 1, _F.TYPE_STRING),)),
    ("PingRequest", (("req", 1, _F.TYPE_STRING),)),
    ("PingResponse", (("value", 1, _F.TYPE_INT32),)),
)

# rpc -> (request message, reply message)
METHODS = (
    ("StartTrain", "TrainRequest", "TrainReply"),
    ("SendModel", "SendModelRequest", "SendModelReply"),
    ("HeartBeat", "Request", "HeartBeatResponse"),
    ("CheckIfPrimaryUp", "PingRequest", "PingResponse"),
)

# 1 GiB both ways, as the reference (src/server.py:42-45, src/client.py:40-47).
MAX_MESSAGE_BYTES = 1024 * 1024 * 1024
CHANNEL_OPTIONS = (
    ("grpc.max_send_message_length", MAX_MESSAGE_BYTES),
    ("grpc.max_receive_message_length", MAX_MESSAGE_BYTES),
)


def file_descriptor_proto() -> descriptor_pb2.FileDescriptorProto:
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "federated.proto"
    fd.package = PACKAGE
    fd.syntax = "proto3"
    for name, fields in MESSAGES:
        m = fd.message_type.add()
        m.name = name
        for fname, number, ftype in fields:
            f = m.field.add()
            f.name, f.number, f.type = fname, number, ftype
            f.label = _F.LABEL_OPTIONAL
    svc = fd.service.add()
    svc.name = SERVICE
    for rpc, req, rep in METHODS:
        meth = svc.method.add()
        meth.name = rpc
        meth.input_type = f".{PACKAGE}.{req}"
        meth.output_type = f".{PACKAGE}.{rep}"
        meth.options.SetInParent()  # `{}` after each rpc produces an empty options msg
    fd.options.java_package = "io.grpc.examples.federated"
    fd.options.java_outer_classname = "FederatedProto"
    fd.options.java_multiple_files = True
    fd.options.objc_class_prefix = "HLW"
    return fd


_POOL = descriptor_pool.DescriptorPool()
DESCRIPTOR = _POOL.Add(file_descriptor_proto()) or _POOL.FindFileByName("federated.proto")
if DESCRIPTOR is None:  # protobuf >= 4 returns None from Add
    DESCRIPTOR = _POOL.FindFileByName("federated.proto")


def _cls(name: str):
    return message_factory.GetMessageClass(_POOL.FindMessageTypeByName(f"{PACKAGE}.{name}"))


Request = _cls("Request")
HeartBeatResponse = _cls("HeartBeatResponse")
TrainRequest = _cls("TrainRequest")
TrainReply = _cls("TrainReply")
SendModelRequest = _cls("SendModelRequest")
SendModelReply = _cls("SendModelReply")
PingRequest = _cls("PingRequest")
PingResponse = _cls("PingResponse")

_BY_NAME = {n: _cls(n) for n, _ in MESSAGES}


def method_path(rpc: str) -> str:
    return f"/{PACKAGE}.{SERVICE}/{rpc}"


class TrainerStub:
    """Client side of ``federated.Trainer``: one unary-unary callable per rpc."""

    def __init__(self, channel: grpc.Channel):
        for rpc, req, rep in METHODS:
            setattr(self, rpc, channel.unary_unary(
                method_path(rpc),
                request_serializer=_BY_NAME[req].SerializeToString,
                response_deserializer=_BY_NAME[rep].FromString))


class TrainerServicer:
    """Server side base class; unimplemented rpcs answer UNIMPLEMENTED like protoc's."""

    def _unimplemented(self, request, context):
        context.set_code(grpc.StatusCode.UNIMPLEMENTED)
        context.set_details("Method not implemented!")
        raise NotImplementedError("Method not implemented!")

    StartTrain = SendModel = HeartBeat = CheckIfPrimaryUp = _unimplemented


def add_TrainerServicer_to_server(servicer, server: grpc.Server) -> None:
    handlers = {
        rpc: grpc.unary_unary_rpc_method_handler(
            getattr(servicer, rpc),
            request_deserializer=_BY_NAME[req].FromString,
            response_serializer=_BY_NAME[rep].SerializeToString)
        for rpc, req, rep in METHODS
    }
    server.add_generic_rpc_handlers(
        (grpc.method_handlers_generic_handler(f"{PACKAGE}.{SERVICE}", handlers),))


def make_channel(target: str, compress: bool = False) -> grpc.Channel:
    """Insecure channel with the 1 GiB caps; gzip iff ``compress`` (reference -c Y,
    src/server.py:103-107)."""
    kw = {"compression": grpc.Compression.Gzip} if compress else {}
    return grpc.insecure_channel(target, options=list(CHANNEL_OPTIONS), **kw)


def make_server(max_workers: int = 10, compress: bool = False) -> grpc.Server:
    from concurrent import futures
    kw = {"compression": grpc.Compression.Gzip} if compress else {}
    return grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers),
                       options=list(CHANNEL_OPTIONS), **kw)
