"""RPC contract (SURVEY work item 1): descriptor equality with the reference and a loop-back call."""
import os
import sys

import grpc
import pytest

from fedb200 import proto
from conftest import reference_src


def test_descriptor_matches_proto_file():
    fd = proto.file_descriptor_proto()
    text = open(os.path.join(os.path.dirname(proto.__file__), "federated.proto")).read()
    assert f"package {fd.package};" in text
    for m in fd.message_type:
        assert f"message {m.name}" in text
        for f in m.field:
            assert f"{f.name} = {f.number};" in text.replace("  ", " ")
    for meth in fd.service[0].method:
        assert f"rpc {meth.name}" in text


def test_descriptor_bytes_equal_reference():
    src = reference_src()
    if src is None:
        pytest.skip("reference not available")
    sys.path.insert(0, src)
    try:
        import federated_pb2 as ref
    finally:
        sys.path.remove(src)
    assert proto.file_descriptor_proto().SerializeToString() == ref.DESCRIPTOR.serialized_pb
    # wire compatibility both ways
    assert ref.TrainRequest.FromString(proto.TrainRequest(rank=3, world=8).SerializeToString()).world == 8
    assert proto.PingRequest.FromString(ref.PingRequest(req="1").SerializeToString()).req == "1"


def test_method_paths():
    assert proto.method_path("StartTrain") == "/federated.Trainer/StartTrain"
    assert [m[0] for m in proto.contract.METHODS] == ["StartTrain", "SendModel", "HeartBeat", "CheckIfPrimaryUp"]


def test_loopback_and_unimplemented():
    class S(proto.TrainerServicer):
        def HeartBeat(self, request, context):
            return proto.HeartBeatResponse(status=1)

    server = proto.make_server(2)
    proto.add_TrainerServicer_to_server(S(), server)
    port = server.add_insecure_port("127.0.0.1:0")
    server.start()
    try:
        ch = proto.make_channel(f"127.0.0.1:{port}", compress=True)
        stub = proto.TrainerStub(ch)
        assert stub.HeartBeat(proto.Request(), timeout=5).status == 1
        with pytest.raises(grpc.RpcError) as e:
            stub.CheckIfPrimaryUp(proto.PingRequest(req="0"), timeout=5)
        assert e.value.code() == grpc.StatusCode.UNIMPLEMENTED
        ch.close()
    finally:
        server.stop(0)
