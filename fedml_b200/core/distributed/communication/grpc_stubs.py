"""The generated-code surface of the reference's gRPC service, written by hand on grpc's generic API.

``grpc_comm_manager.proto`` (reference core/distributed/communication/grpc/proto/grpc_comm_manager.proto:3-16) declares

    service gRPCCommManager { rpc sendMessage(CommRequest) returns (CommResponse);
                              rpc handleReceiveMessage(CommRequest) returns (CommResponse); }

and protoc turns it into ``grpc_comm_manager_pb2_grpc.py`` with a ``gRPCCommManagerStub``, a ``gRPCCommManagerServicer`` base class
and ``add_gRPCCommManagerServicer_to_server``.  protoc's output needs the protobuf runtime of the matching generation; the classes
below give the same names and behaviour with the wire codec of ``grpc_backend`` (``CommRequest`` / ``CommResponse`` encode exactly
the two proto fields), so a program written against the generated module - its own servicer subclass, its own stub calls - runs
unchanged and interoperates with a reference peer.
"""
from __future__ import annotations

import grpc

from .grpc_backend import CommRequest, CommResponse

SERVICE_NAME = "gRPCCommManager"
_METHODS = ("sendMessage", "handleReceiveMessage")


class _Descriptor:
    """Minimal stand-in for the protobuf file descriptor: the names a caller can reasonably ask for."""
    name = "grpc_comm_manager.proto"
    package = ""
    services_by_name = {SERVICE_NAME: _METHODS}
    message_types_by_name = {"CommRequest": CommRequest, "CommResponse": CommResponse}


DESCRIPTOR = _Descriptor()


class gRPCCommManagerStub:  # noqa: N801
    """Client side: ``stub.sendMessage(CommRequest(client_id=..., message=...)) -> CommResponse``."""

    def __init__(self, channel: grpc.Channel):
        for name in _METHODS:
            setattr(self, name, channel.unary_unary(f"/{SERVICE_NAME}/{name}", request_serializer=CommRequest.SerializeToString,
                                                    response_deserializer=CommResponse.FromString))


class gRPCCommManagerServicer:  # noqa: N801
    """Server side base class: override the two methods."""

    def sendMessage(self, request, context):  # noqa: N802
        context.set_code(grpc.StatusCode.UNIMPLEMENTED)
        context.set_details("Method not implemented!")
        raise NotImplementedError("Method not implemented!")

    def handleReceiveMessage(self, request, context):  # noqa: N802
        context.set_code(grpc.StatusCode.UNIMPLEMENTED)
        context.set_details("Method not implemented!")
        raise NotImplementedError("Method not implemented!")


def add_gRPCCommManagerServicer_to_server(servicer, server):  # noqa: N802
    handlers = {name: grpc.unary_unary_rpc_method_handler(getattr(servicer, name), request_deserializer=CommRequest.FromString,
                                                          response_serializer=CommResponse.SerializeToString) for name in _METHODS}
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE_NAME, handlers),))


class gRPCCommManager:  # noqa: N801
    """One-shot calls without keeping a stub (protoc's "experimental API" class): opens an insecure channel per call unless
    ``channel_credentials`` are given."""

    @staticmethod
    def _call(method, request, target, options=(), channel_credentials=None, insecure=False, compression=None, wait_for_ready=None,
              timeout=None, metadata=None):
        ch = grpc.insecure_channel(target, options=options, compression=compression) if (insecure or channel_credentials is None) \
            else grpc.secure_channel(target, channel_credentials, options=options, compression=compression)
        try:
            return getattr(gRPCCommManagerStub(ch), method)(request, timeout=timeout, metadata=metadata, wait_for_ready=wait_for_ready)
        finally:
            ch.close()

    @staticmethod
    def sendMessage(request, target, options=(), channel_credentials=None, call_credentials=None, insecure=False, compression=None,  # noqa: N802
                    wait_for_ready=None, timeout=None, metadata=None):
        return gRPCCommManager._call("sendMessage", request, target, options, channel_credentials, insecure, compression, wait_for_ready,
                                     timeout, metadata)

    @staticmethod
    def handleReceiveMessage(request, target, options=(), channel_credentials=None, call_credentials=None, insecure=False,  # noqa: N802
                             compression=None, wait_for_ready=None, timeout=None, metadata=None):
        return gRPCCommManager._call("handleReceiveMessage", request, target, options, channel_credentials, insecure, compression,
                                     wait_for_ready, timeout, metadata)
