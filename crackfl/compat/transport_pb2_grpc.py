"""``transport_pb2_grpc`` as imported at fl_server.py:9 / fl_client.py:9."""
import grpc

from ..protocol import messages as pb

SERVICE_PATH = f"{pb.PACKAGE}.{pb.SERVICE}"


class TransportServiceServicer(object):
    def transport(self, request_iterator, context):
        context.set_code(grpc.StatusCode.UNIMPLEMENTED)
        raise NotImplementedError("Method not implemented!")


def add_TransportServiceServicer_to_server(servicer, server):
    handler = grpc.stream_stream_rpc_method_handler(
        servicer.transport, request_deserializer=pb.transportRequest.FromString,
        response_serializer=pb.transportResponse.SerializeToString)
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE_PATH, {pb.METHOD: handler}),))


class TransportServiceStub(object):
    def __init__(self, channel):
        self.transport = channel.stream_stream(f"/{SERVICE_PATH}/{pb.METHOD}",
                                               request_serializer=pb.transportRequest.SerializeToString,
                                               response_deserializer=pb.transportResponse.FromString)
