"""``transport_pb2`` as imported at fl_server.py:8 / fl_client.py:8 (runtime-built descriptors)."""
from ..protocol.messages import (FILE_DESCRIPTOR as DESCRIPTOR, ReadyRep, ReadyReq, Scalar, State, UpdateRep,  # noqa: F401
                                 UpdateReq, VersionRep, VersionReq, transportRequest, transportResponse)
