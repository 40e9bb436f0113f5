"""Runtime construction of the ``transport`` protobuf messages (no ``protoc`` on the image).

Builds the descriptors of ``transport.proto`` (same directory) with
``descriptor_pb2`` and exposes the generated classes under the names the reference
imports from its missing ``transport_pb2`` (``fl_server.py:8``, ``fl_client.py:8``):
``Scalar, transportRequest, transportResponse, ReadyReq, ReadyRep, UpdateReq, UpdateRep,
VersionReq, VersionRep, State``.
"""
from __future__ import annotations

from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
from google.protobuf.internal import enum_type_wrapper

PACKAGE = "transport"
SERVICE = "TransportService"
METHOD = "transport"
FILE_NAME = "crackfl/transport.proto"

_F = descriptor_pb2.FieldDescriptorProto
_T = {"string": _F.TYPE_STRING, "bytes": _F.TYPE_BYTES, "int32": _F.TYPE_INT32, "int64": _F.TYPE_INT64,
      "float": _F.TYPE_FLOAT}

STATE_LABELS = ("ON", "TRAINING", "TRAIN_DONE", "WAIT", "NOT_WAIT", "FIN")

# (message, [(field, type, number)])  type: scalar name | "State" | "map" | message name
SCHEMA = [
    ("ReadyReq", [("type", "string", 1), ("cname", "string", 2), ("state", "State", 3), ("config", "map", 4)]),
    ("UpdateReq", [("type", "string", 1), ("buffer_chunk", "bytes", 2), ("title", "string", 3),
                   ("file_len", "int64", 4), ("state", "State", 5), ("cname", "string", 6),
                   ("current_round", "int32", 7), ("config", "map", 8)]),
    ("VersionReq", [("type", "string", 1), ("config", "map", 2)]),
    ("ReadyRep", [("config", "map", 1)]),
    ("UpdateRep", [("type", "string", 1), ("buffer_chunk", "bytes", 2), ("title", "string", 3), ("config", "map", 4)]),
    ("VersionRep", [("state", "State", 1), ("buffer_chunk", "bytes", 2), ("config", "map", 3)]),
]
ONEOF_WRAPPERS = [
    ("transportRequest", "request", [("ready_req", "ReadyReq", 1), ("update_req", "UpdateReq", 2),
                                     ("version_req", "VersionReq", 3)]),
    ("transportResponse", "response", [("ready_rep", "ReadyRep", 1), ("update_rep", "UpdateRep", 2),
                                       ("version_rep", "VersionRep", 3)]),
]


def _build_file() -> descriptor_pb2.FileDescriptorProto:
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = FILE_NAME
    fd.package = PACKAGE
    fd.syntax = "proto3"

    en = fd.enum_type.add()
    en.name = "State"
    for i, lab in enumerate(STATE_LABELS):
        v = en.value.add()
        v.name, v.number = lab, i

    sc = fd.message_type.add()
    sc.name = "Scalar"
    sc.oneof_decl.add().name = "scalar"
    for name, typ, num in (("scint32", "int32", 1), ("scstring", "string", 2), ("scfloat", "float", 3)):
        f = sc.field.add()
        f.name, f.number, f.type, f.label, f.oneof_index = name, num, _T[typ], _F.LABEL_OPTIONAL, 0

    for mname, fields in SCHEMA:
        m = fd.message_type.add()
        m.name = mname
        for name, typ, num in fields:
            f = m.field.add()
            f.name, f.number = name, num
            if typ == "map":
                entry = m.nested_type.add()
                entry.name = "".join(p.capitalize() for p in name.split("_")) + "Entry"
                entry.options.map_entry = True
                k = entry.field.add()
                k.name, k.number, k.type, k.label = "key", 1, _F.TYPE_STRING, _F.LABEL_OPTIONAL
                v = entry.field.add()
                v.name, v.number, v.type, v.label = "value", 2, _F.TYPE_MESSAGE, _F.LABEL_OPTIONAL
                v.type_name = f".{PACKAGE}.Scalar"
                f.type, f.label = _F.TYPE_MESSAGE, _F.LABEL_REPEATED
                f.type_name = f".{PACKAGE}.{mname}.{entry.name}"
            elif typ == "State":
                f.type, f.label, f.type_name = _F.TYPE_ENUM, _F.LABEL_OPTIONAL, f".{PACKAGE}.State"
            else:
                f.type, f.label = _T[typ], _F.LABEL_OPTIONAL

    for mname, oneof, fields in ONEOF_WRAPPERS:
        m = fd.message_type.add()
        m.name = mname
        m.oneof_decl.add().name = oneof
        for name, typ, num in fields:
            f = m.field.add()
            f.name, f.number, f.type, f.label = name, num, _F.TYPE_MESSAGE, _F.LABEL_OPTIONAL
            f.type_name = f".{PACKAGE}.{typ}"
            f.oneof_index = 0

    svc = fd.service.add()
    svc.name = SERVICE
    meth = svc.method.add()
    meth.name = METHOD
    meth.input_type = f".{PACKAGE}.transportRequest"
    meth.output_type = f".{PACKAGE}.transportResponse"
    meth.client_streaming = True
    meth.server_streaming = True
    return fd


_pool = descriptor_pool.DescriptorPool()
FILE_DESCRIPTOR_PROTO = _build_file()
_file = _pool.Add(FILE_DESCRIPTOR_PROTO) if hasattr(_pool, "Add") else None
if _file is None:  # protobuf >= 4: AddSerializedFile returns the FileDescriptor
    _file = _pool.AddSerializedFile(FILE_DESCRIPTOR_PROTO.SerializeToString())
FILE_DESCRIPTOR = _pool.FindFileByName(FILE_NAME)


def _cls(name: str):
    return message_factory.GetMessageClass(_pool.FindMessageTypeByName(f"{PACKAGE}.{name}"))


Scalar = _cls("Scalar")
ReadyReq = _cls("ReadyReq")
UpdateReq = _cls("UpdateReq")
VersionReq = _cls("VersionReq")
ReadyRep = _cls("ReadyRep")
UpdateRep = _cls("UpdateRep")
VersionRep = _cls("VersionRep")
transportRequest = _cls("transportRequest")
transportResponse = _cls("transportResponse")
State = enum_type_wrapper.EnumTypeWrapper(_pool.FindEnumTypeByName(f"{PACKAGE}.State"))

__all__ = ["Scalar", "ReadyReq", "UpdateReq", "VersionReq", "ReadyRep", "UpdateRep", "VersionRep",
           "transportRequest", "transportResponse", "State", "PACKAGE", "SERVICE", "METHOD"]


def scalar(v) -> "Scalar":
    """Python / NumPy value -> Scalar (bool/int -> scint32, float -> scfloat, anything else -> scstring).  An integer that
    does not fit int32 travels as its decimal string (receivers convert with int()/float()), never as an exception."""
    import numbers
    import numpy as np
    if isinstance(v, (bool, numbers.Integral, np.bool_)):
        i = int(v)
        return Scalar(scint32=i) if -(1 << 31) <= i < (1 << 31) else Scalar(scstring=str(i))
    if isinstance(v, numbers.Real):
        return Scalar(scfloat=float(v))
    return Scalar(scstring=str(v))


def scalar_value(s):
    which = s.WhichOneof("scalar")
    return getattr(s, which) if which else None


def to_config(d: dict) -> dict:
    return {k: scalar(v) for k, v in d.items()}


def from_config(cfg) -> dict:
    return {k: scalar_value(v) for k, v in cfg.items()}
