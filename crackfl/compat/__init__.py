"""Drop-in modules under the names the reference imports (``fl_server.py:8-12``, ``fl_client.py:8-11``):

    transport_pb2, transport_pb2_grpc   (lost generated code: rebuilt from crackfl/protocol/transport.proto)
    model_evaluate                      (module the reference imports but does not ship: SURVEY §2.4)
    client_fit_model                    (``learning_fit`` on top of crackfl.trainer.LocalTrainer)

``crackfl.compat.install()`` registers them in ``sys.modules`` so that reference-style code such as
``from transport_pb2 import Scalar`` or ``from client_fit_model import learning_fit`` keeps working.
"""
import importlib
import sys

NAMES = ("transport_pb2", "transport_pb2_grpc", "model_evaluate", "client_fit_model")


def install(overwrite: bool = False):
    for n in NAMES:
        if overwrite or n not in sys.modules:
            sys.modules[n] = importlib.import_module(f"crackfl.compat.{n}")
