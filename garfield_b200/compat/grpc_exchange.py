"""gRPC pull service `MessageExchange` (GetModel / GetGradient) — the TF implementation's transport.

Parity with tensorflow_impl/libs/garfield.proto:3-33 and grpc_message_exchange_servicer.py:39-91:
unary pulls keyed by iteration number; the servicer keeps per-iteration histories and blocks until
the requested iteration has been committed; payloads are raw float32 bytes; 500 MiB message cap.
Implemented with grpc generic handlers and a tiny fixed binary framing (no protoc needed):

    request  = <int32 iter><int32 req_id><utf-8 job>
    reply    = <int32 iter><uint8 init><raw float32 payload>
"""

from __future__ import annotations

import struct
import threading
import time
from concurrent import futures

import numpy as np

try:
    import grpc
except ImportError:  # pragma: no cover
    grpc = None

__all__ = ["MessageExchangeServicer", "serve", "Client", "MAX_MESSAGE"]

MAX_MESSAGE = 500 * 1024 * 1024
_SERVICE = "MessageExchange"


def _pack_request(iteration, job="", req_id=0):
    return struct.pack("<ii", int(iteration), int(req_id)) + job.encode()


def _unpack_request(blob):
    iteration, req_id = struct.unpack_from("<ii", blob)
    return iteration, req_id, blob[8:].decode()


def _pack_reply(iteration, array, init=False):
    return struct.pack("<iB", int(iteration), 1 if init else 0) + np.ascontiguousarray(array, dtype=np.float32).tobytes()


def _unpack_reply(blob):
    iteration, init = struct.unpack_from("<iB", blob)
    return iteration, bool(init), np.frombuffer(blob, dtype=np.float32, offset=5)


class MessageExchangeServicer:
    """History lists indexed by iteration; `GetModel`/`GetGradient` block until committed."""

    def __init__(self, poll=0.001, timeout=60.0):
        self.model_history, self.gradient_history = [], []
        self._cv = threading.Condition()
        self.poll, self.timeout = poll, timeout

    def commit_model(self, flat):
        with self._cv:
            self.model_history.append(np.asarray(flat, dtype=np.float32).copy())
            self._cv.notify_all()

    def commit_gradient(self, flat):
        with self._cv:
            self.gradient_history.append(np.asarray(flat, dtype=np.float32).copy())
            self._cv.notify_all()

    def _wait(self, history, iteration):
        with self._cv:
            ok = self._cv.wait_for(lambda: len(history) > iteration, self.timeout)
            if not ok:
                raise TimeoutError("iteration %d was never committed" % iteration)
            return history[iteration]

    def GetModel(self, request, context=None):
        iteration, _, _ = _unpack_request(request)
        return _pack_reply(iteration, self._wait(self.model_history, iteration), init=(iteration == 0))

    def GetGradient(self, request, context=None):
        iteration, _, _ = _unpack_request(request)
        return _pack_reply(iteration, self._wait(self.gradient_history, iteration))


def serve(servicer, port, max_workers=30):
    """Start a gRPC server for `servicer` on 127.0.0.1/0.0.0.0:`port`; returns the server object."""
    if grpc is None:
        raise RuntimeError("grpcio is not installed")
    handlers = {name: grpc.unary_unary_rpc_method_handler(getattr(servicer, name)) for name in ("GetModel", "GetGradient")}
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers),
                         options=[("grpc.max_send_message_length", MAX_MESSAGE), ("grpc.max_receive_message_length", MAX_MESSAGE)])
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(_SERVICE, handlers),))
    server.add_insecure_port("[::]:%d" % port)
    server.start()
    return server


class Client:
    """Pull models / gradients from a peer; retries while the peer is not up yet (as the reference does)."""

    def __init__(self, address, retries=10, wait=0.5):
        if grpc is None:
            raise RuntimeError("grpcio is not installed")
        self.channel = grpc.insecure_channel(address, options=[("grpc.max_send_message_length", MAX_MESSAGE),
                                                               ("grpc.max_receive_message_length", MAX_MESSAGE)])
        self.retries, self.wait = retries, wait
        self._get_model = self.channel.unary_unary("/%s/GetModel" % _SERVICE)
        self._get_gradient = self.channel.unary_unary("/%s/GetGradient" % _SERVICE)

    def _call(self, stub, iteration, job):
        last = None
        for _ in range(self.retries):
            try:
                return _unpack_reply(stub(_pack_request(iteration, job), timeout=120))
            except grpc.RpcError as err:
                last = err
                time.sleep(self.wait)
        raise last

    def get_model(self, iteration, job="worker"):
        return self._call(self._get_model, iteration, job)[2]

    def get_gradient(self, iteration, job="ps"):
        return self._call(self._get_gradient, iteration, job)[2]

    def close(self):
        self.channel.close()
