"""The original (TF1-era) Garfield wire protocol `TrainMessageExchange` and its worker-side filters.

Parity target: tensorflow_impl/applications/Garfield_legacy/all.proto (service `TrainMessageExchange`:
GetPublicKey, GetUnifiedModel, GetCompleteModel, GetOnlyHash, GetGradHashes, GetGradHash, GetGradients, GetModel),
grpc_service_impl.py:42-66 (the servicer only ever fills `GetUnifiedModel` and `GetGradients`; here every declared
RPC is served), byzWorker.py:398-425 (the "smart" worker's Lipschitz filter and model-drift bound) and
update_model.cc (`model − 0.1·grad`). SURVEY.md §2.4 C17, §2.7 T11.

Design: same transport as `grpc_exchange` — grpc generic handlers, no protoc — with a small tagged binary framing:

    request = <int32 iter><int32 req_id>
    reply   = <int32 iter><uint8 flags><uint32 n_fields>{<uint32 len><bytes>}*

Models/gradients travel as raw float32; hashes are SHA-256 of those bytes; signatures are Ed25519 over the hash when
the `cryptography` package is importable, else HMAC-SHA256 with the node's secret (the reference only declares the
fields). Everything also works in-process (call the servicer methods directly), which is how the tests drive it.
"""

from __future__ import annotations

import hashlib
import hmac
import os
import struct
import threading
from concurrent import futures

import numpy as np

try:
    import grpc
except ImportError:  # pragma: no cover
    grpc = None

try:
    from cryptography.hazmat.primitives import serialization
    from cryptography.hazmat.primitives.asymmetric.ed25519 import Ed25519PrivateKey, Ed25519PublicKey
except ImportError:  # pragma: no cover
    Ed25519PrivateKey = None

__all__ = ["TrainMessageExchangeServicer", "LegacyClient", "serve_legacy", "LipschitzFilter", "drift_bound", "update_model",
           "model_hash", "Signer"]

_SERVICE = "TrainMessageExchange"
_METHODS = ("GetPublicKey", "GetUnifiedModel", "GetCompleteModel", "GetOnlyHash", "GetGradHashes", "GetGradHash", "GetGradients", "GetModel")
MAX_MESSAGE = 500 * 1024 * 1024


def _f32(array):
    return np.ascontiguousarray(np.asarray(array, dtype=np.float32).reshape(-1))


def model_hash(array):
    """SHA-256 of the float32 little-endian bytes of a flat model or gradient."""
    return hashlib.sha256(_f32(array).tobytes()).digest()


def _pack(iteration, fields, flags=0):
    out = [struct.pack("<iBI", int(iteration), flags, len(fields))]
    for f in fields:
        out.append(struct.pack("<I", len(f)))
        out.append(bytes(f))
    return b"".join(out)


def _unpack(blob):
    iteration, flags, count = struct.unpack_from("<iBI", blob)
    at, fields = 9, []
    for _ in range(count):
        (n,) = struct.unpack_from("<I", blob, at)
        fields.append(blob[at + 4:at + 4 + n])
        at += 4 + n
    return iteration, flags, fields


def _request(iteration=0, req_id=0):
    return struct.pack("<ii", int(iteration), int(req_id))


class Signer:
    """Ed25519 (or HMAC fallback) signatures over message hashes; `public_bytes` is what GetPublicKey returns."""

    def __init__(self, secret=None):
        if Ed25519PrivateKey is not None:
            self._key = Ed25519PrivateKey.generate() if secret is None else Ed25519PrivateKey.from_private_bytes(hashlib.sha256(secret).digest())
            self.public_bytes = self._key.public_key().public_bytes(serialization.Encoding.Raw, serialization.PublicFormat.Raw)
            self.scheme = "ed25519"
        else:  # pragma: no cover
            self._key = secret or os.urandom(32)
            self.public_bytes = hashlib.sha256(self._key).digest()
            self.scheme = "hmac"

    def sign(self, digest):
        if self.scheme == "ed25519":
            return self._key.sign(digest)
        return hmac.new(self._key, digest, hashlib.sha256).digest()  # pragma: no cover

    @staticmethod
    def verify(public_bytes, digest, signature):
        """True when `signature` is `public_bytes`' owner's signature of `digest` (Ed25519 scheme only)."""
        if Ed25519PrivateKey is None:  # pragma: no cover
            return False
        try:
            Ed25519PublicKey.from_public_bytes(public_bytes).verify(signature, digest)
            return True
        except Exception:
            return False


class TrainMessageExchangeServicer:
    """State one node publishes: its latest model / gradient (+ Lipschitz coefficient), the batch it was computed on,
    hashes and signatures. Pull RPCs block until the requested iteration has been published (the reference spins with
    `sleep(0.05)`, grpc_service_impl.py:52-66)."""

    def __init__(self, index=0, secret=None, timeout=60.0):
        self.index, self.timeout = index, timeout
        self.signer = Signer(secret)
        self._cv = threading.Condition()
        self._model = None          # (iter, flat)
        self._gradient = None       # (iter, flat, lipschitz)
        self._batch = (b"", b"")
        self._correct = []
        self._peer_signatures = []
        self._grad_hashes = {}      # iter -> {index: hash}

    # -- publishing ---------------------------------------------------------------------------------
    def publish_model(self, iteration, flat, inputs=None, labels=None, correct=(), signatures=()):
        with self._cv:
            self._model = (int(iteration), _f32(flat).copy())
            self._batch = (b"" if inputs is None else np.ascontiguousarray(inputs).tobytes(),
                           b"" if labels is None else np.ascontiguousarray(labels).tobytes())
            self._correct, self._peer_signatures = list(correct), list(signatures)
            self._cv.notify_all()

    def publish_gradient(self, iteration, flat, lipschitz=0.0):
        with self._cv:
            flat = _f32(flat).copy()
            self._gradient = (int(iteration), flat, float(lipschitz))
            self._grad_hashes.setdefault(int(iteration), {})[self.index] = model_hash(flat)
            self._cv.notify_all()

    def record_peer_hash(self, iteration, index, digest):
        with self._cv:
            self._grad_hashes.setdefault(int(iteration), {})[int(index)] = bytes(digest)
            self._cv.notify_all()

    def _await(self, getter, iteration):
        with self._cv:
            if not self._cv.wait_for(lambda: getter() is not None and getter()[0] >= iteration, self.timeout):
                raise TimeoutError("iteration %d was never published" % iteration)
            return getter()

    # -- RPCs (bytes in, bytes out) -------------------------------------------------------------------
    def GetPublicKey(self, request=b"", context=None):
        return _pack(self.index, [self.signer.public_bytes])

    def GetUnifiedModel(self, request=b"", context=None):
        it, flat = self._await(lambda: self._model, -(1 << 30))
        return _pack(it, [flat.tobytes()], flags=1)

    def GetModel(self, request, context=None):
        iteration, _ = struct.unpack_from("<ii", request)
        it, flat = self._await(lambda: self._model, iteration)
        return _pack(it, [flat.tobytes()], flags=1)

    def GetCompleteModel(self, request, context=None):
        """Model + the batch it should be evaluated on + hash, this node's signature and the peers' it collected."""
        iteration, _ = struct.unpack_from("<ii", request)
        it, flat = self._await(lambda: self._model, iteration)
        digest = model_hash(flat)
        fields = [flat.tobytes(), self._batch[0], self._batch[1], digest, self.signer.sign(digest),
                  np.asarray(self._correct, dtype=np.int32).tobytes()] + list(self._peer_signatures)
        return _pack(it, fields, flags=1)

    def GetOnlyHash(self, request, context=None):
        iteration, _ = struct.unpack_from("<ii", request)
        it, flat = self._await(lambda: self._model, iteration)
        digest = model_hash(flat)
        return _pack(it, [digest, self.signer.sign(digest)])

    def GetGradients(self, request, context=None):
        iteration, _ = struct.unpack_from("<ii", request)
        it, flat, lip = self._await(lambda: self._gradient, iteration)
        return _pack(it, [flat.tobytes(), struct.pack("<f", lip)])

    def GetGradHash(self, request, context=None):
        iteration, _ = struct.unpack_from("<ii", request)
        it, flat, _ = self._await(lambda: self._gradient, iteration)
        return _pack(it, [struct.pack("<i", self.index), model_hash(flat)])

    def GetGradHashes(self, request, context=None):
        iteration, _ = struct.unpack_from("<ii", request)
        with self._cv:
            known = dict(self._grad_hashes.get(iteration, {}))
        fields = []
        for index in sorted(known):
            fields += [struct.pack("<i", index), known[index]]
        return _pack(iteration, fields)


def serve_legacy(servicer, port, max_workers=30):
    if grpc is None:
        raise RuntimeError("grpcio is not installed")
    handlers = {name: grpc.unary_unary_rpc_method_handler(getattr(servicer, name)) for name in _METHODS}
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers),
                         options=[("grpc.max_send_message_length", MAX_MESSAGE), ("grpc.max_receive_message_length", MAX_MESSAGE)])
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(_SERVICE, handlers),))
    bound = server.add_insecure_port("127.0.0.1:%d" % port)
    server.start()
    server.bound_port = bound
    return server


class LegacyClient:
    """Typed wrappers over the eight RPCs; `target` is an address ("host:port") or a servicer object (in-process)."""

    def __init__(self, target):
        self._local = target if not isinstance(target, str) else None
        if self._local is None:
            if grpc is None:
                raise RuntimeError("grpcio is not installed")
            self.channel = grpc.insecure_channel(target, options=[("grpc.max_send_message_length", MAX_MESSAGE),
                                                                  ("grpc.max_receive_message_length", MAX_MESSAGE)])
            self._stubs = {name: self.channel.unary_unary("/%s/%s" % (_SERVICE, name)) for name in _METHODS}

    def _call(self, name, iteration=0, req_id=0):
        blob = _request(iteration, req_id)
        reply = getattr(self._local, name)(blob) if self._local is not None else self._stubs[name](blob, timeout=120)
        return _unpack(reply)

    def get_public_key(self):
        index, _, fields = self._call("GetPublicKey")
        return index, fields[0]

    def get_unified_model(self):
        it, _, fields = self._call("GetUnifiedModel")
        return it, np.frombuffer(fields[0], dtype=np.float32)

    def get_model(self, iteration, req_id=0):
        it, _, fields = self._call("GetModel", iteration, req_id)
        return it, np.frombuffer(fields[0], dtype=np.float32)

    def get_complete_model(self, iteration, req_id=0):
        it, _, f = self._call("GetCompleteModel", iteration, req_id)
        return dict(iter=it, model=np.frombuffer(f[0], dtype=np.float32), inputs=f[1], labels=f[2], msg_hash=f[3], signature=f[4],
                    correct=np.frombuffer(f[5], dtype=np.int32).tolist(), signatures=f[6:])

    def get_only_hash(self, iteration, req_id=0):
        it, _, f = self._call("GetOnlyHash", iteration, req_id)
        return it, f[0], f[1]

    def get_gradients(self, iteration, req_id=0):
        it, _, f = self._call("GetGradients", iteration, req_id)
        return it, np.frombuffer(f[0], dtype=np.float32), struct.unpack("<f", f[1])[0]

    def get_grad_hash(self, iteration, req_id=0):
        it, _, f = self._call("GetGradHash", iteration, req_id)
        return struct.unpack("<i", f[0])[0], f[1]

    def get_grad_hashes(self, iteration, req_id=0):
        _, _, f = self._call("GetGradHashes", iteration, req_id)
        return {struct.unpack("<i", f[i])[0]: f[i + 1] for i in range(0, len(f), 2)}

    def close(self):
        if self._local is None:
            self.channel.close()


class LipschitzFilter:
    """The "smart" worker's model filter (byzWorker.py:398-425): the empirical Lipschitz coefficient
    ‖g_t − g_{t−1}‖ / ‖x_t − x_{t−1}‖ of the model a server sent is compared with the (num_ps − num_byzps)/num_ps
    percentile of the coefficients seen so far; the history is trimmed to its last 50 entries when it reaches 100."""

    def __init__(self, num_ps, num_byzps):
        self.quantile = 100.0 * (num_ps - num_byzps) / num_ps
        self.history = []

    @staticmethod
    def coefficient(grad, past_grad, model, past_model):
        den = float(np.linalg.norm(np.asarray(model, dtype=np.float64) - np.asarray(past_model, dtype=np.float64)))
        num = float(np.linalg.norm(np.asarray(grad, dtype=np.float64) - np.asarray(past_grad, dtype=np.float64)))
        return num / den if den > 0 else float("inf")

    def observe(self, value):
        """Record a coefficient; returns (accepted, percentile) — accepted iff value ≤ the current percentile."""
        self.history.append(float(value))
        bound = float(np.percentile(self.history, self.quantile))
        if len(self.history) == 100:
            self.history = self.history[-50:]
        return value <= bound, bound


def drift_bound(rate, grad_norm, period, num_workers, num_byzwrks, iteration):
    """Largest distance a served model may be from the worker's own speculative model (byzWorker.py:424):
    rate·‖g‖·((3T+2)(n−f)/(4f) + 2((i−1) mod T)). Only meaningful for f > 0."""
    if num_byzwrks <= 0:
        return float("inf")
    return rate * grad_norm * (((3 * period + 2) * (num_workers - num_byzwrks)) / (4.0 * num_byzwrks) + 2 * ((iteration - 1) % period))


def update_model(model, grad, lr=0.1):
    """The legacy `UpdateModel` custom op (update_model.cc): model − lr·grad, lr = 0.1. Device tensors use the
    library's axpby kernel; numpy / CPU tensors are computed in place of TensorFlow's CPU kernel."""
    try:
        import torch
    except ImportError:  # pragma: no cover
        torch = None
    if torch is not None and isinstance(model, torch.Tensor):
        if model.is_cuda:
            from ..ops import _lib
            out = grad.to(torch.float32).clone().reshape(-1)           # y := 1·model + (−lr)·y
            with torch.cuda.device(model.device):
                _lib.check(_lib.require().gb_axpby(model.contiguous().data_ptr(), out.data_ptr(), out.numel(), 1.0, -float(lr),
                                                   _lib.stream_ptr(model.device)))
            _lib.count()
            return out.view_as(model)
        return model - lr * grad
    return np.asarray(model, dtype=np.float32) - np.float32(lr) * np.asarray(grad, dtype=np.float32)
