"""Message envelope exchanged by FL managers (reference: communication/message.py:5-80).

A dict with ``msg_type/sender/receiver`` plus arbitrary params.  Beyond the reference's JSON
round-trip this adds a compact binary codec (``to_bytes``/``from_bytes``): tensors are shipped
as raw little-endian buffers after a msgpack header instead of pickled objects or JSON nested
lists, which is what the gRPC / socket / file backends put on the wire.
"""
import io
import json
import struct

import numpy as np

try:
    import msgpack
except Exception:  # pragma: no cover
    msgpack = None


class Message(object):
    MSG_ARG_KEY_OPERATION = "operation"
    MSG_ARG_KEY_TYPE = "msg_type"
    MSG_ARG_KEY_SENDER = "sender"
    MSG_ARG_KEY_RECEIVER = "receiver"

    MSG_OPERATION_SEND = "send"
    MSG_OPERATION_RECEIVE = "receive"
    MSG_OPERATION_BROADCAST = "broadcast"
    MSG_OPERATION_REDUCE = "reduce"

    MSG_ARG_KEY_MODEL_PARAMS = "model_params"
    MSG_ARG_KEY_MODEL_PARAMS_URL = "model_params_url"

    def __init__(self, type=0, sender_id=0, receiver_id=0):  # noqa: A002 - reference signature
        self.type = type
        self.sender_id = sender_id
        self.receiver_id = receiver_id
        self.msg_params = {
            Message.MSG_ARG_KEY_TYPE: type,
            Message.MSG_ARG_KEY_SENDER: sender_id,
            Message.MSG_ARG_KEY_RECEIVER: receiver_id,
        }

    # -- construction ------------------------------------------------------
    def init(self, msg_params):
        self.msg_params = msg_params
        self._sync_header()

    def init_from_json_string(self, json_string):
        self.init(json.loads(json_string))

    def init_from_json_object(self, json_object):
        self.init(json_object)

    def _sync_header(self):
        self.type = self.msg_params.get(Message.MSG_ARG_KEY_TYPE, self.type)
        self.sender_id = self.msg_params.get(Message.MSG_ARG_KEY_SENDER, self.sender_id)
        self.receiver_id = self.msg_params.get(Message.MSG_ARG_KEY_RECEIVER, self.receiver_id)

    # -- accessors ---------------------------------------------------------
    def get_sender_id(self):
        return self.sender_id

    def get_receiver_id(self):
        return self.receiver_id

    def add_params(self, key, value):
        self.msg_params[key] = value

    add = add_params

    def get_params(self):
        return self.msg_params

    def get(self, key, default=None):
        return self.msg_params.get(key, default)

    def get_type(self):
        return self.msg_params[Message.MSG_ARG_KEY_TYPE]

    def to_string(self):
        return self.msg_params

    def to_json(self):
        return json.dumps(_jsonable(self.msg_params))

    def get_content(self):
        return f"{self.get_type()}: {{{', '.join(map(str, self.msg_params))}}}"

    # -- binary codec ------------------------------------------------------
    def to_buffers(self) -> list:
        """The wire form as a list of buffers whose concatenation is ``to_bytes()``: tensor payloads are views of the tensors'
        own memory (no copy), so a socket transport can hand them to the kernel directly."""
        blobs = []
        header = _encode(self.msg_params, blobs)
        head = msgpack.packb(header, use_bin_type=True) if msgpack else json.dumps(header).encode()
        out = [struct.pack("<4sBIQ", b"FB2M", 1 if msgpack else 0, len(blobs), len(head)) + head]
        for b in blobs:
            out.append(struct.pack("<Q", b.nbytes if isinstance(b, memoryview) else len(b)))
            out.append(b)
        return out

    def to_bytes(self) -> bytes:
        return b"".join(self.to_buffers())

    @classmethod
    def from_bytes(cls, data: bytes) -> "Message":
        magic, packed, nblobs, hlen = struct.unpack_from("<4sBIQ", data, 0)
        if magic != b"FB2M":
            raise ValueError("not a fedml_b200 message")
        off = struct.calcsize("<4sBIQ")
        head = data[off:off + hlen]
        off += hlen
        header = msgpack.unpackb(head, raw=False, strict_map_key=False) if packed else json.loads(head)
        blobs = []
        for _ in range(nblobs):
            (n,) = struct.unpack_from("<Q", data, off)
            off += 8
            blobs.append(memoryview(data)[off:off + n])
            off += n
        msg = cls()
        msg.init(_decode(header, blobs))
        return msg


def _jsonable(obj):
    import torch
    if isinstance(obj, dict):
        return {str(k): _jsonable(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return [_jsonable(v) for v in obj]
    if isinstance(obj, torch.Tensor):
        return obj.detach().cpu().tolist()
    if isinstance(obj, np.ndarray):
        return obj.tolist()
    if isinstance(obj, (np.integer,)):
        return int(obj)
    if isinstance(obj, (np.floating,)):
        return float(obj)
    if isinstance(obj, (bytes, bytearray)):
        import base64
        return {"__b64": base64.b64encode(bytes(obj)).decode()}
    return obj


def _byte_view(a: np.ndarray):
    """The array's memory as a flat byte view (keeps the array alive); bytes for empty / exotic layouts."""
    if a.size == 0:
        return b""
    try:
        return memoryview(a.reshape(-1)).cast("B")
    except (TypeError, ValueError):
        return a.tobytes()


def _encode(obj, blobs):
    import torch
    if isinstance(obj, dict):
        return {"__d": [[_encode(k, blobs), _encode(v, blobs)] for k, v in obj.items()]}
    if isinstance(obj, (list, tuple)):
        return {"__l": [_encode(v, blobs) for v in obj], "t": isinstance(obj, tuple)}
    if isinstance(obj, torch.Tensor):
        t = obj.detach().cpu().contiguous()
        raw = t.view(torch.int16) if t.dtype == torch.bfloat16 else t
        blobs.append(_byte_view(raw.numpy()))
        return {"__t": len(blobs) - 1, "dt": str(t.dtype).replace("torch.", ""), "sh": list(t.shape)}
    if isinstance(obj, np.ndarray):
        blobs.append(_byte_view(np.ascontiguousarray(obj)))
        return {"__n": len(blobs) - 1, "dt": str(obj.dtype), "sh": list(obj.shape)}
    if isinstance(obj, np.generic):
        return obj.item()
    if isinstance(obj, bytes):
        blobs.append(obj)
        return {"__b": len(blobs) - 1}
    return obj


def _decode(obj, blobs):
    import torch
    if isinstance(obj, dict):
        if "__d" in obj:
            return {_hashable(_decode(k, blobs)): _decode(v, blobs) for k, v in obj["__d"]}
        if "__l" in obj:
            seq = [_decode(v, blobs) for v in obj["__l"]]
            return tuple(seq) if obj.get("t") else seq
        if "__t" in obj:
            buf = bytearray(blobs[obj["__t"]])
            if obj["dt"] == "bfloat16":
                return torch.frombuffer(buf, dtype=torch.int16).view(torch.bfloat16).reshape(obj["sh"]) \
                    if len(buf) else torch.zeros(obj["sh"], dtype=torch.bfloat16)
            dt = getattr(torch, obj["dt"])
            return torch.frombuffer(buf, dtype=dt).reshape(obj["sh"]) if len(buf) else torch.zeros(obj["sh"], dtype=dt)
        if "__n" in obj:
            return np.frombuffer(bytes(blobs[obj["__n"]]), dtype=np.dtype(obj["dt"])).reshape(obj["sh"]).copy()
        if "__b" in obj:
            return bytes(blobs[obj["__b"]])
    return obj


def _hashable(k):
    return tuple(k) if isinstance(k, list) else k
