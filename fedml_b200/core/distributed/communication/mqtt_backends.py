"""MQTT, MQTT_S3 and MQTT_S3_MNN communication managers and the MLOps status channel.

Topic scheme and behaviour follow the reference:
  * MQTT (mqtt/mqtt_comm_manager.py:47-124): topics ``fedml<cid>`` (client -> server) and ``fedml0_<cid>``
    (server -> client), JSON payload with the model as nested lists;
  * MQTT_S3 (mqtt_s3/mqtt_s3_multi_clients_comm_manager.py:78-323): control plane on
    ``fedml_<run_id>_<cid>`` / ``fedml_<run_id>_0_<cid>``; ``model_params`` is written to the object store and
    replaced by its key (+ ``model_params_url``) on send, and fetched back on receive; a synthetic message of
    type 0 ("connection ready") is emitted once subscriptions are in place; a last-will ``W/topic`` carries
    ``{ID, stat: "Online"}``;
  * MQTT_S3_MNN (mqtt_s3_mnn/mqtt_s3_comm_manager.py:197-235): the payload is a model *file* path.
Broker: any MQTT 3.1.1 broker at ``mqtt_host:mqtt_port`` (defaults 127.0.0.1:1883) — the bundled :class:`LocalBroker`
or a production one; the client is the bundled protocol implementation (``broker.BrokerClient``) or paho
(``args.mqtt_client: "paho"``), both speak the standard wire protocol."""
import json
import os
import queue
import uuid
from typing import List

import yaml

from .base_com_manager import BaseCommunicationManager
from .broker import BrokerClient, make_client
from .message import Message, _jsonable
from .observer import Observer
from .remote_storage import create_store

_STOP = object()


def _tls_kwargs(args):
    """``mqtt_tls: true`` / ``mqtt_ca_certs: <bundle>`` / ``mqtt_tls_insecure: true`` in the YAML -> TLS options of the native client."""
    if not (getattr(args, "mqtt_tls", False) or getattr(args, "mqtt_ca_certs", None)) or str(getattr(args, "mqtt_client", "auto")).lower() == "paho":
        return {}
    return {"tls": True, "ca_certs": getattr(args, "mqtt_ca_certs", None), "tls_insecure": bool(getattr(args, "mqtt_tls_insecure", False))}


def _broker_endpoint(args):
    host, port = getattr(args, "mqtt_host", "127.0.0.1"), int(getattr(args, "mqtt_port", 1883))
    path = getattr(args, "mqtt_config_path", None)
    if path and os.path.exists(str(path)):
        with open(path) as fh:
            cfg = yaml.safe_load(fh) or {}
        host, port = cfg.get("BROKER_HOST", host), int(cfg.get("BROKER_PORT", port))
    return host, port


class _MqttBase(BaseCommunicationManager):
    def __init__(self, args, topic="fedml", client_rank=0, client_num=0):
        self.args, self.rank, self.client_num = args, int(client_rank), int(client_num)
        self.run_id = str(getattr(args, "run_id", 0))
        self._topic = f"{topic}_{self.run_id}_"
        self._observers: List[Observer] = []
        self._inbox: "queue.Queue" = queue.Queue()
        self._running = False
        ids = args.client_ids() if hasattr(args, "client_ids") else []
        self.client_real_ids = ids or list(range(1, self.client_num + 1))
        self.edge_id = 0 if self.rank == 0 else (self.client_real_ids[self.rank - 1]
                                                 if self.rank - 1 < len(self.client_real_ids) else self.rank)
        host, port = _broker_endpoint(args)
        self._client = make_client(host, port, kind=getattr(args, "mqtt_client", "auto"), client_id=f"{self.run_id}-{self.edge_id}", will_topic="W/topic",
                                    will_payload={"ID": self.edge_id, "stat": "Online"}, **_tls_kwargs(args))
        self._client.on_message = self._on_message
        if self.rank == 0:
            for cid in self.client_real_ids:
                self._client.subscribe(self._topic + str(cid))
        else:
            self._client.subscribe(self._topic + "0_" + str(self.edge_id))
        self._inbox.put(Message(0, self.rank, self.rank))      # "connection ready"

    # -- hooks ---------------------------------------------------------------
    def _encode(self, msg: Message) -> bytes:
        return json.dumps(_jsonable(msg.get_params())).encode()

    def _decode(self, payload: bytes) -> Message:
        m = Message()
        m.init_from_json_string(payload.decode())
        return m

    def _on_message(self, topic, payload):
        self._inbox.put(self._decode(payload))

    def _topic_for(self, msg: Message):
        recv = int(msg.get_receiver_id())
        if self.rank == 0:
            return self._topic + "0_" + str(recv)
        return self._topic + str(self.edge_id)

    def send_message(self, msg: Message):
        self._client.publish(self._topic_for(msg), self._encode(msg))

    def add_observer(self, observer: Observer):
        self._observers.append(observer)

    def remove_observer(self, observer: Observer):
        self._observers.remove(observer)

    def handle_receive_message(self):
        self._running = True
        while self._running:
            msg = self._inbox.get()
            if msg is _STOP:
                break
            for obs in list(self._observers):
                obs.receive_message(msg.get_type(), msg.get_params())

    def stop_receive_message(self):
        self._running = False
        self._inbox.put(_STOP)
        self._client.disconnect()


class MqttCommManager(_MqttBase):
    """Pure MQTT.  Default payload: the whole message as JSON with tensors as nested lists - the reference's format
    (mqtt_comm_manager.py:100-124), readable by its mobile SDK.  ``mqtt_binary: true`` publishes the binary tensor codec instead
    (``Message.to_bytes``; ~100x the throughput, see profiles/r2/comm_latency_cpu.md); receivers recognise either form by the
    payload's first bytes, so peers with different settings interoperate."""

    def _encode(self, msg: Message) -> bytes:
        if bool(getattr(self.args, "mqtt_binary", False)):
            return msg.to_bytes()
        return super()._encode(msg)

    def _decode(self, payload):
        import torch
        if bytes(payload[:4]) == b"FB2M":
            return Message.from_bytes(bytes(payload))
        m = super()._decode(payload)
        mp = m.get(Message.MSG_ARG_KEY_MODEL_PARAMS)
        if isinstance(mp, dict):
            m.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, {k: torch.tensor(v) for k, v in mp.items()})
        return m


class MqttS3MultiClientsCommManager(_MqttBase):
    def __init__(self, args, topic="fedml", client_rank=0, client_num=0):
        self.store = create_store(args)
        super().__init__(args, topic, client_rank, client_num)

    def _encode(self, msg: Message) -> bytes:
        params = dict(msg.get_params())
        model = params.get(Message.MSG_ARG_KEY_MODEL_PARAMS)
        if model is not None and not isinstance(model, str):
            key = self._topic_for(msg) + "_" + uuid.uuid4().hex
            url = self.store.write_model(key, model)
            params[Message.MSG_ARG_KEY_MODEL_PARAMS] = key
            params[Message.MSG_ARG_KEY_MODEL_PARAMS_URL] = url
            msg.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS_URL, url)   # visible to the sender (MLOps reporting)
        return json.dumps(_jsonable(params)).encode()

    def _decode(self, payload):
        m = _MqttBase._decode(self, payload)
        key = m.get(Message.MSG_ARG_KEY_MODEL_PARAMS)
        if isinstance(key, str):
            m.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, self.store.read_model(key))
        return m


class MqttS3MNNCommManager(MqttS3MultiClientsCommManager):
    """Server-side variant for cross-device: ``model_params`` is a path to a model file that is uploaded as a
    file object; on receive the file is downloaded into ``model_file_cache_folder`` and the local path is
    handed to the handler."""

    def _encode(self, msg: Message) -> bytes:
        params = dict(msg.get_params())
        path = params.get(Message.MSG_ARG_KEY_MODEL_PARAMS)
        if isinstance(path, str) and os.path.exists(path):
            key = self._topic_for(msg) + "_" + uuid.uuid4().hex + ".mnn"
            params[Message.MSG_ARG_KEY_MODEL_PARAMS_URL] = self.store.upload_file(path, key)
            params[Message.MSG_ARG_KEY_MODEL_PARAMS] = key
        return json.dumps(_jsonable(params)).encode()

    def _decode(self, payload):
        m = _MqttBase._decode(self, payload)
        key = m.get(Message.MSG_ARG_KEY_MODEL_PARAMS)
        if isinstance(key, str) and key.endswith(".mnn"):
            folder = getattr(self.args, "model_file_cache_folder", "./model_file_cache")
            dst = os.path.join(folder, os.path.basename(key))
            m.add_params(Message.MSG_ARG_KEY_MODEL_PARAMS, self.store.download_file(key, dst))
        return m


class MqttS3StatusManager:
    """Side channel for MLOps status / metrics JSON (reference: mqtt_s3/mqtt_s3_status_manager.py:17-160)."""

    def __init__(self, args, topic="fedml_status"):
        host, port = _broker_endpoint(args)
        self._client = make_client(host, port, kind=getattr(args, "mqtt_client", "auto"),
                                   client_id=f"status-{getattr(args, 'run_id', 0)}-{uuid.uuid4().hex[:6]}", **_tls_kwargs(args))
        self._client.on_message = self._on_message
        self._observers = []

    def _on_message(self, topic, payload):
        for obs in list(self._observers):
            obs.receive_message(topic, payload.decode())

    def add_observer(self, observer):
        self._observers.append(observer)

    def remove_observer(self, observer):
        self._observers.remove(observer)

    def subscribe(self, topic):
        self._client.subscribe(topic)

    def send_message_json(self, topic_name, json_message):
        self._client.publish(topic_name, json_message if isinstance(json_message, str) else json.dumps(json_message))

    def stop(self):
        self._client.disconnect()
