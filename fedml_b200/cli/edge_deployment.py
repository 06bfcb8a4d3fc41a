"""Helpers of the edge (client) agent under the names the reference's ``cli/edge_deployment`` package exports.

* ``MqttManager`` - topic -> listener dispatch over one MQTT connection with the agent's three control topics subscribed on
  connect (reference mqtt_manager.py:8-77).  Runs on paho when that is importable, else on the bundled MQTT 3.1.1 client; both
  speak the standard protocol, so either side of a deployment can be the reference.
* ``load_yaml_config`` (yaml_utils.py:4-12), ``upload_file`` / ``download_file`` / ``delete_s3_zip`` over the configured object store
  (remote_storage.py:19-70; a bucket with ``FEDML_S3_CONFIG=<yaml>`` and boto3, else the directory ``FEDML_OBJECT_STORE_DIR``),
  ``save_edge_infos`` / ``login`` / ``logout`` (login.py:482-560: bind the device, remember the ids, start the agent loop), ``login_with_docker_mode`` (login_with_docker.py:9-60: the same agent inside a container - the command line is built,
  docker is only invoked when it exists).
"""
from __future__ import annotations

import json
import logging
import os
import platform
import shutil
import subprocess
import threading
import uuid
from typing import Callable, Dict, Optional

import yaml


class MqttManager:
    MSG_MLOPS_CLIENT_STATUS_IDLE = "IDLE"
    MSG_MLOPS_CLIENT_STATUS_UPGRADING = "UPGRADING"
    MSG_MLOPS_CLIENT_STATUS_INITIALIZING = "INITIALIZING"
    MSG_MLOPS_CLIENT_STATUS_TRAINING = "TRAINING"
    MSG_MLOPS_CLIENT_STATUS_STOPPING = "STOPPING"
    MSG_MLOPS_CLIENT_STATUS_FINISHED = "FINISHED"

    def __init__(self, host, port, user, pwd, keepalive_time, id):  # noqa: A002  (the reference's parameter name)
        self._host, self._port, self._client_id = host, int(port), str(id)
        self._listeners: Dict[str, Callable[[str, str], None]] = {}
        self.mqtt_connection_id = uuid.uuid4().hex[:22]
        self._stopped = threading.Event()
        will = json.dumps({"ID": self.mqtt_connection_id, "stat": "Online"})
        try:
            import paho.mqtt.client as mqtt
            try:
                self._paho = mqtt.Client(client_id=self.mqtt_connection_id, clean_session=True)
            except (TypeError, ValueError):                     # paho >= 2.0 wants the callback API version first
                self._paho = mqtt.Client(mqtt.CallbackAPIVersion.VERSION1, client_id=self.mqtt_connection_id, clean_session=True)
            self._paho.on_connect = self.on_connect
            self._paho.on_message = self.on_message
            self._paho.on_publish, self._paho.on_disconnect = self.on_publish, self.on_disconnect
            if user is not None:
                self._paho.username_pw_set(user, pwd)
            self._paho.will_set("W/topic", payload=will, qos=0, retain=True)
            self._paho.connect(self._host, self._port, keepalive_time)
            self._paho.loop_start()                              # callbacks run on paho's network thread, like the native client
            self._native = None
        except ImportError:
            from ..core.distributed.communication.broker import BrokerClient
            self._paho = None
            self._native = BrokerClient(self._host, self._port, client_id=self.mqtt_connection_id, will_topic="W/topic", will_payload=will,
                                        keepalive=int(keepalive_time or 0), username=user, password=pwd)
            self._native.on_message = lambda topic, payload: self._dispatch(topic, payload)
            self.on_connect(self._native, None, None, 0)

    # ---- topics of this edge
    def control_topics(self):
        cid = self._client_id
        return [f"flserver_agent/{cid}/start_train", f"flserver_agent/{cid}/stop_train", f"fl_client/mlops/{cid}/status"]

    def on_connect(self, client, userdata, flags, rc):
        for topic in self.control_topics():
            client.subscribe(topic)
            logging.info("subscribe: %s", topic)

    def _dispatch(self, topic: str, payload: bytes):
        logging.info("on_message(%s, %r)", topic, payload[:200])
        listener = self._listeners.get(topic)
        if callable(listener):
            listener(topic, payload.decode("utf-8") if isinstance(payload, (bytes, bytearray)) else str(payload))

    def on_message(self, client, userdata, msg):
        self._dispatch(msg.topic, msg.payload)

    @staticmethod
    def on_publish(client, obj, mid):
        logging.info("on_publish mid=%s", mid)

    @staticmethod
    def on_disconnect(client, userdata, rc):
        logging.info("on_disconnect code=%s", rc)

    # ---- API
    def add_message_listener(self, topic, listener):
        self._listeners[topic] = listener
        if topic not in self.control_topics():
            (self._paho or self._native).subscribe(topic)

    def remove_message_listener(self, topic):
        del self._listeners[topic]

    def send_message(self, topic, message):
        if self._paho is not None:
            self._paho.publish(topic, payload=message, qos=2)
        else:
            self._native.publish(topic, message if isinstance(message, (bytes, bytearray)) else str(message).encode("utf-8"))

    def send_message_json(self, topic, message):
        logging.info(message)
        self.send_message(topic, message)

    def loop_forever(self):
        self._stopped.wait()

    def disconnect(self):
        self._stopped.set()
        if self._paho is not None:
            self._paho.loop_stop()
            self._paho.disconnect()
        else:
            self._native.disconnect()


def load_yaml_config(yaml_path):
    """Parsed YAML file (a dict); an unreadable file raises ``ValueError("Yaml error - check yaml file")`` like the reference."""
    with open(yaml_path) as stream:
        try:
            return yaml.safe_load(stream)
        except yaml.YAMLError as exc:
            raise ValueError("Yaml error - check yaml file") from exc


# ---- object store of the packages (an S3 bucket when boto3 and credentials exist, else FEDML_OBJECT_STORE_DIR / ~/.fedml_b200/store)
def _store():
    from ..arguments import Arguments
    from ..core.distributed.communication.remote_storage import create_store
    root = os.environ.get("FEDML_OBJECT_STORE_DIR", os.path.join(os.path.expanduser("~"), ".fedml_b200", "store"))
    return create_store(Arguments.from_dict({"object_store_dir": root, "s3_config_path": os.environ.get("FEDML_S3_CONFIG", "")}))


def upload_file(src_local_path, dest_s3_path) -> bool:
    try:
        _store().upload_file(src_local_path, dest_s3_path)
        return True
    except Exception as exc:  # noqa: BLE001 - the reference reports and returns False
        logging.error("Upload data failed. | src: %s | dest: %s | Exception: %s", src_local_path, dest_s3_path, exc)
        return False


def download_file(path_s3, path_local) -> None:
    _store().download_file(path_s3, path_local)


def delete_s3_zip(path_s3) -> None:
    _store().delete(path_s3)


# ---- login / logout
def _home() -> str:
    return os.path.join(os.path.expanduser("~"), "fedml-client")


def save_edge_infos(unique_device_id, edge_id, home: Optional[str] = None) -> str:
    """Remembers which edge this machine is bound as (``<home>/fedml/data/runner_infos/runner_infos.yaml``)."""
    d = os.path.join(home or _home(), "fedml", "data", "runner_infos")
    os.makedirs(d, exist_ok=True)
    path = os.path.join(d, "runner_infos.yaml")
    with open(path, "w") as fh:
        yaml.safe_dump({"unique_device_id": str(unique_device_id), "edge_id": edge_id}, fh)
    return path


def login(userid, version="release", binding_url: Optional[str] = None, broker: Optional[str] = None, block: bool = True):
    """Binds this device to the account and runs the agent loop (returns the runner when ``block`` is False)."""
    from .edge_agent import FedMLClientRunner, bind_account_and_device_id, get_device_id
    device_id = get_device_id()
    edge_id = bind_account_and_device_id(binding_url, userid, device_id) if binding_url else 0
    save_edge_infos(device_id, edge_id)
    runner = FedMLClientRunner(userid, broker or os.environ.get("FEDML_MQTT_BROKER", "127.0.0.1:1883"), version=version, binding_url=binding_url)
    if block:
        runner.loop()
    return runner


def logout(home: Optional[str] = None) -> bool:
    """Forgets the binding; True when there was one."""
    path = os.path.join(home or _home(), "fedml", "data", "runner_infos", "runner_infos.yaml")
    if os.path.exists(path):
        os.remove(path)
        return True
    return False


def login_with_docker_mode(userid, version="release", image: Optional[str] = None, run: bool = True):
    """The agent in a container: returns the docker command line; executes it only when docker is installed and ``run``."""
    sys_name = {"Darwin": "MacOS"}.get(platform.system(), platform.system())
    if sys_name == "Windows":
        raise RuntimeError("docker mode is not available on Windows")
    home = _home()
    os.makedirs(os.path.join(home, "fedml"), exist_ok=True)
    image = image or f"fedml-b200/client-agent:{version or 'release'}"
    cmd = ["docker", "run", "--rm", "--name", f"fedml-client-{userid}", "-v", f"{home}:/home/fedml/fedml-client",
           "--env", f"ACCOUNT_ID={userid}", "--env", f"FEDML_VERSION={version or 'release'}", "--env", f"CLIENT_OS_NAME={sys_name}",
           image]
    if run and shutil.which("docker"):
        subprocess.Popen(cmd)
    return cmd
