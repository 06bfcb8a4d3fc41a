"""Edge agent process started by ``fedml login`` (reference: cli/edge_deployment/login.py FedMLClientRunner :31-560):
binds the account / device pair to an edge id at the MLOps service (:482-497), listens on the broker for start / stop
requests, fetches and unzips the run's package (http(s):// or file://, :119-140), rewrites its configuration with the
run's ``${FEDSYS.*}`` values (:142-204), starts the training process, reports status changes and ships the run log to the
log server while the run is alive."""
import argparse
import json
import os
import signal
import subprocess
import sys
import time

import platform
import shutil
import urllib.request
import uuid

from ..core.distributed.communication.broker import BrokerClient
from .cli import HOME, materialise_package


def get_device_id() -> str:
    """Stable id of this machine (the reference uses the MAC address / platform serial, login.py:441-480)."""
    return f"{uuid.getnode():012x}"


def bind_account_and_device_id(url: str, account_id, device_id: str, os_name: str = None) -> int:
    """POST ``{accountid, deviceid, type, gpu, processor, network}`` to ``<server>/fedmlOpsServer/edges/binding``; the
    answer ``{"code": "SUCCESS", "data": {"id": edge_id}}`` carries the edge id (0 when the service refuses)."""
    from ..core.mlops import _post_json
    params = {"accountid": account_id, "deviceid": device_id, "type": os_name or platform.system(), "gpu": "None",
              "processor": platform.processor(), "network": ""}
    resp = _post_json(url, params)
    if resp.get("code") != "SUCCESS":
        return 0
    return int((resp.get("data") or {}).get("id", 0))


def retrieve_package(package_url: str, dest_dir: str) -> str:
    """Download (http / https / file URL or plain path) into ``dest_dir``; returns the local zip path."""
    os.makedirs(dest_dir, exist_ok=True)
    if "://" not in package_url:
        package_url = "file://" + os.path.abspath(package_url)
    local = os.path.join(dest_dir, os.path.basename(package_url.split("?")[0]) or "package.zip")
    if package_url.startswith("file://"):
        shutil.copyfile(package_url[len("file://"):], local)
    else:
        urllib.request.urlretrieve(package_url, local)
    return local

STATUS = ("IDLE", "INITIALIZING", "TRAINING", "STOPPING", "FINISHED")


class FedMLClientRunner:
    def __init__(self, account_id, broker, version="release", binding_url=None, log_server_url=None):
        host, port = broker.split(":")
        self.account_id, self.device_id = account_id, get_device_id()
        self.edge_id = 0
        if binding_url:
            try:
                self.edge_id = bind_account_and_device_id(binding_url, account_id, self.device_id)
            except Exception as exc:  # noqa: BLE001 - the service is optional offline
                print(f"device binding at {binding_url} failed: {exc}", file=sys.stderr)
        if not self.edge_id:     # offline: the account id (or a stable hash of it) is the edge id
            import zlib
            self.edge_id = int(account_id) if str(account_id).isdigit() else zlib.crc32(str(account_id).encode()) % 100000
        self.version, self.log_server_url, self.log = version, log_server_url, None
        self.client = BrokerClient(host, int(port), client_id=f"agent-{self.edge_id}", will_topic="W/topic",
                                   will_payload={"ID": self.edge_id, "stat": "Offline"})
        self.client.on_message = self.on_message
        self.proc = None
        self.status = "IDLE"
        self.client.subscribe(f"flserver_agent/{self.edge_id}/start_train")
        self.client.subscribe(f"flserver_agent/{self.edge_id}/stop_train")
        self.report("IDLE")

    def report(self, status):
        self.status = status
        self.client.publish("fl_client/mlops/status", {"edge_id": self.edge_id, "status": status})
        self.client.publish(f"fl_client/mlops/{self.edge_id}/status", {"edge_id": self.edge_id, "status": status})

    def on_message(self, topic, payload):
        msg = json.loads(payload.decode() or "{}")
        if topic.endswith("start_train"):
            self.start_train(msg)
        elif topic.endswith("stop_train"):
            self.stop_train()

    def start_train(self, msg):
        self.report("INITIALIZING")
        run_id = msg.get("runId", 0)
        pkg = msg.get("run_config", {}).get("packages_config", {}).get("linuxClientUrl") or msg.get("package")
        edges = [e["id"] if isinstance(e, dict) else e for e in msg.get("edges", [self.edge_id])]
        fedsys = {"RUN_ID": run_id, "CLIENT_INDEX": edges.index(self.edge_id) + 1 if self.edge_id in edges else 1,
                  "CLIENT_ID_LIST": json.dumps(edges), "CLIENT_NUM": len(edges), "MQTT_CONFIG_PATH": "", "S3_CONFIG_PATH": "",
                  "LOG_FILE_DIR": os.path.join(HOME, "logs")}
        work = os.path.join(HOME, "fedml-package", f"run_{run_id}")
        local_zip = retrieve_package(pkg, os.path.join(HOME, "fedml_packages"))
        entry, conf, dyn = materialise_package(local_zip, work, fedsys)
        if self.log_server_url:
            from types import SimpleNamespace
            from ..core.mlops import MLOpsRuntimeLog
            self.log = MLOpsRuntimeLog(SimpleNamespace(run_id=run_id, rank=fedsys["CLIENT_INDEX"], log_server_url=self.log_server_url,
                                                       log_upload_interval=float(msg.get("log_upload_interval", 10.0))))
            self.log.edge_id = self.edge_id
            self.log.log_file_path = os.path.join(fedsys["LOG_FILE_DIR"], f"fedavg-cross-silo-run-{run_id}-edge-{self.edge_id}.log")
            self.log.start_uploader()
        self.proc = subprocess.Popen([sys.executable, entry, "--cf", conf, "--rank", str(dyn.get("rank", fedsys["CLIENT_INDEX"])),
                                      "--run_id", str(run_id)], start_new_session=True)
        self.report("TRAINING")

    def stop_train(self):
        self.report("STOPPING")
        if self.proc is not None and self.proc.poll() is None:
            os.killpg(os.getpgid(self.proc.pid), signal.SIGTERM)
        if self.log is not None:
            self.log.stop_uploader()
            self.log = None
        self.report("FINISHED")

    def loop(self):
        while True:
            if self.proc is not None and self.proc.poll() is not None and self.status == "TRAINING":
                self.report("FINISHED" if self.proc.returncode == 0 else "FAILED")
                self.proc = None
                if self.log is not None:
                    self.log.stop_uploader()
                    self.log = None
            time.sleep(0.5)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--account"); ap.add_argument("--version", default="release"); ap.add_argument("--broker", default="127.0.0.1:1883")
    ap.add_argument("--binding-url", default=None, help="MLOps edge-binding endpoint (.../fedmlOpsServer/edges/binding)")
    ap.add_argument("--log-server-url", default=None, help="where run logs are posted (.../fedmlLogsServer/logs/update)")
    a = ap.parse_args()
    FedMLClientRunner(a.account, a.broker, a.version, a.binding_url, a.log_server_url).loop()
