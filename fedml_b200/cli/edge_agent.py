"""Edge agent process started by ``fedml login`` (reference: cli/edge_deployment/login.py FedMLClientRunner :31-440)."""
import argparse
import json
import os
import signal
import subprocess
import sys
import time

from ..core.distributed.communication.broker import BrokerClient
from .cli import HOME, materialise_package

STATUS = ("IDLE", "INITIALIZING", "TRAINING", "STOPPING", "FINISHED")


class FedMLClientRunner:
    def __init__(self, account_id, broker, version="release"):
        host, port = broker.split(":")
        self.edge_id = int(account_id) if str(account_id).isdigit() else abs(hash(account_id)) % 100000
        self.version = version
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
        entry, conf, dyn = materialise_package(pkg.replace("file://", ""), work, fedsys)
        self.proc = subprocess.Popen([sys.executable, entry, "--cf", conf, "--rank", str(dyn.get("rank", fedsys["CLIENT_INDEX"])),
                                      "--run_id", str(run_id)], start_new_session=True)
        self.report("TRAINING")

    def stop_train(self):
        self.report("STOPPING")
        if self.proc is not None and self.proc.poll() is None:
            os.killpg(os.getpgid(self.proc.pid), signal.SIGTERM)
        self.report("FINISHED")

    def loop(self):
        while True:
            if self.proc is not None and self.proc.poll() is not None and self.status == "TRAINING":
                self.report("FINISHED" if self.proc.returncode == 0 else "FAILED")
                self.proc = None
            time.sleep(0.5)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--account"); ap.add_argument("--version", default="release"); ap.add_argument("--broker", default="127.0.0.1:1883")
    a = ap.parse_args()
    FedMLClientRunner(a.account, a.broker, a.version).loop()
