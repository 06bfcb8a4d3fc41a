"""MLOps observability singletons (reference: core/mlops/*.py, 714 LoC).

* ``MLOpsProfilerEvent`` — ``log_event_started/ended(event_name, event_value, edge_id)`` publishing
  ``{run_id, edge_id, event_name, event_value, started_time|ended_time}`` on ``/mlops/events``
  (mlops_profiler_event.py:35-101; the reference sends "ended" twice, :77-78 — once here).
* ``MLOpsMetrics`` — client/server status FSM topics, training progress, model URLs, system performance
  (mlops_metrics.py:31-140).
* ``MLOpsRuntimeLog`` — the log format + optional file sink (mlops_runtime_log.py:70-160); the HTTP
  uploader to open.fedml.ai cannot work offline and is replaced by a local rolling file.
* ``SysStats`` — cpu / memory / disk / net / GPU utilisation via psutil + pynvml (system_stats.py:8-165).
All sinks are no-ops unless ``args.using_mlops`` is true; the messenger is any object with
``send_message_json(topic, json_str)`` (``MqttS3StatusManager``) — or an in-memory list for tests.
"""
import json
import logging
import time


class _ListMessenger:
    def __init__(self):
        self.sent = []

    def send_message_json(self, topic, payload):
        self.sent.append((topic, payload))


class MLOpsProfilerEvent:
    EVENT_TYPE_STARTED, EVENT_TYPE_ENDED = 0, 1
    _instance = None

    def __init__(self, args=None):
        self.args = args
        self.run_id = getattr(args, "run_id", 0) if args is not None else 0
        self.edge_id = getattr(args, "edge_id", 0) if args is not None else 0
        self.enabled = bool(getattr(args, "using_mlops", False)) if args is not None else False
        self.messenger = _ListMessenger()
        self.open_spans = {}

    @classmethod
    def get_instance(cls, args=None):
        if cls._instance is None or (args is not None and cls._instance.args is not args):
            cls._instance = cls(args)
        return cls._instance

    def set_messenger(self, messenger):
        self.messenger = messenger

    def _emit(self, kind, name, value, edge_id):
        msg = {"run_id": self.run_id, "edge_id": self.edge_id if edge_id is None else edge_id, "event_name": name,
               "event_value": value, ("started_time" if kind == 0 else "ended_time"): int(time.time())}
        if self.enabled:
            self.messenger.send_message_json("/mlops/events", json.dumps(msg))
        return msg

    def log_event_started(self, event_name, event_value=None, event_edge_id=None):
        self.open_spans[(event_name, str(event_value))] = time.perf_counter()
        return self._emit(0, event_name, "" if event_value is None else str(event_value), event_edge_id)

    def log_event_ended(self, event_name, event_value=None, event_edge_id=None):
        t0 = self.open_spans.pop((event_name, str(event_value)), None)
        if t0 is not None:
            logging.debug("event %s took %.4fs", event_name, time.perf_counter() - t0)
        return self._emit(1, event_name, "" if event_value is None else str(event_value), event_edge_id)


class MLOpsMetrics:
    _instance = None
    STATUS = ("IDLE", "INITIALIZING", "TRAINING", "STOPPING", "FINISHED", "OFFLINE", "FAILED", "RUNNING", "KILLED")

    def __init__(self, args=None):
        self.args = args
        self.run_id = getattr(args, "run_id", 0) if args is not None else 0
        self.edge_id = getattr(args, "edge_id", 0) if args is not None else 0
        self.enabled = bool(getattr(args, "using_mlops", False)) if args is not None else False
        self.messenger = _ListMessenger()

    @classmethod
    def get_instance(cls, args=None):
        if cls._instance is None or (args is not None and cls._instance.args is not args):
            cls._instance = cls(args)
        return cls._instance

    def set_messenger(self, messenger):
        self.messenger = messenger

    def _pub(self, topic, obj):
        if self.enabled:
            self.messenger.send_message_json(topic, json.dumps(obj))
        return topic, obj

    def report_client_training_status(self, edge_id, status):
        self._pub("fl_client/mlops/status", {"edge_id": edge_id, "status": status})
        return self._pub(f"fl_client/mlops/{edge_id}/status", {"edge_id": edge_id, "status": status})

    def report_server_training_status(self, run_id, status):
        self._pub("fl_server/mlops/status", {"run_id": run_id, "status": status})
        return self._pub("fl_server/mlops/id/status", {"run_id": run_id, "status": status})

    def report_client_training_metric(self, metric_json):
        return self._pub("fl_client/mlops/training_metrics", metric_json)

    def report_server_training_metric(self, metric_json):
        return self._pub("fl_server/mlops/training_progress_and_eval", metric_json)

    def report_server_training_round_info(self, round_info):
        return self._pub("fl_client/mlops/training_roundx", round_info)

    def report_client_model_info(self, model_info_json):
        return self._pub("fl_server/mlops/client_model", model_info_json)

    def report_aggregated_model_info(self, model_info_json):
        return self._pub("fl_server/mlops/global_aggregated_model", model_info_json)

    def report_system_metric(self, metric_json=None):
        return self._pub("fl_client/mlops/system_performance", metric_json or SysStats().produce_info(self.run_id, self.edge_id))


class SysStats:
    def produce_info(self, run_id=0, edge_id=0):
        info = {"run_id": run_id, "edge_id": edge_id, "timestamp": time.time()}
        try:
            import psutil
            vm, du, net = psutil.virtual_memory(), psutil.disk_usage("/"), psutil.net_io_counters()
            info.update(cpu_utilization=psutil.cpu_percent(interval=None), SystemMemoryUtilization=vm.percent,
                        process_memory_in_use=psutil.Process().memory_info().rss / 2 ** 20,
                        disk_utilization=du.percent, network_traffic=net.bytes_sent + net.bytes_recv)
        except Exception:  # noqa: BLE001
            pass
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(0)
            u = pynvml.nvmlDeviceGetUtilizationRates(h)
            info.update(gpu_utilization=u.gpu, gpu_memory_allocated=u.memory,
                        gpu_temp=pynvml.nvmlDeviceGetTemperature(h, pynvml.NVML_TEMPERATURE_GPU),
                        gpu_power_usage=pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0)
        except Exception:  # noqa: BLE001
            pass
        return info


CA_BUNDLE = None      # path of a certificate bundle for https:// endpoints (the reference ships one per platform version under
#                       core/mlops/ssl/ and points requests at it: mlops_configs.py:50-56); None = the system trust store


def set_ca_bundle(path):
    """``args.mlops_ca_bundle`` / ``FEDML_CA_BUNDLE``: verify the MLOps https endpoints against this bundle."""
    global CA_BUNDLE
    CA_BUNDLE = path or None


def _post_json(url: str, obj: dict, timeout: float = 10.0) -> dict:
    """POST a JSON document; ``file://path`` appends it as one line to a local file (offline sink)."""
    import json
    if url.startswith("file://"):
        path = url[len("file://"):]
        import os
        os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
        with open(path, "a") as fh:
            fh.write(json.dumps(obj) + "\n")
        return {"code": "SUCCESS"}
    import urllib.request
    req = urllib.request.Request(url, data=json.dumps(obj).encode(), headers={"Content-Type": "application/json", "Connection": "close"})
    ctx = None
    import os as _os
    bundle = CA_BUNDLE or _os.environ.get("FEDML_CA_BUNDLE")
    if url.startswith("https://") and bundle:
        import ssl
        ctx = ssl.create_default_context(cafile=bundle)
    with urllib.request.urlopen(req, timeout=timeout, context=ctx) as resp:
        body = resp.read().decode() or "{}"
        if resp.status != 200:
            raise OSError(f"HTTP {resp.status}")
    try:
        return json.loads(body)
    except ValueError:
        return {}


class MLOpsRuntimeLog:
    """Run log + its background uploader (reference: core/mlops/mlops_runtime_log.py:29-200): the log file of the run is
    tailed by a daemon thread that every ``log_upload_interval`` seconds (10 in the reference) posts the new lines as
    ``{run_id, edge_id, logs, create_time, update_time, created_by, updated_by}`` to ``args.log_server_url``; a failed post
    is retried with the same lines at the next tick."""
    _instance = None

    def __init__(self, args):
        self.args = args
        self.run_id = getattr(args, "run_id", 0)
        self.edge_id = self._edge_id(args)
        self.log_server_url = getattr(args, "log_server_url", None)
        self.log_line_index = 0
        self._thread = None
        self._stop = None
        self.log_file_path = None

    @staticmethod
    def _edge_id(args):
        rank = int(getattr(args, "rank", 0) or 0)
        if rank == 0:
            return 0
        try:
            ids = args.client_ids() if hasattr(args, "client_ids") else []
            return ids[rank - 1] if rank - 1 < len(ids) else rank
        except Exception:  # noqa: BLE001
            return rank

    @classmethod
    def get_instance(cls, args):
        if cls._instance is None or cls._instance.args is not args:
            cls._instance = cls(args)
        return cls._instance

    def init_logs(self):
        from ...utils.logging import init_logs
        self.log_file_path = init_logs(self.args)
        if self.log_server_url and bool(getattr(self.args, "using_mlops", False)):
            self.start_uploader()
        return self.log_file_path

    # ---- uploader
    def log_read(self):
        import os
        if not self.log_file_path or not os.path.exists(self.log_file_path):
            return []
        with open(self.log_file_path, errors="replace") as fh:
            lines = fh.readlines()
        return lines[self.log_line_index:]

    def log_upload(self, run_id=None, edge_id=None) -> int:
        """One upload tick; returns the number of lines the server accepted."""
        import time
        lines = self.log_read()
        if not lines or not self.log_server_url:
            return 0
        now = time.time()
        eid = self.edge_id if edge_id is None else edge_id
        req = {"run_id": self.run_id if run_id is None else run_id, "edge_id": eid, "logs": lines, "create_time": now,
               "update_time": now, "created_by": str(eid), "updated_by": str(eid)}
        try:
            _post_json(self.log_server_url, req)
        except Exception:  # noqa: BLE001 - keep the lines for the next tick
            return 0
        self.log_line_index += len(lines)
        return len(lines)

    def start_uploader(self, interval: float = None):
        import threading
        if self._thread is not None:
            return
        self._stop = threading.Event()
        every = float(interval if interval is not None else getattr(self.args, "log_upload_interval", 10.0))

        def loop():
            while not self._stop.wait(every):
                self.log_upload()
            self.log_upload()                       # final flush

        self._thread = threading.Thread(target=loop, daemon=True, name="mlops-log-upload")
        self._thread.start()

    def stop_uploader(self):
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=15)
            self._thread = None


class MLOpsConfigs:
    """Broker / S3 credentials.  The reference fetches them from open.fedml.ai
    (mlops_configs.py:29-68); offline they come from ``mqtt_config_path`` / ``s3_config_path`` YAMLs."""
    _instance = None

    def __init__(self, args):
        self.args = args

    @classmethod
    def get_instance(cls, args):
        cls._instance = cls(args)
        return cls._instance

    CONFIG_URLS = {"release": "https://open.fedml.ai/fedmlOpsServer/configs/fetch",
                   "test": "http://open-test.fedml.ai/fedmlOpsServer/configs/fetch",
                   "dev": "http://open-dev.fedml.ai/fedmlOpsServer/configs/fetch",
                   "local": "http://localhost:9000/fedmlOpsServer/configs/fetch"}

    def fetch_remote(self, url=None):
        """(mqtt_config, s3_config) from the MLOps config service (mlops_configs.py:29-68): POST
        ``{"config_name": ["mqtt_config", "s3_config"]}``, answer ``{"code": "SUCCESS", "data": {...}}``."""
        url = url or getattr(self.args, "mlops_config_url", None) or \
            self.CONFIG_URLS.get(str(getattr(self.args, "config_version", "release")), self.CONFIG_URLS["release"])
        resp = _post_json(url, {"config_name": ["mqtt_config", "s3_config"]})
        if resp.get("code") != "SUCCESS":
            raise RuntimeError(f"config service at {url} answered {resp.get('code')!r}")
        data = resp.get("data", {})
        return data.get("mqtt_config", {}), data.get("s3_config", {})

    def fetch_configs(self):
        import os
        import yaml
        if getattr(self.args, "mlops_config_url", None):
            return self.fetch_remote()
        out = []
        for key in ("mqtt_config_path", "s3_config_path"):
            p = getattr(self.args, key, None)
            if p and os.path.exists(str(p)):
                with open(p) as fh:
                    out.append(yaml.safe_load(fh) or {})
            else:
                out.append({})
        return tuple(out)
