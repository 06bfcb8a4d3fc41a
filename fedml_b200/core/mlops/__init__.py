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


class MLOpsRuntimeLog:
    _instance = None

    def __init__(self, args):
        self.args = args

    @classmethod
    def get_instance(cls, args):
        if cls._instance is None or cls._instance.args is not args:
            cls._instance = cls(args)
        return cls._instance

    def init_logs(self):
        from ...utils.logging import init_logs
        init_logs(self.args)


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

    def fetch_configs(self):
        import os
        import yaml
        out = []
        for key in ("mqtt_config_path", "s3_config_path"):
            p = getattr(self.args, key, None)
            if p and os.path.exists(str(p)):
                with open(p) as fh:
                    out.append(yaml.safe_load(fh) or {})
            else:
                out.append({})
        return tuple(out)
