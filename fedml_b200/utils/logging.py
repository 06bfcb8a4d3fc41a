"""Logging set-up with the reference's line format (core/mlops/mlops_runtime_log.py:77-120)."""
import logging
import os


class LoggerCreator:
    @staticmethod
    def create_logger(name=None, level=logging.INFO, args=None):
        logger = logging.getLogger(name)
        logger.setLevel(level)
        return logger


def init_logs(args, role=None):
    rank = int(getattr(args, "rank", 0) or 0)
    role = role or ("Server" if rank == 0 else "Client")
    fmt = (f"[FedML-{role}({rank}) @device-id-{getattr(args, 'edge_id', rank)}] [%(asctime)s] [%(levelname)s] "
           "[%(filename)s:%(lineno)d:%(funcName)s] %(message)s")
    level = getattr(logging, str(getattr(args, "log_level", "INFO")).upper(), logging.INFO)
    root = logging.getLogger()
    for h in list(root.handlers):
        if getattr(h, "_fedml_b200", False):
            root.removeHandler(h)
    h = logging.StreamHandler(); h._fedml_b200 = True
    h.setFormatter(logging.Formatter(fmt, datefmt="%a, %d %b %Y %H:%M:%S"))
    root.addHandler(h)
    root.setLevel(level)
    if getattr(args, "using_mlops", False) or getattr(args, "log_to_file", False):
        d = getattr(args, "log_file_dir", "./log") or "./log"
        os.makedirs(d, exist_ok=True)
        path = os.path.join(d, f"fedavg-cross-silo-run-{getattr(args, 'run_id', 0)}-edge-{getattr(args, 'edge_id', rank)}.log")
        fh = logging.FileHandler(path)
        fh._fedml_b200 = True
        fh.setFormatter(logging.Formatter(fmt))
        root.addHandler(fh)
        return path
    return None


log_levels = {"debug": logging.DEBUG, "info": logging.INFO, "warning": logging.WARNING, "error": logging.ERROR, "critical": logging.CRITICAL}
logger = logging.getLogger("FedML")          # module-level logger of the reference (utils/logging.py:33)
