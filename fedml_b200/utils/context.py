"""Error-handling contexts (reference: utils/context.py:8-38, which aborts MPI_COMM_WORLD)."""
import logging
import threading
import traceback
from contextlib import contextmanager


@contextmanager
def raise_MPI_error(abort_fn=None):  # noqa: N802 - reference name
    """On an exception: log it, tear the job down (``torch.distributed`` process group instead of
    ``MPI.COMM_WORLD.Abort()``), re-raise."""
    try:
        yield
    except Exception:
        logging.error(traceback.format_exc())
        if abort_fn is not None:
            abort_fn()
        else:
            try:
                import torch.distributed as dist
                if dist.is_available() and dist.is_initialized():
                    dist.destroy_process_group()
            except Exception:
                pass
        raise


raise_error_and_teardown = raise_MPI_error

_locks = {}


@contextmanager
def get_lock(lock: threading.Lock):
    lock.acquire()
    try:
        yield
    finally:
        lock.release()
