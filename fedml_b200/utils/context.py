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


@contextmanager
def raise_error_without_process():
    """Logs an exception of the block with its traceback and swallows it (reference utils/context.py:21-30: the variant used
    where no MPI world exists to abort)."""
    try:
        yield
    except Exception as exc:  # noqa: BLE001
        logging.info("%s", exc)
        logging.info("traceback.format_exc():\n%s", traceback.format_exc())


def post_complete_message_to_sweep_process(args, pipe_path: str = "./tmp/fedml"):
    """Tells a hyper-parameter sweep driver that the run is over: one line into the named pipe the reference's sweep scripts read
    (cross_silo/horizontal/utils.py:21-29).  Without a reader on the pipe nothing is written (the reference blocks forever there)."""
    import os
    os.makedirs(os.path.dirname(pipe_path) or ".", exist_ok=True)
    if not os.path.exists(pipe_path):
        os.mkfifo(pipe_path)
    try:
        fd = os.open(pipe_path, os.O_WRONLY | os.O_NONBLOCK)
    except OSError:                      # no reader: ENXIO
        return False
    with os.fdopen(fd, "w") as pipe:
        pipe.write("training is finished! \n%s\n" % (str(args)))
    return True
