"""Message-handler managers (reference: core/distributed/client/client_manager.py:20-148,
server/server_manager.py:19-143): pick a transport by name, register ``msg_type -> handler``, run a blocking
receive loop that dispatches through the Observer interface, ``finish()`` stops the loop.  The reference's
``finish`` calls ``MPI.COMM_WORLD.Abort()``; here every backend stops cleanly."""
import logging
from abc import abstractmethod

from .communication import create_comm_manager
from .communication.message import Message
from .communication.observer import Observer


class _Manager(Observer):
    announce_ready = False

    def __init__(self, args, comm=None, rank=0, size=0, backend="MPI"):
        self.args, self.size, self.rank, self.backend = args, int(size), int(rank), backend
        self.comm = comm
        self.com_manager = comm if (comm is not None and hasattr(comm, "send_message")) else \
            create_comm_manager(args, self.rank, self.size, backend, self.announce_ready, comm=comm)
        self.com_manager.add_observer(self)
        self.message_handler_dict = {}
        self.com_manager_status = "INIT"

    def run(self):
        self.register_message_receive_handlers()
        self.com_manager_status = "RUNNING"
        self.com_manager.handle_receive_message()
        self.com_manager_status = "STOPPED"

    def get_sender_id(self):
        return self.rank

    def receive_message(self, msg_type, msg_params) -> None:
        handler = self.message_handler_dict.get(msg_type)
        if handler is None:
            logging.debug("rank %d: no handler for message type %s", self.rank, msg_type)
            return
        handler(msg_params)

    def send_message(self, message: Message):
        self.com_manager.send_message(message)

    @abstractmethod
    def register_message_receive_handlers(self) -> None:
        ...

    def register_message_receive_handler(self, msg_type, handler_callback_func):
        self.message_handler_dict[msg_type] = handler_callback_func

    def finish(self):
        logging.info("__finish rank %d", self.rank)
        self.com_manager.stop_receive_message()


class ClientManager(_Manager):
    pass


class ServerManager(_Manager):
    pass
