from .base_com_manager import BaseCommunicationManager
from .message import Message
from .observer import Observer

BACKENDS = ("INPROC", "MPI", "TCP", "GRPC", "TRPC", "MQTT", "MQTT_S3", "MQTT_S3_MNN")


def create_comm_manager(args, rank, size, backend, announce_ready=False, comm=None):
    """Backend-by-name factory shared by ClientManager / ServerManager
    (reference: client_manager.py:27-94, server_manager.py:27-96)."""
    b = str(backend).upper()
    run_id = getattr(args, "run_id", 0)
    if b == "INPROC":
        from .inproc import InProcCommManager
        return InProcCommManager(run_id, rank, size, announce_ready)
    if b == "MPI" and comm is not None and hasattr(comm, "Get_rank"):
        # a real MPI communicator (mpi4py, job started by mpirun): the reference's transport
        from .mpi_backend import MpiCommManager
        return MpiCommManager(comm, rank, size, announce_ready=announce_ready)
    if b in ("MPI", "TCP"):
        # no MPI communicator (mpi4py absent or not under mpirun): the same star topology over sockets
        from .tcp import TcpCommManager
        hosts = getattr(args, "host_table", None)
        return TcpCommManager(rank, size, hosts, int(getattr(args, "tcp_base_port", 50000)), announce_ready)
    if b == "GRPC":
        from .grpc_backend import GRPCCommManager
        base = int(getattr(args, "grpc_base_port", 8888))
        return GRPCCommManager("0.0.0.0", base + rank, getattr(args, "grpc_ipconfig_path", None), client_id=rank,
                               client_num=size - 1, base_port=base, announce_ready=announce_ready,
                               wire_format=getattr(args, "grpc_wire_format", "fedml_b200"),
                               allow_pickle=bool(getattr(args, "grpc_allow_pickle", False)))
    if b == "TRPC":
        from .trpc_backend import TRPCCommManager
        return TRPCCommManager(getattr(args, "trpc_master_config_path", None), rank, size, args, announce_ready)
    if b == "MQTT":
        from .mqtt_backends import MqttCommManager
        return MqttCommManager(args, client_rank=rank, client_num=size - 1)
    if b == "MQTT_S3":
        from .mqtt_backends import MqttS3MultiClientsCommManager
        return MqttS3MultiClientsCommManager(args, client_rank=rank, client_num=size - 1)
    if b == "MQTT_S3_MNN":
        from .mqtt_backends import MqttS3MNNCommManager
        return MqttS3MNNCommManager(args, client_rank=rank, client_num=size - 1)
    raise ValueError(f"unknown communication backend {backend!r}")
