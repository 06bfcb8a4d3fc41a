"""the public API surface of the reference that scripts rely on, item by item."""
import inspect
import os


def test_public_api_checklist():
    import fedml
    assert isinstance(fedml.__version__, str)
    for fn in ("init", "run_simulation", "run_cross_silo_server", "run_cross_silo_client", "run_hierarchical_cross_silo_server",
               "run_hierarchical_cross_silo_client", "run_mnn_server", "run_distributed", "load_arguments"):
        assert callable(getattr(fedml, fn)), fn
    assert "backend" in inspect.signature(fedml.run_simulation).parameters
    assert list(inspect.signature(fedml.load_arguments).parameters)[:2] == ["training_type", "comm_backend"]
    assert inspect.isclass(fedml.ClientTrainer) and inspect.isclass(fedml.ServerAggregator) and inspect.isclass(fedml.FedMLRunner)
    assert callable(fedml.device.get_device) and callable(fedml.data.load) and callable(fedml.data.load_cross_silo)
    assert callable(fedml.model.create) and callable(fedml.model.create_mnn_lenet5_model) and callable(fedml.model.create_mnn_resnet20_model)
    for cls in ("SimulatorSingleProcess", "SimulatorMPI", "SimulatorNCCL"):
        params = list(inspect.signature(getattr(fedml.simulation, cls).__init__).parameters)
        assert params[1:5] == ["args", "device", "dataset", "model"], (cls, params)
    for cls in (fedml.cross_silo.Client, fedml.cross_silo.Server, fedml.cross_silo.hierarchical.Client, fedml.cross_silo.hierarchical.Server,
                fedml.cross_device.ServerMNN):
        assert "model_trainer" in inspect.signature(cls.__init__).parameters or "kw" in str(inspect.signature(cls.__init__)), cls
        assert callable(getattr(cls, "run"))
    assert fedml.logging.info is __import__("logging").info
    from fedml.core.distributed.client.client_manager import ClientManager
    from fedml.core.distributed.communication.message import Message
    from fedml.core.distributed.communication.observer import Observer
    from fedml.core.distributed.server.server_manager import ServerManager
    from fedml.core.distributed.topology.asymmetric_topology_manager import AsymmetricTopologyManager
    from fedml.core.distributed.topology.symmetric_topology_manager import SymmetricTopologyManager
    from fedml.core.non_iid_partition.noniid_partition import (non_iid_partition_with_dirichlet_distribution,
                                                               partition_class_samples_with_dirichlet_distribution, record_data_stats)
    from fedml.core.robustness.robust_aggregation import RobustAggregator
    for obj in (ClientManager, ServerManager, Message, Observer, SymmetricTopologyManager, AsymmetricTopologyManager, RobustAggregator,
                non_iid_partition_with_dirichlet_distribution, partition_class_samples_with_dirichlet_distribution, record_data_stats):
        assert obj is not None
    for m in ("register_message_receive_handler", "send_message", "run", "finish"):
        assert callable(getattr(ClientManager, m)) and callable(getattr(ServerManager, m))
    # custom-trainer contract
    for m in ("get_model_params", "set_model_params", "train", "test"):
        assert m in fedml.ClientTrainer.__abstractmethods__ or callable(getattr(fedml.ClientTrainer, m))
    assert callable(getattr(fedml.ClientTrainer, "test_on_the_server"))
    # CLI commands
    from fedml.cli.cli import cli
    assert {"version", "login", "logout", "build"} <= set(cli.commands)


def test_no_undefined_names_in_the_tree():
    """The reference's CI ran pyflakes; the equivalent check here (tools/check_names.py) needs no third-party linter."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "check_names.py")], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout
