"""String constants shared by the config loader, runners and algorithm registry.

Mirrors the vocabulary of the reference (python/fedml/constants.py:1-34) so that
existing ``fedml_config.yaml`` files keep working, and adds the names the reference
advertises but never wires (FedNova, robust aggregators, Cheetah).
"""

FEDML_DATA_CACHE_FOLDER = "fedml_data"
FEDML_TRAINING_PLATFORM_SIMULATION = "simulation"
FEDML_TRAINING_PLATFORM_CROSS_SILO = "cross_silo"
FEDML_TRAINING_PLATFORM_CROSS_DEVICE = "cross_device"
FEDML_TRAINING_PLATFORM_DISTRIBUTED = "distributed"

FEDML_CROSS_SILO_SCENARIO_HORIZONTAL = "horizontal"
FEDML_CROSS_SILO_SCENARIO_HIERARCHICAL = "hierarchical"

FEDML_SIMULATION_TYPE_SP = "single_process"
FEDML_SIMULATION_TYPE_MPI = "MPI"
FEDML_SIMULATION_TYPE_NCCL = "NCCL"

FedML_FEDERATED_OPTIMIZER_BASE_FRAMEWORK = "base_framework"
FedML_FEDERATED_OPTIMIZER_FEDAVG = "FedAvg"
FedML_FEDERATED_OPTIMIZER_FEDOPT = "FedOpt"
FedML_FEDERATED_OPTIMIZER_FEDPROX = "FedProx"
FedML_FEDERATED_OPTIMIZER_FEDNOVA = "FedNova"
FedML_FEDERATED_OPTIMIZER_CLASSICAL_VFL = "classical_vertical"
FedML_FEDERATED_OPTIMIZER_SPLIT_NN = "split_nn"
FedML_FEDERATED_OPTIMIZER_DECENTRALIZED_FL = "decentralized_fl"
FedML_FEDERATED_OPTIMIZER_FEDGAN = "FedGAN"
FedML_FEDERATED_OPTIMIZER_FEDAVG_ROBUST = "FedAvg_robust"
FedML_FEDERATED_OPTIMIZER_FEDGKT = "FedGKT"
FedML_FEDERATED_OPTIMIZER_FEDNAS = "FedNAS"
FedML_FEDERATED_OPTIMIZER_FEDSEG = "FedSeg"
FedML_FEDERATED_OPTIMIZER_TURBO_AGGREGATE = "turbo_aggregate"
FedML_FEDERATED_OPTIMIZER_HIERARCHICAL_FL = "HierarchicalFL"
FedML_FEDERATED_OPTIMIZER_S_FEDAVG = "S_FedAvg"
FedML_FEDERATED_OPTIMIZER_HS_FEDAVG = "HS_FedAvg"

FEDML_DATA_MNIST_URL = "https://fedcv.s3.us-west-1.amazonaws.com/MNIST.zip"


def canonical_optimizer_name(name: str) -> str:
    """Case/underscore-insensitive lookup of a federated optimizer name.

    The reference compares strings exactly, which is how its own robust example YAML
    ("fedavg_robust") fails against the constant "FedAvg_robust" (SURVEY §0.3 item 8).
    """
    table = {
        v.lower().replace("_", ""): v
        for k, v in globals().items()
        if k.startswith("FedML_FEDERATED_OPTIMIZER_")
    }
    key = str(name).lower().replace("_", "").replace("-", "")
    aliases = {"fedadam": "FedOpt", "fedyogi": "FedOpt", "fedadagrad": "FedOpt",
               "hierarchical": "HierarchicalFL", "hierarchicalfl": "HierarchicalFL",
               "decentralized": "decentralized_fl", "vfl": "classical_vertical",
               "turboaggregate": "turbo_aggregate", "splitnn": "split_nn"}
    if key in aliases:
        return aliases[key]
    return table.get(key, name)
