from .topology_managers import SymmetricTopologyManager  # reference import path
