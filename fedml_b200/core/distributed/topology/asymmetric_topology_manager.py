from .topology_managers import AsymmetricTopologyManager  # reference import path
