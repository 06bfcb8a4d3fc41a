from .simulator import SimulatorMPI, SimulatorNCCL, SimulatorSingleProcess

__all__ = ["SimulatorSingleProcess", "SimulatorMPI", "SimulatorNCCL"]
