"""Worker-per-process simulator.  Without a launcher the ranks run as threads over the in-process transport; under
``run.sh`` every rank is its own process and the ranks talk over TCP (the mpirun replacement: mpi4py is used when importable)."""
import fedml
from fedml.constants import FEDML_SIMULATION_TYPE_MPI

if __name__ == "__main__":
    fedml.run_simulation(backend=FEDML_SIMULATION_TYPE_MPI)
