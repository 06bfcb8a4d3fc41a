"""The B200 product path: the reference's ``backend: "NCCL"`` simulator (a stub there).  One process per GPU; the clients
that share a GPU train as one client-batched network on the tcgen05 kernels and a round ends in one fused aggregation kernel
over NVLink peer memory.  ``compute_dtype: tf32`` = the reference's precision (fp32 tensors, TF32 tensor-core products);
``bf16`` is ~1.5x faster."""
import fedml
from fedml.constants import FEDML_SIMULATION_TYPE_NCCL

if __name__ == "__main__":
    fedml.run_simulation(backend=FEDML_SIMULATION_TYPE_NCCL)
