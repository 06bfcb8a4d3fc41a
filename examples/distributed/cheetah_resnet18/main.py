"""Cheetah: data-parallel training (``fedml.run_distributed``, a ``pass`` in the reference).  Gradients live in one flat
peer-visible bucket; a step is one fused reduce-scatter + SGD + all-gather kernel over NVLink instead of NCCL + optimizer."""
import fedml

if __name__ == "__main__":
    fedml.run_distributed()
