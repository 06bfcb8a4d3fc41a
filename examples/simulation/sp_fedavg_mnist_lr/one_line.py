"""One-line API (reference: examples/simulation/sp_fedavg_mnist_lr_example, one-line variant)."""
import fedml

if __name__ == "__main__":
    fedml.run_simulation()
