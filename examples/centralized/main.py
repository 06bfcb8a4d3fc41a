"""Centralized baseline on the union of all clients' data (reference: examples/centralized/main.py): with one client in
total the data loader returns the whole training set and the trainer is plain SGD."""
import fedml
from fedml.centralized.centralized_trainer import CentralizedTrainer

if __name__ == "__main__":
    args = fedml.init()
    device = fedml.device.get_device(args)
    dataset, output_dim = fedml.data.load(args)
    model = fedml.model.create(args, output_dim)
    CentralizedTrainer(dataset, model, device, args).train()
