"""Five-line API: init, device, data, model, runner."""
import fedml
from fedml import FedMLRunner

if __name__ == "__main__":
    args = fedml.init()
    device = fedml.device.get_device(args)
    dataset, output_dim = fedml.data.load(args)
    model = fedml.model.create(args, output_dim)
    FedMLRunner(args, device, dataset, model).run()
