"""Your own data and model: build the reference's 8-tuple dataset and any ``nn.Module``, hand both to the runner."""
import torch

import fedml
from fedml import FedMLRunner


def load_data(args):
    """Two blobs per client; returns the 8-tuple ``[train_num, test_num, train_global, test_global, local_num_dict,
    train_local_dict, test_local_dict, class_num]`` the reference's loaders produce (data/data_loader.py:59-81)."""
    g = torch.Generator().manual_seed(0)
    n_clients, per, dim, classes = int(args.client_num_in_total), 40, 20, 4
    centers = torch.randn(classes, dim, generator=g) * 3

    def batches(n):
        y = torch.randint(0, classes, (n,), generator=g)
        x = centers[y] + torch.randn(n, dim, generator=g)
        bs = int(args.batch_size)
        return [(x[i:i + bs], y[i:i + bs]) for i in range(0, n, bs)]

    train_local = {c: batches(per) for c in range(n_clients)}
    test_local = {c: batches(per // 4) for c in range(n_clients)}
    flat = lambda d: [b for c in d for b in d[c]]      # noqa: E731
    nums = {c: per for c in range(n_clients)}
    return [per * n_clients, per // 4 * n_clients, flat(train_local), flat(test_local), nums, train_local, test_local, classes], classes


class TwoLayer(torch.nn.Module):
    def __init__(self, dim, classes):
        super().__init__()
        self.net = torch.nn.Sequential(torch.nn.Linear(dim, 32), torch.nn.ReLU(), torch.nn.Linear(32, classes))

    def forward(self, x):
        return self.net(x)


if __name__ == "__main__":
    args = fedml.init()
    args.client_num_in_total, args.client_num_per_round = 8, 4
    device = fedml.device.get_device(args)
    dataset, classes = load_data(args)
    FedMLRunner(args, device, dataset, TwoLayer(20, classes)).run()
