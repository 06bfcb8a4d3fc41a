"""Per-dataset loader functions with the reference's names, signatures and return tuples.

Scripts written against the reference often skip ``fedml.data.load(args)`` and call the dataset's own loader, e.g.
``load_partition_data_cifar10(dataset, data_dir, partition_method, partition_alpha, client_number, batch_size)``
(data/cifar10/data_loader.py:312) or ``load_partition_data_mnist(batch_size, train_path, test_path)``
(data/MNIST/data_loader.py:109).  All of them are thin fronts of :func:`fedml_b200.data.load`, so they read the real files when
``data_dir`` holds them and otherwise produce the synthetic set of the same shape.  Return conventions:

* partitioned image sets (CIFAR-10/100, CINIC-10, ImageNet, Landmarks): ``(train_num, test_num, train_global, test_global,
  local_num_dict, train_local_dict, test_local_dict, class_num)``;
* natural federated sets (MNIST, FEMNIST, fed_cifar100, Shakespeare, StackOverflow, synthetic): the same with ``client_num`` in
  front;
* ``load_partition_data_distributed_*`` (one process per client): ``(train_num, train_global, test_global, local_num,
  train_local, test_local, class_num)`` - globals on process 0, the client's loaders on process ``i > 0``.
"""
from __future__ import annotations

from ..arguments import Arguments

SYNTHETIC_TRAIN_SIZE = 0          # > 0: cap on the size of the synthetic stand-in (tests, demos); real files are never truncated


def _load(dataset, data_dir=None, batch_size=10, partition_method="hetero", partition_alpha=0.5, client_number=None, **extra):
    from . import load
    kw = dict(dataset=dataset, batch_size=int(batch_size), partition_method=partition_method or "hetero",
              partition_alpha=0.5 if partition_alpha is None else float(partition_alpha), log_level="WARNING")
    if data_dir:
        kw["data_cache_dir"] = data_dir
    if client_number:
        kw["client_num_in_total"] = int(client_number)
        kw.setdefault("client_num_per_round", min(int(client_number), 10))
    if SYNTHETIC_TRAIN_SIZE:
        kw["synthetic_train_size"] = SYNTHETIC_TRAIN_SIZE
    kw.update(extra)
    args = Arguments.from_dict(kw)
    ds, class_num = load(args)
    return args, ds, class_num


def _eight(ds, class_num):
    return tuple(ds[:7]) + (class_num,)


def _nine(args, ds, class_num):
    return (int(args.client_num_in_total),) + _eight(ds, class_num)


def _distributed(process_id, ds, class_num):
    train_num, _, train_g, test_g, nums, train_l, test_l = ds[:7]
    if int(process_id) == 0:
        return train_num, train_g, test_g, 0, None, None, class_num
    c = int(process_id) - 1
    return train_num, None, None, nums[c], train_l[c], test_l[c], class_num


# ---- partitioned image sets --------------------------------------------------------------------------------------------------
def _partitioned(name):
    def loader(dataset, data_dir, partition_method, partition_alpha, client_number, batch_size, n_proc_in_silo=0):
        _, ds, k = _load(name, data_dir, batch_size, partition_method, partition_alpha, client_number)
        return _eight(ds, k)

    def distributed(process_id, dataset, data_dir, partition_method, partition_alpha, client_number, batch_size):
        _, ds, k = _load(name, data_dir, batch_size, partition_method, partition_alpha, client_number)
        return _distributed(process_id, ds, k)

    loader.__name__, distributed.__name__ = f"load_partition_data_{name}", f"load_partition_data_distributed_{name}"
    return loader, distributed


load_partition_data_cifar10, load_partition_data_distributed_cifar10 = _partitioned("cifar10")
load_partition_data_cifar100, load_partition_data_distributed_cifar100 = _partitioned("cifar100")
load_partition_data_cinic10, load_partition_data_distributed_cinic10 = _partitioned("cinic10")


def load_partition_data_ImageNet(dataset, data_dir, partition_method=None, partition_alpha=None, client_number=100, batch_size=10):  # noqa: N802
    _, ds, k = _load("ILSVRC2012", data_dir, batch_size, partition_method or "homo", partition_alpha, client_number)
    return _eight(ds, k)


def load_partition_data_landmarks(dataset, data_dir, fed_train_map_file=None, fed_test_map_file=None, partition_method=None,
                                  partition_alpha=None, client_number=233, batch_size=10):
    _, ds, k = _load(dataset if str(dataset).startswith("gld") else "gld23k", data_dir, batch_size, partition_method, partition_alpha)
    return _eight(ds, k)


# ---- natural federated sets -----------------------------------------------------------------------------------------------------
def load_partition_data_mnist(batch_size, train_path="./MNIST/train", test_path="./MNIST/test"):
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(train_path)))       # <root>/MNIST/train -> <root>
    a, ds, k = _load("mnist", root, batch_size)
    return _nine(a, ds, k)


def load_partition_data_mnist_by_device_id(batch_size, device_id, train_path="MNIST_mobile", test_path="MNIST_mobile"):
    import os
    return load_partition_data_mnist(batch_size, os.path.join(train_path, str(device_id), "train"), os.path.join(test_path, str(device_id), "test"))


def _natural(name, fn_name):
    def loader(dataset=None, data_dir=None, batch_size=20):
        a, ds, k = _load(name, data_dir, batch_size)
        return _nine(a, ds, k)

    def distributed(process_id, dataset=None, data_dir=None, batch_size=20):
        _, ds, k = _load(name, data_dir, batch_size)
        return _distributed(process_id, ds, k)

    loader.__name__, distributed.__name__ = f"load_partition_data_{fn_name}", f"load_partition_data_distributed_{fn_name}"
    return loader, distributed


load_partition_data_federated_emnist, load_partition_data_distributed_federated_emnist = _natural("femnist", "federated_emnist")
load_partition_data_federated_cifar100, load_partition_data_distributed_federated_cifar100 = _natural("fed_cifar100", "federated_cifar100")
load_partition_data_federated_shakespeare, load_partition_data_distributed_federated_shakespeare = _natural(
    "fed_shakespeare", "federated_shakespeare")
load_partition_data_federated_stackoverflow_lr, load_partition_data_distributed_federated_stackoverflow_lr = _natural(
    "stackoverflow_lr", "federated_stackoverflow_lr")
load_partition_data_federated_stackoverflow_nwp, load_partition_data_distributed_federated_stackoverflow_nwp = _natural(
    "stackoverflow_nwp", "federated_stackoverflow_nwp")


def load_partition_data_shakespeare(batch_size):
    a, ds, k = _load("shakespeare", None, batch_size)
    return _nine(a, ds, k)


def load_partition_data_federated_synthetic_1_1(data_dir=None, batch_size=10):
    a, ds, k = _load("synthetic_1_1", data_dir, batch_size)
    return _nine(a, ds, k)
