"""``fedml.data`` — dataset entry points.

``load(args) -> (dataset8, class_num)`` keeps the reference contract
(data/data_loader.py:29-322): ``dataset8 = [train_num, test_num, train_global, test_global,
train_local_num_dict, train_local_dict, test_local_dict, class_num]`` and side effects on
``args`` (``client_num_in_total`` is overwritten for naturally-partitioned sets).  The packed
:class:`FederatedDataset` behind the tuple is attached as ``args._federated_dataset`` /
returned by :func:`load_federated` for the device-resident simulators.
"""
from __future__ import annotations

import json
import logging
import os
import pickle

import numpy as np
import torch

from . import synthetic as S
from .federated_dataset import FederatedDataset, combine_batches, merge_to_single_client

__all__ = ["load", "load_federated", "load_cross_silo", "FederatedDataset", "split_data_for_dist_trainers",
           "load_poisoned_dataset_from_edge_case_examples"]

_NATURAL = {"mnist", "femnist", "fed_cifar100", "shakespeare", "fed_shakespeare", "stackoverflow_lr",
            "stackoverflow_nwp", "gld23k", "gld160k", "synthetic_1_1", "synthetic_0_0", "synthetic_0.5_0.5",
            "synthetic_iid"}


def _try_real_mnist(cache_dir):
    """LEAF JSON ``{users, user_data{x,y}}`` if the files are present (MNIST/data_loader.py:32-72)."""
    tr, te = os.path.join(cache_dir, "MNIST", "train"), os.path.join(cache_dir, "MNIST", "test")
    if not (os.path.isdir(tr) and os.path.isdir(te)):
        return None

    def read(d):
        users, data = [], {}
        for f in sorted(os.listdir(d)):
            if f.endswith(".json"):
                with open(os.path.join(d, f)) as fh:
                    blob = json.load(fh)
                users.extend(blob["users"]); data.update(blob["user_data"])
        return users, data

    users, trd = read(tr)
    _, ted = read(te)
    xs = [np.asarray(trd[u]["x"], dtype=np.float32) for u in users]
    ys = [np.asarray(trd[u]["y"], dtype=np.int64) for u in users]
    txs = [np.asarray(ted[u]["x"], dtype=np.float32) for u in users]
    tys = [np.asarray(ted[u]["y"], dtype=np.int64) for u in users]
    off = lambda a: np.concatenate([[0], np.cumsum([len(v) for v in a])])
    cat = lambda a: torch.from_numpy(np.concatenate(a))
    return FederatedDataset(cat(xs), cat(ys), off(ys), cat(txs), cat(tys), off(tys), 10, None, "mnist")


def _try_real_cifar10(cache_dir):
    d = os.path.join(cache_dir, "cifar-10-batches-py")
    if not os.path.isdir(d):
        return None
    def rd(name):
        with open(os.path.join(d, name), "rb") as fh:
            blob = pickle.load(fh, encoding="bytes")
        x = np.asarray(blob[b"data"], dtype=np.uint8).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)
        return x, np.asarray(blob[b"labels"], dtype=np.int64)
    parts = [rd(f"data_batch_{i}") for i in range(1, 6)]
    xtr = np.concatenate([p[0] for p in parts]); ytr = np.concatenate([p[1] for p in parts])
    xte, yte = rd("test_batch")
    return tuple(torch.from_numpy(np.ascontiguousarray(a)) for a in (xtr, ytr, xte, yte))


def load_federated(args) -> FederatedDataset:
    name = str(args.dataset).lower()
    seed = int(getattr(args, "random_seed", 0))
    n_clients = int(getattr(args, "client_num_in_total", 10))
    pm, pa = getattr(args, "partition_method", "homo"), float(getattr(args, "partition_alpha", 0.5))
    cache = getattr(args, "data_cache_dir", "./data")
    small = int(getattr(args, "synthetic_train_size", 0) or 0)
    if name == "mnist":
        ds = None if getattr(args, "synthetic_data", True) is True and not os.path.isdir(os.path.join(cache, "MNIST")) \
            else _try_real_mnist(cache)
        if ds is None:
            nat = int(getattr(args, "mnist_client_num", 1000))
            ds = S.make_mnist_like(client_num=nat, total_train=small or 61664, seed=seed)
    elif name == "femnist":
        from . import readers as R
        ds = R.load_leaf_images(cache, "femnist", 62, "femnist", shape=(28, 28)) \
            or R.load_tff_h5(os.path.join(cache, "fed_emnist_train.h5"), os.path.join(cache, "fed_emnist_test.h5"),
                             "pixels", "label", 62, "femnist")
        if ds is None:
            ds = S.make_femnist_like(client_num=int(getattr(args, "femnist_client_num", 3400)), seed=seed,
                                     total_train=small or 340000)
    elif name in ("cifar10", "cifar100", "cinic10"):
        cnum = 100 if name == "cifar100" else 10
        mean, std = (S.CIFAR100_MEAN, S.CIFAR100_STD) if name == "cifar100" else (S.CIFAR10_MEAN, S.CIFAR10_STD)
        ntr = small or (90000 if name == "cinic10" else 50000)
        nte = max(100, ntr // 5)
        real = _try_real_cifar10(cache) if name == "cifar10" else None
        if name == "cifar100":
            from . import readers as R
            real = R.load_cifar100_pickles(cache)
        elif name == "cinic10":
            from . import readers as R
            folder = R.load_image_folder(os.path.join(cache, "cinic10"), 32, "train", "test")
            real = folder[:4] if folder is not None else None
        if real is not None:
            from ..core.non_iid_partition import homo_partition, non_iid_partition_with_dirichlet_distribution
            xtr, ytr, xte, yte = real
            st = np.random.get_state(); np.random.seed(10)
            part = homo_partition(len(ytr), n_clients) if pm == "homo" else \
                non_iid_partition_with_dirichlet_distribution(ytr.numpy(), n_clients, cnum, pa)
            np.random.set_state(st)
            x, y, offs = S._pack_by_partition(xtr, ytr, part)
            ds = FederatedDataset(x, y, offs, xte, yte, None, cnum, S.ImageNormalize(mean, std), name)
        else:
            ds = S.make_cifar_like(n_clients, pm, pa, cnum, ntr, nte, 32, seed, name, mean, std)
    elif name == "fed_cifar100" and os.path.isfile(os.path.join(cache, "fed_cifar100_train.h5")):
        from . import readers as R
        ds = R.load_tff_h5(os.path.join(cache, "fed_cifar100_train.h5"), os.path.join(cache, "fed_cifar100_test.h5"),
                           "image", "label", 100, "fed_cifar100", S.ImageNormalize(S.CIFAR100_MEAN, S.CIFAR100_STD), to_uint8=True)
        if bool(getattr(args, "fed_cifar100_tff_preprocess", False)):
            # the reference's recipe (fed_cifar100/utils.py): one 24x24 crop (+ flip) per training image fixed at load time, centre
            # crop for the test images, every image standardised by its own statistics -> float NCHW, no batch-time transform
            from .dataset_utils import preprocess_cifar_img
            g = torch.Generator().manual_seed(seed)
            ds.x, ds.x_test = preprocess_cifar_img(ds.x, True, g), preprocess_cifar_img(ds.x_test, False)
            ds.transform = None
    elif name == "fed_cifar100":
        ds = S.make_cifar_like(int(getattr(args, "fed_cifar100_client_num", 500)), "hetero", 0.1, 100,
                               small or 50000, 10000 if not small else max(100, small // 5), 24, seed,
                               name, S.CIFAR100_MEAN, S.CIFAR100_STD)
    elif name in ("ilsvrc2012", "imagenet") and os.path.isdir(os.path.join(cache, "train")):
        from . import readers as R
        from ..core.non_iid_partition import homo_partition, non_iid_partition_with_dirichlet_distribution
        hw = int(getattr(args, "image_size", 224))
        xtr, ytr, xva, yva, classes = R.load_image_folder(cache, hw, "train", "val", int(getattr(args, "max_images_per_class", 0)))
        st = np.random.get_state(); np.random.seed(10)
        part = homo_partition(len(ytr), n_clients) if pm == "homo" else \
            non_iid_partition_with_dirichlet_distribution(ytr.numpy(), n_clients, len(classes), pa)
        np.random.set_state(st)
        x, y, offs = S._pack_by_partition(xtr, ytr, part)
        ds = FederatedDataset(x, y, offs, xva, yva, None, len(classes), S.ImageNormalize(S.IMAGENET_MEAN, S.IMAGENET_STD), name)
    elif name in ("gld23k", "gld160k") and os.path.isfile(os.path.join(cache, f"{name}_user_dict_train.csv")):
        from . import readers as R
        ds = R.load_landmarks_csv(os.path.join(cache, "images"), os.path.join(cache, f"{name}_user_dict_train.csv"),
                                  os.path.join(cache, f"{name}_user_dict_test.csv"), int(getattr(args, "image_size", 224)))
    elif name in ("ilsvrc2012", "imagenet", "gld23k", "gld160k"):
        clients = {"gld23k": 233, "gld160k": 1262}.get(name, n_clients)
        classes = {"gld23k": 203, "gld160k": 2028}.get(name, 1000)
        hw = int(getattr(args, "image_size", 224))
        ntr = small or 2048
        ds = S.make_cifar_like(clients, "homo" if name.startswith("il") or name == "imagenet" else "hetero", pa,
                               classes, ntr, max(64, ntr // 8), hw, seed, name, S.IMAGENET_MEAN, S.IMAGENET_STD)
    elif name == "shakespeare" and os.path.isdir(os.path.join(cache, "shakespeare", "train")):
        from . import readers as R
        ds = R.load_leaf_shakespeare(cache)
    elif name == "fed_shakespeare" and os.path.isfile(os.path.join(cache, "shakespeare_train.h5")):
        from . import readers as R
        ds = R.load_tff_shakespeare_h5(os.path.join(cache, "shakespeare_train.h5"), os.path.join(cache, "shakespeare_test.h5"),
                                       int(getattr(args, "shakespeare_client_num", 0)))
    elif name in ("stackoverflow_nwp", "stackoverflow_lr") and os.path.isfile(os.path.join(cache, "stackoverflow_train.h5")):
        from . import readers as R
        ds = R.load_tff_stackoverflow_h5(cache, "nwp" if name.endswith("nwp") else "lr",
                                         int(getattr(args, "stackoverflow_client_num", 1000 if name.endswith("nwp") else 200)))
    elif name in ("shakespeare", "fed_shakespeare"):
        ds = S.make_char_lm(715 if not small else max(4, small // 64), 80, 90, 64, seed, name,
                            next_token_only=(name == "shakespeare"))
    elif name == "stackoverflow_nwp":
        ds = S.make_char_lm(int(getattr(args, "stackoverflow_client_num", 1000)), 20, 10004, 32, seed, name,
                            next_token_only=False)
    elif name == "stackoverflow_lr":
        ds = S.make_stackoverflow_lr(int(getattr(args, "stackoverflow_client_num", 200)), seed=seed)
    elif name.startswith("synthetic"):
        parts = name.replace("synthetic", "").strip("_").split("_")
        iid = "iid" in parts
        ab = [float(p) for p in parts if p.replace(".", "", 1).isdigit()] or [1.0, 1.0]
        ds = S.make_synthetic_ab(ab[0], ab[-1], 30, seed=seed, iid=iid)
    elif name in ("fednlp", "20news", "agnews", "sst2", "text_classification", "fednlp_text"):
        classes = {"20news": 20, "agnews": 4, "sst2": 2}.get(name, int(getattr(args, "num_labels", 4)))
        ds = S.make_text_classification(n_clients, int(getattr(args, "max_seq_length", 128)), 30522, classes,
                                        int(getattr(args, "samples_per_client", 32)), seed, name)
    elif name in ("pascal_voc", "coco_seg", "cityscapes"):
        ds = S.make_segmentation(n_clients, seed=seed, name=name)
    else:
        raise ValueError(f"unknown dataset {args.dataset!r}")
    if name in _NATURAL or name.startswith("synthetic"):
        args.client_num_in_total = ds.client_num
    return ds


def _augment_enabled(args) -> bool:
    """``augment: true`` (or "cifar"): the reference's CIFAR-family train transform, RandomCrop(32, 4) + flip + Cutout(16)
    (data/cifar10/data_loader.py:58-103).  Off by default: the synthetic benchmarks measure training, not input noise."""
    v = getattr(args, "augment", False)
    return bool(v) and str(v).lower() not in ("0", "false", "none", "no")


def load(args):
    logging.info("load_data. dataset_name = %s", args.dataset)
    centralized = int(args.client_num_in_total) == 1 and getattr(args, "training_type", "") != "cross_silo"
    ds = load_federated(args)
    bs = int(args.batch_size)
    image_like = ds.transform is not None           # CIFAR-style loaders use drop_last=True (cifar10/data_loader.py:209-214)
    if image_like and _augment_enabled(args) and ds.x.dtype == torch.uint8 and ds.x.dim() == 4:
        from .augment import TrainAugment
        ds.train_augment = TrainAugment(int(getattr(args, "random_seed", 0)) + 7919, pad=int(getattr(args, "augment_pad", 4)),
                                        cutout=int(getattr(args, "augment_cutout", 16)))
    dl = getattr(args, "drop_last", None)             # explicit override; default follows the reference's loaders
    dataset8 = ds.as_tuple(bs, drop_last=(image_like if dl is None else bool(dl)) and bs > 0)
    if centralized:
        dataset8 = merge_to_single_client(dataset8)
        args.client_num_in_total = 1
    args._federated_dataset = ds
    return dataset8, ds.class_num


def load_synthetic_data(args):
    """The reference's ``load`` is a thin wrapper around this name (data/data_loader.py:29-42); both are kept."""
    return load(args)


def split_data_for_dist_trainers(batches, n_dist_trainer):
    """Chunk one client's batch list over the processes of a silo
    (reference: data/data_loader_cross_silo.py:9-55)."""
    batches = list(batches)
    n = max(1, int(n_dist_trainer))
    return [batches[i::n] for i in range(n)]


def split_array(array, n_dist_trainer):
    """Contiguous chunks of a batch list, the last batch repeated until the length divides (data_loader_cross_silo.py:9-19)."""
    array, n = list(array), max(1, int(n_dist_trainer))
    while len(array) % n:
        array.append(array[-1])
    size = len(array) // n
    return [array[i * size:(i + 1) * size] for i in range(n)]


def split_dl(dl, n_dist_trainer):
    """One DataLoader per process of a silo over a DistributedSampler (data_loader_cross_silo.py:22-47)."""
    from torch.utils import data as tud
    shuffle = isinstance(dl.sampler, tud.RandomSampler)
    return [tud.DataLoader(dl.dataset, batch_size=dl.batch_size, drop_last=dl.drop_last,
                           sampler=tud.distributed.DistributedSampler(dl.dataset, num_replicas=int(n_dist_trainer), rank=r, shuffle=shuffle))
            for r in range(int(n_dist_trainer))]


def load_synthetic_data_cross_silo(args):
    return load_cross_silo(args)


def load_cross_silo(args):
    """Like :func:`load`, but each client's train data becomes a list indexed by
    ``proc_rank_in_silo`` when the silo runs several trainer processes (hierarchical scenario)."""
    dataset8, class_num = load(args)
    n_proc = int(getattr(args, "n_proc_in_silo", 1) or 1)
    if getattr(args, "scenario", "horizontal") == "hierarchical" and n_proc > 1:
        train_local = dataset8[5]
        dataset8[5] = {c: split_data_for_dist_trainers(train_local[c], n_proc) for c in train_local.keys()}
    return dataset8, class_num


def load_poisoned_dataset_from_edge_case_examples(args):
    """Backdoor ("edge-case") data for the robust-FedAvg runner: a poisoned train loader, a clean
    test loader and a targeted-task test loader (edge_case_examples/data_loader.py:393-).  The
    real Southwest-airline / ARDIS sets are used when their files are under ``data_cache_dir`` (``data/edge_case.py``); they cannot
    be downloaded here, so otherwise a synthetic stand-in stamps a trigger patch on samples and relabels them to ``target_label``."""
    ds = getattr(args, "_federated_dataset", None) or load_federated(args)
    bs = int(args.batch_size)
    # the published sets (Southwest airplanes -> "truck" on CIFAR-10, ARDIS sevens -> "1" on MNIST) when their files are present
    from .edge_case import load_edge_case_arrays
    real = load_edge_case_arrays(str(args.dataset), str(getattr(args, "data_cache_dir", "") or ""), ds.x.numpy(), ds.y.numpy(),
                                 str(getattr(args, "poison_type", "southwest")), str(getattr(args, "attack_case", "edge-case")),
                                 int(getattr(args, "random_seed", 0))) if getattr(args, "data_cache_dir", None) else None
    if real is not None:
        xa, ya, xt, yt = (torch.from_numpy(np.ascontiguousarray(a)) for a in real)
        return (ds.batches(xa, ya.long(), bs, drop_last=False, train=True), ds.batches(ds.x_test, ds.y_test, bs),
                ds.batches(xt.to(ds.x_test.dtype), yt.long(), bs), int(xa.shape[0]))
    target = int(getattr(args, "target_label", 9)) % ds.class_num
    g = torch.Generator().manual_seed(int(getattr(args, "random_seed", 0)) + 17)
    n = min(512, ds.x.shape[0])
    idx = torch.randperm(ds.x.shape[0], generator=g)[:n]
    xp, yp = ds.x[idx].clone(), torch.full((n,), target, dtype=torch.int64)
    if xp.dim() == 4:            # NHWC uint8 images: bright 4x4 corner patch
        xp[:, :4, :4, :] = 255
    else:
        xp[:, :16] = xp.max()
    nt = min(256, ds.x_test.shape[0])
    xt = ds.x_test[:nt].clone()
    if xt.dim() == 4:
        xt[:, :4, :4, :] = 255
    else:
        xt[:, :16] = xt.max()
    yt = torch.full((nt,), target, dtype=torch.int64)
    poisoned_train = ds.batches(xp, yp, bs, drop_last=False)
    clean_test = ds.batches(ds.x_test, ds.y_test, bs)
    targeted_test = ds.batches(xt, yt, bs)
    return poisoned_train, clean_test, targeted_test, n
