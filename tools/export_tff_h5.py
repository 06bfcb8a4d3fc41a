"""Writes synthetic benchmark data in the TFF .h5 layout (the files the reference's federated loaders open).

    python tools/export_tff_h5.py femnist ./data/femnist --clients 200 --train-size 20000
    python tools/export_tff_h5.py fed_cifar100 ./data/fed_cifar100
    python tools/export_tff_h5.py fed_shakespeare ./data/fed_shakespeare --clients 64
    python tools/export_tff_h5.py stackoverflow_nwp ./data/stackoverflow --clients 500      # also serves stackoverflow_lr

Afterwards ``fedml_b200.data.load`` with ``data_cache_dir`` pointing at the directory reads the files back (through h5py when
installed, else through fedml_b200.data.minihdf5), and so does the reference.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from fedml_b200.data import synthetic as S, tff_export  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("dataset", choices=["femnist", "fed_cifar100", "fed_shakespeare", "stackoverflow_nwp", "stackoverflow_lr"])
    ap.add_argument("out_dir")
    ap.add_argument("--clients", type=int, default=0)
    ap.add_argument("--train-size", type=int, default=0)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    if a.dataset == "femnist":
        ds = S.make_femnist_like(client_num=a.clients or 3400, seed=a.seed, total_train=a.train_size or 340000)
    elif a.dataset == "fed_cifar100":
        n = a.train_size or 50000
        ds = S.make_cifar_like(a.clients or 500, "hetero", 0.1, 100, n, max(100, n // 5), 32, a.seed, "fed_cifar100",
                               S.CIFAR100_MEAN, S.CIFAR100_STD)
    elif a.dataset == "fed_shakespeare":
        ds = S.make_char_lm(a.clients or 715, 80, 87, 64, a.seed, "fed_shakespeare", next_token_only=False)
    else:
        ds = S.make_char_lm(a.clients or 1000, 20, 10001, 32, a.seed, "stackoverflow_nwp", next_token_only=False)
    for p in tff_export.export(a.dataset, ds, a.out_dir):
        print(p, os.path.getsize(p), "bytes")


if __name__ == "__main__":
    main()
