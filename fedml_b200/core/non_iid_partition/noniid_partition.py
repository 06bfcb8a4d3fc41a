"""Client data partitioners.

``non_iid_partition_with_dirichlet_distribution`` follows the LDA procedure of the reference
(core/non_iid_partition/noniid_partition.py:6-109): per class draw p ~ Dir(alpha), zero the
clients already at the N/client_num cap, renormalise, cut the shuffled class indices at the
cumulative proportions, and redo the whole thing until every client holds >= 10 samples.
The NumPy global RNG call order (shuffle, dirichlet, ..., final per-client shuffle) is the
same, so a given ``np.random.seed`` produces the same split as the reference.
"""
import logging

import numpy as np


def partition_class_samples_with_dirichlet_distribution(N, alpha, client_num, idx_batch, idx_k):
    np.random.shuffle(idx_k)
    proportions = np.random.dirichlet(np.repeat(alpha, client_num))
    cap = N / client_num
    keep = np.array([len(held) < cap for held in idx_batch], dtype=np.float64)
    proportions = proportions * keep
    proportions = proportions / proportions.sum()
    cuts = (np.cumsum(proportions) * len(idx_k)).astype(int)[:-1]
    pieces = np.split(idx_k, cuts)
    idx_batch = [held + piece.tolist() for held, piece in zip(idx_batch, pieces)]
    return idx_batch, min(len(h) for h in idx_batch)


def non_iid_partition_with_dirichlet_distribution(label_list, client_num, classes, alpha,
                                                  task="classification", min_required=10):
    segmentation = task == "segmentation"
    N = len(label_list) if segmentation else label_list.shape[0]
    min_size = 0
    idx_batch = [[] for _ in range(client_num)]
    # the reference redraws until every client holds min_required samples (noniid_partition.py:44); that cannot end when there
    # are fewer samples than clients need, so the requirement is lowered to what the sample count allows
    min_required = min(int(min_required), N // max(int(client_num), 1))
    tries = 0
    while tries == 0 or min_size < min_required:
        tries += 1
        if tries % 200 == 0:                 # still unlucky after many draws (tiny alpha, many clients): relax by one
            min_required = max(0, min_required - 1)
        idx_batch = [[] for _ in range(client_num)]
        if segmentation:
            cats = list(classes)
            for c, cat in enumerate(cats):
                has = np.asarray([np.any(np.asarray(lbl) == cat) for lbl in label_list])
                if c > 0:
                    earlier = np.asarray([np.any(np.isin(np.asarray(lbl), cats[:c])) for lbl in label_list])
                    has = has & ~earlier
                idx_k = np.where(has)[0]
                idx_batch, min_size = partition_class_samples_with_dirichlet_distribution(
                    N, alpha, client_num, idx_batch, idx_k)
        else:
            for k in range(int(classes)):
                idx_k = np.where(label_list == k)[0]
                idx_batch, min_size = partition_class_samples_with_dirichlet_distribution(
                    N, alpha, client_num, idx_batch, idx_k)
    out = {}
    for i in range(client_num):
        np.random.shuffle(idx_batch[i])
        out[i] = idx_batch[i]
    return out


def homo_partition(n_samples, client_num):
    """Equal split of a random permutation (reference: data/cifar10/data_loader.py:127-131)."""
    idxs = np.random.permutation(n_samples)
    return {i: part for i, part in enumerate(np.array_split(idxs, client_num))}


def power_law_sizes(client_num, total, rng, min_size=10, exponent=1.5):
    """Client sample counts following a power law, as in the LEAF/FedProx MNIST split
    described in data/MNIST/README.md ("1000 devices ... power law")."""
    raw = rng.pareto(exponent, size=client_num) + 1.0
    sizes = np.maximum(min_size, np.floor(raw / raw.sum() * (total - min_size * client_num)).astype(np.int64) + min_size)
    return sizes


def record_data_stats(y_train, net_dataidx_map, task="classification"):
    stats = {}
    for cid, idx in net_dataidx_map.items():
        ys = np.concatenate([np.asarray(y_train[i]).ravel() for i in idx]) if task == "segmentation" \
            else np.asarray(y_train)[np.asarray(idx, dtype=np.int64)]
        unq, cnt = np.unique(ys, return_counts=True)
        stats[cid] = {int(u): int(c) for u, c in zip(unq, cnt)}
    logging.debug("Data statistics: %s", stats)
    return stats
