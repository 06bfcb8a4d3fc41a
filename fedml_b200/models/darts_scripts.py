"""Centralised DARTS on one machine: architecture search and training of a derived network.

The reference ships these as two scripts, model/cv/darts/train_search.py (search: split the training set by ``train_portion``, per
batch one architecture step on the held-out part and one weight step, cosine learning rate, genotype logged every epoch) and
model/cv/darts/train.py (train ``NetworkCIFAR`` of ``--arch`` with Cutout, auxiliary loss and a drop-path probability that grows
linearly with the epoch).  Same options, same flow, as functions: ``search_main(argv)`` / ``train_main(argv)`` return the history
instead of only logging it, and the data comes from ``data.dataset_utils`` (real CIFAR-10 under ``--data`` when present, otherwise
the synthetic set; ``--train_size`` caps it for quick runs).

    python -m fedml_b200.models.darts_scripts search --epochs 2 --layers 5 --init_channels 8 --train_size 512
    python -m fedml_b200.models.darts_scripts train --arch FedNAS_V1 --epochs 2 --layers 5 --init_channels 8 --train_size 512
"""
from __future__ import annotations

import argparse
import logging
import os
import sys
import time
from typing import List, Optional

import numpy as np
import torch
import torch.nn as nn
from torch.utils import data as torch_data

from . import darts as D

CIFAR_CLASSES = 10
log_format = "%(asctime)s %(message)s"


def build_parser(search: bool) -> argparse.ArgumentParser:
    p = argparse.ArgumentParser("darts " + ("search" if search else "train"))
    p.add_argument("--data", type=str, default="./data", help="directory holding cifar-10-batches-py (synthetic data if absent)")
    p.add_argument("--train_size", type=int, default=0, help="cap on the number of training images (0 = all)")
    p.add_argument("--batch_size", type=int, default=64 if search else 96)
    p.add_argument("--learning_rate", type=float, default=0.025)
    p.add_argument("--learning_rate_min", type=float, default=0.001)
    p.add_argument("--momentum", type=float, default=0.9)
    p.add_argument("--weight_decay", type=float, default=3e-4)
    p.add_argument("--report_freq", type=int, default=50)
    p.add_argument("--gpu", type=str, default="0")
    p.add_argument("--epochs", type=int, default=50 if search else 600)
    p.add_argument("--init_channels", type=int, default=16 if search else 36)
    p.add_argument("--layers", type=int, default=8 if search else 20)
    p.add_argument("--cutout", action="store_true", default=False)
    p.add_argument("--cutout_length", type=int, default=16)
    p.add_argument("--save", type=str, default="EXP")
    p.add_argument("--seed", type=int, default=2 if search else 0)
    p.add_argument("--grad_clip", type=float, default=5.0)
    if search:
        p.add_argument("--train_portion", type=float, default=0.5)
        p.add_argument("--unrolled", action="store_true", default=False, help="second-order architecture step")
        p.add_argument("--arch_learning_rate", type=float, default=3e-4)
        p.add_argument("--arch_weight_decay", type=float, default=1e-3)
        p.add_argument("--optimization", type=str, default="DARTS", help="DARTS | MiLeNAS (mixed-level) | single-level | AOS")
        p.add_argument("--arch_search_method", type=str, default="DARTS", help="DARTS | GDAS")
        p.add_argument("--lambda_train_regularizer", type=float, default=1.0)
        p.add_argument("--lambda_valid_regularizer", type=float, default=1.0)
        p.add_argument("--tau_max", type=float, default=10.0)
        p.add_argument("--tau_min", type=float, default=1.0)
    else:
        p.add_argument("--arch", type=str, default="DARTS")
        p.add_argument("--auxiliary", action="store_true", default=False)
        p.add_argument("--auxiliary_weight", type=float, default=0.4)
        p.add_argument("--drop_path_prob", type=float, default=0.2)
        p.add_argument("--model_path", type=str, default="")
    return p


search_parser, train_parser = build_parser(True), build_parser(False)


def _device(args) -> torch.device:
    return torch.device(f"cuda:{str(args.gpu).split(',')[0]}") if torch.cuda.is_available() else torch.device("cpu")


def _datasets(args):
    from ..data import dataset_utils as U
    train_t, test_t = U.data_transforms("cifar10")
    if not args.cutout:
        train_t.cutout = None
    elif train_t.cutout is not None:
        train_t.cutout.length = args.cutout_length
    old = U.SYNTHETIC_TRAIN_SIZE
    try:
        if args.train_size and not os.path.isdir(os.path.join(args.data, "cifar-10-batches-py")):
            U.SYNTHETIC_TRAIN_SIZE = args.train_size
        train = U.CIFAR10_truncated(args.data, train=True, transform=train_t)
        test = U.CIFAR10_truncated(args.data, train=False, transform=test_t)
    finally:
        U.SYNTHETIC_TRAIN_SIZE = old
    if args.train_size and len(train) > args.train_size:
        train = U.CIFAR10_truncated(args.data, dataidxs=list(range(args.train_size)), train=True, transform=train_t)
        test = U.CIFAR10_truncated(args.data, dataidxs=list(range(max(64, args.train_size // 5))), train=False, transform=test_t)
    return train, test


def infer(valid_queue, model, criterion, device=None, report_freq: int = 0):
    """Top-1 accuracy (percent) and mean loss of ``model`` over ``valid_queue``."""
    device = device or next(model.parameters()).device
    objs, top1 = D.AvgrageMeter(), D.AvgrageMeter()
    model.eval()
    with torch.no_grad():
        for step, (x, y) in enumerate(valid_queue):
            x, y = x.to(device), y.to(device)
            out = model(x)
            logits = out[0] if isinstance(out, tuple) else out
            objs.update(float(criterion(logits, y)), x.size(0))
            top1.update(float(D.accuracy(logits, y)[0]), x.size(0))
            if report_freq and step % report_freq == 0:
                logging.info("valid %03d %e %f", step, objs.avg, top1.avg)
    return top1.avg, objs.avg


def train_search_epoch(train_queue, valid_queue, model, architect, criterion, optimizer, lr, args, device=None):
    """One search epoch: for every training batch, an architecture step on the next held-out batch (rule per
    ``args.optimization`` / ``args.unrolled``), then an SGD step on the weights with gradient clipping."""
    device = device or next(model.parameters()).device
    objs, top1 = D.AvgrageMeter(), D.AvgrageMeter()
    held_out = iter(valid_queue)
    rule = str(getattr(args, "optimization", "DARTS")).lower()
    weights = model.weight_parameters()
    for step, (x, y) in enumerate(train_queue):
        model.train()
        x, y = x.to(device), y.to(device)
        try:
            xs, ys = next(held_out)
        except StopIteration:
            held_out = iter(valid_queue)
            xs, ys = next(held_out)
        xs, ys = xs.to(device), ys.to(device)
        if rule == "milenas":
            if args.unrolled:
                architect.step_v2_2ndorder(x, y, xs, ys, lr, optimizer, args.lambda_train_regularizer, args.lambda_valid_regularizer)
            else:
                architect.step_v2(x, y, xs, ys, args.lambda_train_regularizer, args.lambda_valid_regularizer)
        elif rule == "single-level":
            architect.step_single_level(x, y)
        elif rule == "aos":
            architect.step_AOS(x, y, xs, ys)
        else:
            architect.step(x, y, xs, ys, lr, optimizer, unrolled=args.unrolled)
        optimizer.zero_grad(set_to_none=True)
        logits = model(x)
        loss = criterion(logits, y)
        loss.backward()
        nn.utils.clip_grad_norm_(weights, args.grad_clip)
        optimizer.step()
        objs.update(float(loss), x.size(0))
        top1.update(float(D.accuracy(logits.detach(), y)[0]), x.size(0))
        if args.report_freq and step % args.report_freq == 0:
            logging.info("train %03d %e %f", step, objs.avg, top1.avg)
    return top1.avg, objs.avg


def search_main(argv: Optional[List[str]] = None) -> List[dict]:
    args = search_parser.parse_args(argv)
    np.random.seed(args.seed)
    torch.manual_seed(args.seed)
    device = _device(args)
    D.create_exp_dir(args.save)
    criterion = nn.CrossEntropyLoss().to(device)
    cls = D.NetworkGDAS if args.arch_search_method.upper() == "GDAS" else D.Network
    model = cls(args.init_channels, CIFAR_CLASSES, args.layers, criterion, device).to(device)
    logging.info("param size = %fMB", D.count_parameters_in_MB(model))
    optimizer = torch.optim.SGD(model.weight_parameters(), args.learning_rate, momentum=args.momentum, weight_decay=args.weight_decay)
    train_data, _ = _datasets(args)
    split = int(np.floor(args.train_portion * len(train_data)))
    idx = list(range(len(train_data)))
    mk = lambda part: torch_data.DataLoader(train_data, batch_size=args.batch_size, sampler=torch_data.SubsetRandomSampler(part))  # noqa: E731
    train_queue, valid_queue = mk(idx[:split]), mk(idx[split:])
    scheduler = torch.optim.lr_scheduler.CosineAnnealingLR(optimizer, float(args.epochs), eta_min=args.learning_rate_min)
    architect = D.Architect(model, criterion, args, device)
    history = []
    for epoch in range(args.epochs):
        lr = scheduler.get_last_lr()[0]
        if hasattr(model, "set_tau"):
            model.set_tau(args.tau_max - (args.tau_max - args.tau_min) * epoch / max(1, args.epochs - 1))
        t0 = time.time()
        train_acc, train_loss = train_search_epoch(train_queue, valid_queue, model, architect, criterion, optimizer, lr, args, device)
        valid_acc, valid_loss = infer(valid_queue, model, criterion, device)
        scheduler.step()
        genotype, n_cnn, r_cnn = model.genotype()
        history.append({"epoch": epoch, "lr": lr, "train_acc": train_acc, "train_loss": train_loss, "valid_acc": valid_acc,
                        "valid_loss": valid_loss, "genotype": genotype, "normal_cnn_count": n_cnn, "reduce_cnn_count": r_cnn,
                        "model_size_mb": model.get_current_model_size(), "seconds": time.time() - t0})
        logging.info("epoch %d lr %e train_acc %f valid_acc %f genotype = %s", epoch, lr, train_acc, valid_acc, genotype)
        D.save(model, os.path.join(args.save, "weights.pt"))
    return history


def train_epoch(train_queue, model, criterion, optimizer, args, device=None):
    """One epoch of a derived network: cross-entropy plus ``auxiliary_weight`` times the auxiliary head's loss."""
    device = device or next(model.parameters()).device
    objs, top1 = D.AvgrageMeter(), D.AvgrageMeter()
    model.train()
    for step, (x, y) in enumerate(train_queue):
        x, y = x.to(device), y.to(device)
        optimizer.zero_grad(set_to_none=True)
        logits, logits_aux = model(x)
        loss = criterion(logits, y)
        if args.auxiliary and logits_aux is not None:
            loss = loss + args.auxiliary_weight * criterion(logits_aux, y)
        loss.backward()
        nn.utils.clip_grad_norm_(model.parameters(), args.grad_clip)
        optimizer.step()
        objs.update(float(loss), x.size(0))
        top1.update(float(D.accuracy(logits.detach(), y)[0]), x.size(0))
        if args.report_freq and step % args.report_freq == 0:
            logging.info("train %03d %e %f", step, objs.avg, top1.avg)
    return top1.avg, objs.avg


def train_main(argv: Optional[List[str]] = None) -> List[dict]:
    args = train_parser.parse_args(argv)
    np.random.seed(args.seed)
    torch.manual_seed(args.seed)
    device = _device(args)
    D.create_exp_dir(args.save)
    genotype = getattr(D, args.arch)
    model = D.NetworkCIFAR(args.init_channels, CIFAR_CLASSES, args.layers, args.auxiliary, genotype).to(device)
    if args.model_path:
        D.load(model, args.model_path)
    logging.info("param size = %fMB", D.count_parameters_in_MB(model))
    criterion = nn.CrossEntropyLoss().to(device)
    optimizer = torch.optim.SGD(model.parameters(), args.learning_rate, momentum=args.momentum, weight_decay=args.weight_decay)
    train_data, valid_data = _datasets(args)
    train_queue = torch_data.DataLoader(train_data, batch_size=args.batch_size, shuffle=True)
    valid_queue = torch_data.DataLoader(valid_data, batch_size=args.batch_size, shuffle=False)
    scheduler = torch.optim.lr_scheduler.CosineAnnealingLR(optimizer, float(args.epochs))
    history, best = [], 0.0
    for epoch in range(args.epochs):
        model.drop_path_prob = args.drop_path_prob * epoch / args.epochs
        train_acc, train_loss = train_epoch(train_queue, model, criterion, optimizer, args, device)
        valid_acc, valid_loss = infer(valid_queue, model, criterion, device)
        scheduler.step()
        best = max(best, valid_acc)
        history.append({"epoch": epoch, "train_acc": train_acc, "train_loss": train_loss, "valid_acc": valid_acc, "valid_loss": valid_loss,
                        "drop_path_prob": model.drop_path_prob})
        logging.info("epoch %d train_acc %f valid_acc %f (best %f)", epoch, train_acc, valid_acc, best)
        D.save(model, os.path.join(args.save, "weights.pt"))
    return history


def main(argv: Optional[List[str]] = None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] not in ("search", "train"):
        raise SystemExit("usage: python -m fedml_b200.models.darts_scripts {search|train} [options]")
    logging.basicConfig(stream=sys.stdout, level=logging.INFO, format=log_format)
    history = (search_main if argv[0] == "search" else train_main)(argv[1:])
    print(history[-1])


if __name__ == "__main__":
    main()
