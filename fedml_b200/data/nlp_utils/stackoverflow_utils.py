"""Tokenisers and the lazy per-client data set of the TFF StackOverflow export (reference: data/stackoverflow_nwp/utils.py:19-90,
data/stackoverflow_lr/utils.py:19-156, data/stackoverflow_{nwp,lr}/dataset.py).  Tables are cached per directory."""
from __future__ import annotations

from collections import OrderedDict
from functools import lru_cache
from typing import List, Sequence

import numpy as np
from torch.utils import data as torch_data

from .. import readers as R

DEFAULT_WORD_COUNT_FILE, DEFAULT_TAG_COUNT_FILE = "stackoverflow.word_count", "stackoverflow.tag_count"
DEFAULT_TRAIN_CLIENTS_NUM, DEFAULT_TEST_CLIENTS_NUM = 342477, 204088
DEFAULT_BATCH_SIZE = 100
DEFAULT_TRAIN_FILE, DEFAULT_TEST_FILE, DEFAULT_CACHE_FILE = "stackoverflow_train.h5", "stackoverflow_test.h5", "stackoverflow_nwp.pkl"
_PAD, _BOS, _EOS = "<pad>", "<bos>", "<eos>"


def get_word_count_file(data_dir: str) -> str:
    import os
    return os.path.join(data_dir, DEFAULT_WORD_COUNT_FILE)


def get_tag_count_file(data_dir: str) -> str:
    import os
    return os.path.join(data_dir, DEFAULT_TAG_COUNT_FILE)


@lru_cache(maxsize=8)
def _words(data_dir: str, vocab_size: int):
    return tuple(R.stackoverflow_vocab(data_dir, vocab_size))


@lru_cache(maxsize=8)
def _tags(data_dir: str, tag_size: int):
    return tuple(R.stackoverflow_tags(data_dir, tag_size))


def get_most_frequent_words(data_dir: str, vocab_size: int = 10000) -> List[str]:
    return list(_words(data_dir, vocab_size))


def get_tags(data_dir: str, tag_size: int = 500) -> List[str]:
    return list(_tags(data_dir, tag_size))


def get_word_dict(data_dir: str, with_specials: bool = True) -> "OrderedDict[str, int]":
    """Next-word table: ``<pad>``, the 10,000 words, ``<bos>``, ``<eos>``.  ``with_specials=False`` gives the bag-of-words table
    of the tag-prediction task (words only, ids from 0)."""
    words = get_most_frequent_words(data_dir)
    names = ([_PAD] + words + [_BOS, _EOS]) if with_specials else words
    return OrderedDict((w, i) for i, w in enumerate(names))


def get_tag_dict(data_dir: str) -> "OrderedDict[str, int]":
    return OrderedDict((t, i) for i, t in enumerate(get_tags(data_dir)))


def get_word_list(data_dir: str) -> List[str]:
    return list(get_word_dict(data_dir))


def id_to_word(idx: int, data_dir: str) -> str:
    return get_word_list(data_dir)[idx]


def tokenizer(sentence: str, data_dir: str, max_seq_len: int = 20) -> List[int]:
    """``max_seq_len + 1`` ids of one sentence (next-word task)."""
    table = {w: i for i, w in enumerate(get_most_frequent_words(data_dir))}
    return R.stackoverflow_nwp_ids([sentence], table, max_seq_len)[0].tolist()


def split(dataset):
    """x = all but the last id, y = the last id (the reference's helper; the data set class shifts by one instead)."""
    ds = np.asarray(dataset)
    return ds[:, :-1], ds[:, -1]


def preprocess_inputs(sentences: Sequence[str], data_dir: str) -> List[np.ndarray]:
    return list(R.stackoverflow_bag_of_words(list(sentences), get_word_dict(data_dir, with_specials=False)))


def preprocess_input(sentence: str, data_dir: str) -> np.ndarray:
    return preprocess_inputs([sentence], data_dir)[0]


def preprocess_targets(tags: Sequence[str], data_dir: str) -> List[np.ndarray]:
    return list(R.stackoverflow_tag_targets(list(tags), get_tag_dict(data_dir)))


def preprocess_target(tag: str, data_dir: str) -> np.ndarray:
    return preprocess_targets([tag], data_dir)[0]


class StackOverflowDataset(torch_data.Dataset):
    """One client of ``stackoverflow_{train,test}.h5``, read on demand.  ``preprocess``: a callable on the sentence (next-word
    task: returns the 21 ids, the item is ``(ids[:-1], ids[1:])``) or ``{"input": fn, "target": fn}`` (tag prediction: the item
    is ``(fn(tokens + " " + title), fn(tags))``)."""

    _ids = {}

    def __init__(self, h5_path: str, client_idx: int, datast: str = "train", preprocess=None):
        if datast not in ("train", "test"):
            raise ValueError("datast must be 'train' or 'test'")
        self.h5_path, self.datast, self.preprocess = h5_path, datast, preprocess
        self.client_id = self.get_client_id_list()[client_idx]
        self._rows = None

    def get_client_id_list(self) -> List[str]:
        key = (self.h5_path, self.datast)
        if key not in StackOverflowDataset._ids:
            with R.open_h5(self.h5_path) as f:
                StackOverflowDataset._ids[key] = list(R._examples(f).keys())
        return StackOverflowDataset._ids[key]

    def _load(self):
        if self._rows is None:
            with R.open_h5(self.h5_path) as f:
                g = R._examples(f)[self.client_id]
                self._rows = {k: R._decode_all(g[k]) for k in ("tokens", "title", "tags") if k in g}
        return self._rows

    def __len__(self):
        return len(self._load()["tokens"])

    def __getitem__(self, idx):
        rows = self._load()
        if isinstance(self.preprocess, dict):
            sample = " ".join([rows["tokens"][idx], rows["title"][idx]])
            return self.preprocess["input"](sample), self.preprocess["target"](rows["tags"][idx])
        sample = rows["tokens"][idx]
        if self.preprocess is None:
            return sample
        ids = np.asarray(self.preprocess(sample))
        return ids[:-1], ids[1:]
