"""Tokenisation of the TFF Shakespeare export (reference: data/fed_shakespeare/utils.py:15-84, data_loader.py:12-20)."""
from __future__ import annotations

from collections import OrderedDict
from typing import List, Sequence, Tuple

import numpy as np

from ..readers import shakespeare_char_ids, shakespeare_sequences

SEQUENCE_LENGTH = 80
DEFAULT_TRAIN_CLIENTS_NUM = DEFAULT_TEST_CLIENTS_NUM = 715
DEFAULT_BATCH_SIZE = 4
DEFAULT_TRAIN_FILE, DEFAULT_TEST_FILE = "shakespeare_train.h5", "shakespeare_test.h5"


def get_word_dict() -> "OrderedDict[str, int]":
    return OrderedDict(shakespeare_char_ids())


def get_word_list() -> List[str]:
    return list(shakespeare_char_ids())


def id_to_word(idx: int) -> str:
    return get_word_list()[idx]


def char_to_id(char: str) -> int:
    table = shakespeare_char_ids()
    return table.get(char, len(table))


def preprocess(sentences: Sequence[str], max_seq_len: int = SEQUENCE_LENGTH) -> List[List[int]]:
    """Rows of ``max_seq_len + 1`` ids: ``<bos> text <eos>`` padded to a whole number of rows."""
    return shakespeare_sequences(sentences, max_seq_len).tolist()


def split(dataset) -> Tuple[np.ndarray, np.ndarray]:
    ds = np.asarray(dataset)
    return ds[:, :-1], ds[:, 1:]
