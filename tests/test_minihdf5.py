"""Pure-Python HDF5 reader / writer and the TFF .h5 dataset readers built on it."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from fedml_b200.arguments import Arguments
from fedml_b200.data import minihdf5 as H
from fedml_b200.data import readers as R
from fedml_b200.data import synthetic as S
from fedml_b200.data import tff_export

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reads_a_file_written_by_the_hdf5_library():
    """scipy ships a MATLAB v7.3 file = an HDF5 file behind a 512-byte user block, written by the C library."""
    import scipy.io
    path = os.path.join(os.path.dirname(scipy.io.__file__), "matlab", "tests", "data", "testhdf5_7.4_GLNX86.mat")
    if not os.path.isfile(path):
        pytest.skip("scipy test data not installed")
    with H.File(path) as f:
        assert f.keys() == ["testdouble"]
        d = f["testdouble"]
        assert d.shape == (9, 1) and d.dtype == np.dtype("<f8")
        np.testing.assert_allclose(d[()].ravel(), np.arange(9) * np.pi / 4, rtol=1e-15)
        assert d.attrs["MATLAB_class"] == b"double"
        assert np.asarray(d).shape == (9, 1) and d[2:4, 0].tolist() == [np.pi / 2, 3 * np.pi / 4]


def test_write_read_round_trip_of_every_supported_storage_form(tmp_path):
    rs = np.random.RandomState(0)
    clients = {f"c{i:04d}": {"pixels": rs.rand(1 + i % 4, 5, 5).astype(np.float32), "label": rs.randint(0, 62, (1 + i % 4,)).astype(np.int32)}
               for i in range(300)}                                   # 300 links: three levels of the group B-tree
    img = rs.randint(0, 255, (37, 19, 3)).astype(np.uint8)
    big = np.arange(70 * 40, dtype=np.int64).reshape(70, 40)       # 70 chunks: two levels of the chunk B-tree
    words = ["hello world", "", "Ünïcode ✓", "x" * 5000]
    tree = {"examples": clients, "text": {"snippets": words, "raw": [b"ab", b"cde"]},
            "img": H.H5Data(img, chunks=(8, 8, 3), compression=4, shuffle=True),
            "big": H.H5Data(big, chunks=(1, 40), compression=1), "plain_chunks": H.H5Data(big.astype(np.float64), chunks=(16, 16)),
            "f16": np.linspace(0, 1, 7).astype(np.float16), "i16": np.array([[1, -2], [3, 4]], dtype=np.int16),
            "u64": np.array([2 ** 63 + 5], dtype=np.uint64), "fixed": np.array([b"ab", b"cdef"]), "flag": np.array([True, False]),
            "scalar": np.float32(3.5), "empty_group": {}, "empty_data": np.zeros((0, 4), np.float32),
            "chunked_text": H.H5Data([f"s{i}" for i in range(10)], chunks=(4,), compression=6)}
    path = H.write_hdf5(str(tmp_path / "t.h5"), tree)
    with H.File(path) as f:
        assert f.keys() == sorted(tree) and len(f["examples"]) == 300 and "c0007" in f["examples"] and "nope" not in f
        assert f["examples"].keys() == sorted(clients)
        for k in ("c0000", "c0008", "c0150", "c0299"):
            np.testing.assert_array_equal(f["examples"][k]["pixels"][()], clients[k]["pixels"])
            np.testing.assert_array_equal(f[f"/examples/{k}/label"][()], clients[k]["label"])
        assert [b.decode("utf8") for b in f["text/snippets"][()]] == words
        assert list(f["text/snippets"].asstr()) == words and list(f["text"]["raw"][()]) == [b"ab", b"cde"]
        np.testing.assert_array_equal(f["img"][()], img)
        np.testing.assert_array_equal(f["big"][()], big)
        np.testing.assert_array_equal(f["big"][3:5], big[3:5])
        np.testing.assert_array_equal(f["plain_chunks"][()], big.astype(np.float64))
        np.testing.assert_array_equal(f["f16"][()], tree["f16"])
        np.testing.assert_array_equal(f["i16"][()], tree["i16"])
        assert f["u64"][()][0] == 2 ** 63 + 5 and f["fixed"][()].tolist() == [b"ab", b"cdef"] and f["flag"][()].tolist() == [1, 0]
        assert f["scalar"][()] == np.float32(3.5) and f["scalar"].shape == ()
        assert f["empty_group"].keys() == [] and f["empty_data"][()].shape == (0, 4)
        assert [b.decode() for b in f["chunked_text"][()]] == [f"s{i}" for i in range(10)]
        seen = []
        f["text"].visit(seen.append)
        assert seen == ["raw", "snippets"]
        with pytest.raises(KeyError):
            f["examples"]["c9999"]
    with pytest.raises(H.Hdf5Error):
        (tmp_path / "junk.h5").write_bytes(b"not an hdf5 file at all" * 10)
        H.File(str(tmp_path / "junk.h5"))


def _args(**kw):
    base = dict(dataset="femnist", batch_size=8, client_num_in_total=6, client_num_per_round=2, log_level="WARNING")
    base.update(kw)
    return Arguments.from_dict(base)


def test_fed_emnist_and_fed_cifar100_files_load_through_the_public_loader(tmp_path):
    import fedml_b200
    src = S.make_femnist_like(client_num=6, seed=1, total_train=240)
    d = tmp_path / "femnist"
    paths = tff_export.export("femnist", src, str(d))
    assert [os.path.basename(p) for p in paths] == ["fed_emnist_train.h5", "fed_emnist_test.h5"]
    ds, classes = fedml_b200.data.load(_args(data_cache_dir=str(d)))
    fed = fedml_b200.data.load_federated(_args(data_cache_dir=str(d)))
    assert classes == 62 and fed.client_num == 6 and fed.name == "femnist"
    np.testing.assert_array_equal(fed.client_offsets, src.client_offsets)
    np.testing.assert_allclose(fed.x.reshape(len(fed.x), -1).numpy(), src.x.reshape(len(src.x), -1).numpy())
    np.testing.assert_array_equal(fed.y.numpy(), src.y.numpy())
    assert ds[0] == len(src.x)

    src = S.make_cifar_like(5, "hetero", 0.5, 100, 200, 50, 32, 0, "fed_cifar100", S.CIFAR100_MEAN, S.CIFAR100_STD)
    d = tmp_path / "fc100"
    tff_export.export("fed_cifar100", src, str(d))
    fed = fedml_b200.data.load_federated(_args(dataset="fed_cifar100", data_cache_dir=str(d)))
    assert fed.client_num == 5 and fed.x.dtype == torch.uint8 and tuple(fed.x.shape[1:]) == (32, 32, 3)
    np.testing.assert_array_equal(fed.x.numpy(), src.x.numpy())
    np.testing.assert_array_equal(fed.y.numpy(), src.y.numpy())


def test_fed_shakespeare_snippets_become_81_id_rows(tmp_path):
    import fedml_b200
    snippets = {"PLAY_A": ["Yonder comes my master, your brother.", "x" * 100], "PLAY_B": ["Is 'old dog' my reward? ~"]}
    for split in ("train", "test"):
        H.write_hdf5(str(tmp_path / f"shakespeare_{split}.h5"), {"examples": {k: {"snippets": v} for k, v in snippets.items()}})
    fed = fedml_b200.data.load_federated(_args(dataset="fed_shakespeare", data_cache_dir=str(tmp_path)))
    table = {c: i + 1 for i, c in enumerate(R.CHAR_VOCAB)}
    assert fed.class_num == 90 and fed.client_num == 2 and fed.task == "nwp"
    assert fed.client_offsets.tolist() == [0, 1 + 2, 4]               # 39 ids -> 1 row; 102 ids -> 2 rows; 28 ids -> 1 row
    first = [87] + [table[c] for c in snippets["PLAY_A"][0]] + [88]
    row = first + [0] * (81 - len(first))
    assert fed.x[0].tolist() == row[:-1] and fed.y[0].tolist() == row[1:]
    second = [87] + [table["x"]] * 100 + [88]
    assert fed.x[1].tolist() == second[:80] and fed.x[2].tolist()[:21] == second[81:] and fed.x[2].tolist()[21:] == [0] * 59
    assert fed.x[3].tolist()[-4:-1] == [0, 0, 0] and 89 in fed.x[3].tolist()      # "~" is out of vocabulary
    # the exporter's inverse table round-trips plain text
    assert tff_export.ids_to_snippet(first) == snippets["PLAY_A"][0]


def test_stackoverflow_next_word_and_tag_prediction_from_h5(tmp_path):
    import fedml_b200
    vocab = [f"w{i}" for i in range(10000)]
    (tmp_path / "stackoverflow.word_count").write_text("".join(f"{w} {20000 - i}\n" for i, w in enumerate(vocab)) + "rare 1\n")
    (tmp_path / "stackoverflow.tag_count").write_text(json.dumps({f"t{i}": 1000 - i for i in range(600)}))
    clients = {"00001": {"tokens": ["w0 w5 zzz", " ".join(["w1"] * 25)], "title": ["w7", "w2 w2"], "tags": ["t0|t499|t599", "t3"]},
               "00002": {"tokens": ["w9999"], "title": [""], "tags": ["nothing"]}}
    for split in ("train", "test"):
        H.write_hdf5(str(tmp_path / f"stackoverflow_{split}.h5"), {"examples": clients})
    fed = fedml_b200.data.load_federated(_args(dataset="stackoverflow_nwp", data_cache_dir=str(tmp_path)))
    assert fed.class_num == 10004 and fed.client_offsets.tolist() == [0, 2, 3] and tuple(fed.x.shape) == (3, 20)
    row = [10001, 1, 6, 10003, 10002] + [0] * 16                      # <bos> w0 w5 oov <eos> pad...
    assert fed.x[0].tolist() == row[:-1] and fed.y[0].tolist() == row[1:]
    assert fed.x[1].tolist() == [10001] + [2] * 19 and fed.y[1].tolist() == [2] * 20        # 20 words: no <eos>
    assert fed.x[2].tolist()[:3] == [10001, 10000, 10002]

    fed = fedml_b200.data.load_federated(_args(dataset="stackoverflow_lr", data_cache_dir=str(tmp_path)))
    assert tuple(fed.x.shape) == (3, 10000) and tuple(fed.y.shape) == (3, 500) and fed.task == "tag_prediction"
    x0 = fed.x[0]
    assert x0[0] == 0.25 and x0[5] == 0.25 and x0[7] == 0.25 and float(x0.sum()) == 0.75        # 4 words, one unknown
    assert abs(float(fed.x[1][1]) - 25 / 27) < 1e-6 and abs(float(fed.x[1][2]) - 2 / 27) < 1e-6
    assert fed.y[0].nonzero().flatten().tolist() == [0, 499] and fed.y[1].nonzero().flatten().tolist() == [3] and fed.y[2].sum() == 0


def test_export_tool_and_stackoverflow_round_trip(tmp_path):
    out = tmp_path / "so"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "export_tff_h5.py"), "stackoverflow_nwp", str(out), "--clients", "12"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-800:]
    assert sorted(os.listdir(out)) == ["stackoverflow.tag_count", "stackoverflow.word_count", "stackoverflow_test.h5", "stackoverflow_train.h5"]
    src = S.make_char_lm(12, 20, 10001, 32, 0, "stackoverflow_nwp", next_token_only=False)
    fed = R.load_tff_stackoverflow_h5(str(out), "nwp", 0)
    assert fed.client_num == 12 and np.array_equal(fed.client_offsets, src.client_offsets)
    # sentence = the 20 input ids; the loader puts <bos> in front, so its targets are the source's inputs
    np.testing.assert_array_equal(fed.y.numpy(), src.x.numpy())
    lr = R.load_tff_stackoverflow_h5(str(out), "lr", 4)
    assert lr.client_num == 4 and lr.x.shape[1] == 10000 and lr.y.shape[1] == 500 and float(lr.y.sum()) > 0


def test_reference_named_text_helpers(tmp_path):
    """Same names and call signatures as data/{shakespeare,fed_shakespeare,stackoverflow_*}/..., importable through ``fedml.data``."""
    from fedml.data.fed_shakespeare.utils import SEQUENCE_LENGTH, char_to_id, get_word_dict, id_to_word, preprocess, split
    from fedml.data.shakespeare.language_utils import VOCAB_SIZE, bag_of_words, letter_to_index, letter_to_vec, line_to_indices, split_line
    from fedml.data.stackoverflow_lr.utils import get_tag_dict, preprocess_input, preprocess_target
    from fedml.data.stackoverflow_nwp.dataset import StackOverflowDataset
    from fedml.data.stackoverflow_nwp.utils import get_word_dict as so_words, tokenizer
    from fedml.data.synthetic_1_1.generate_synthetic import generate_synthetic, main as write_synthetic, softmax
    assert VOCAB_SIZE == 90 and letter_to_index("d") == 0 and letter_to_index("~") == -1 and sum(letter_to_vec("h")) == 1
    assert split_line("it's ok, right?") == ["it's", "ok", ",", "right", "?"]
    assert line_to_indices("b a zzz", {"a": 0, "b": 1}, 5) == [1, 0, 2, 2, 2] and bag_of_words("a a b c", {"a": 0, "b": 1}) == [2, 1]
    table = get_word_dict()
    assert SEQUENCE_LENGTH == 80 and len(table) == 89 and table["<pad>"] == 0 and table["<bos>"] == 87 and char_to_id("~") == 89
    assert id_to_word(1) == "d"
    rows = preprocess(["abc", "x" * 100])
    x, y = split(rows)
    assert len(rows) == 3 and x.shape == (3, 80) and y.shape == (3, 80) and rows[0][:5] == [87, char_to_id("a"), char_to_id("b"), char_to_id("c"), 88]

    (tmp_path / "stackoverflow.word_count").write_text("".join(f"w{i} {9 - i}\n" for i in range(8)))
    (tmp_path / "stackoverflow.tag_count").write_text(json.dumps({"python": 5, "c++": 3, "sql": 1}))
    d = str(tmp_path)
    assert list(so_words(d))[:2] == ["<pad>", "w0"] and so_words(d)["<eos>"] == 10 and get_tag_dict(d) == {"python": 0, "c++": 1, "sql": 2}
    assert tokenizer("w0 w7 nope", d, max_seq_len=5) == [9, 1, 8, 11, 10, 0]
    np.testing.assert_allclose(preprocess_input("w1 w1 nope w3", d), [0, 0.5, 0, 0.25, 0, 0, 0, 0])
    assert preprocess_target("sql|python|rust", d).tolist() == [1, 0, 1]
    H.write_hdf5(str(tmp_path / "stackoverflow_train.h5"),
                 {"examples": {"u1": {"tokens": ["w0 w1", "w2"], "title": ["w3", "w4"], "tags": ["sql", "c++|python"]}, "u2": {"tokens": ["w5"], "title": [""], "tags": [""]}}})
    ds = StackOverflowDataset(str(tmp_path / "stackoverflow_train.h5"), 0, "train", lambda s: tokenizer(s, d, 4))
    assert len(ds) == 2 and ds.client_id == "u1" and [a.tolist() for a in ds[0]] == [[9, 1, 2, 10], [1, 2, 10, 0]]
    lr = StackOverflowDataset(str(tmp_path / "stackoverflow_train.h5"), 0, "train",
                              {"input": lambda s: preprocess_input(s, d), "target": lambda t: preprocess_target(t, d)})
    xin, tgt = lr[1]
    assert xin[2] == 0.5 and xin[4] == 0.5 and tgt.tolist() == [1, 1, 0]

    assert abs(softmax([1.0, 2.0, 3.0]).sum() - 1) < 1e-12
    xs, ys = generate_synthetic(1, 1, 0)
    assert len(xs) == 30 and len(xs[0][0]) == 60 and set(ys[0]) <= set(float(k) for k in range(10)) and min(len(u) for u in ys) >= 50
    tr, te = write_synthetic(str(tmp_path / "syn"))
    blob = json.load(open(tr))
    assert len(blob["users"]) == 30 and blob["users"][0] == "f_00000" and blob["num_samples"][0] == len(blob["user_data"]["f_00000"]["y"])
    from fedml.data.synthetic_1_1.stats import print_dataset_stats
    stats = print_dataset_stats("syn", str(tmp_path))
    assert stats["users"] == 30 and stats["samples"] == sum(blob["num_samples"])


def test_fed_cifar100_h5_with_the_tff_preprocessing_switch(tmp_path):
    import fedml_b200
    src = S.make_cifar_like(4, "homo", 0.5, 100, 40, 12, 32, 0, "fed_cifar100", S.CIFAR100_MEAN, S.CIFAR100_STD)
    tff_export.export("fed_cifar100", src, str(tmp_path))
    fed = fedml_b200.data.load_federated(_args(dataset="fed_cifar100", data_cache_dir=str(tmp_path), fed_cifar100_tff_preprocess=True))
    assert fed.x.dtype == torch.float32 and tuple(fed.x.shape) == (40, 3, 24, 24) and tuple(fed.x_test.shape[1:]) == (3, 24, 24)
    assert fed.transform is None and abs(float(fed.x_test[0].mean())) < 0.5
    ds, k = fedml_b200.data.load(_args(dataset="fed_cifar100", data_cache_dir=str(tmp_path), fed_cifar100_tff_preprocess=True))
    xb, yb = next(iter(ds[5][0]))
    assert xb.shape[1:] == (3, 24, 24) and k == 100


def test_row_reads_touch_only_the_requested_rows_and_imagenet_hdf5(tmp_path):
    rs = np.random.RandomState(2)
    big = rs.randint(0, 255, (50, 12, 10, 3)).astype(np.uint8)
    labels = rs.randint(0, 4, 50).astype(np.int64)
    words = [f"w{i}" for i in range(23)]
    path = H.write_hdf5(str(tmp_path / "imagenet.h5"), {
        "train_img": big, "train_labels": labels, "val_img": H.H5Data(big[:20], chunks=(3, 12, 10, 3), compression=2), "val_labels": labels[:20],
        "ragged": H.H5Data(np.arange(7 * 5).reshape(7, 5), chunks=(2, 3)), "words": H.H5Data(words, chunks=(4,)), "scalar": np.int32(5)})
    with H.File(path) as f:
        d = f["train_img"]
        np.testing.assert_array_equal(d[7], big[7])
        np.testing.assert_array_equal(d[-1], big[-1])
        np.testing.assert_array_equal(d[10:31:5], big[10:31:5])
        np.testing.assert_array_equal(d[3:9, 2:4, :, 1], big[3:9, 2:4, :, 1])
        np.testing.assert_array_equal(d[4, 1:3], big[4, 1:3])
        np.testing.assert_array_equal(d[[1, 5, 9]], big[[1, 5, 9]])              # fancy index: whole-array path
        with pytest.raises(IndexError):
            d[50]
        v = f["val_img"]                                                           # chunked + deflate: rows across chunk borders
        for key in (0, 2, 3, 19, slice(2, 8), slice(17, 20), slice(0, 20, 7)):
            np.testing.assert_array_equal(v[key], big[:20][key])
        r = f["ragged"]
        for key in (0, 6, slice(1, 6), slice(5, 7)):
            np.testing.assert_array_equal(r[key], np.arange(35).reshape(7, 5)[key])
        assert [b.decode() for b in f["words"][5:11]] == words[5:11] and f["words"][22] == b"w22" and f["scalar"][()] == 5
        # only the requested rows are decoded: reading one row of the chunked set decodes one chunk
        calls = []
        orig = v._unfilter
        v._unfilter = lambda raw, filters, mask: (calls.append(1), orig(raw, filters, mask))[1]
        v[4]
        assert len(calls) == 1
    from fedml_b200.data.folder_datasets import DatasetHDF5, ImageNet_hdf5, ImageNet_truncated_hdf5
    raw = DatasetHDF5(path, "train")
    img, lab = raw[11]
    assert len(raw) == 50 and np.array_equal(img, big[11]) and lab == labels[11]
    ds = ImageNet_hdf5(path, dataidxs=None, train=True)
    assert len(ds) == 50 and sum(ds.get_data_local_num_dict().values()) == 50
    cls = int(labels[0])
    one = ImageNet_hdf5(path, dataidxs=cls, train=True)
    assert len(one) == int((labels == cls).sum()) and all(one[i][1] == cls for i in range(len(one)))
    x, y = one[0]
    assert x.shape == (3, 12, 10) and abs(float(x[0, 0, 0]) - big[np.where(labels == cls)[0][0], 0, 0, 0] / 255) < 1e-6
    view = ImageNet_truncated_hdf5(ds, [0, 1], ds.get_net_dataidx_map())
    assert len(view) == int(((labels == 0) | (labels == 1)).sum())
    assert len(ImageNet_hdf5(path, train=False)) == 20
