"""Readers of the reference's tabular datasets (UCI streams, NUS-WIDE, Lending Club) on small files written in the original
formats (reference: data/UCI/data_loader_for_susy_and_ro.py, data/NUS_WIDE/nus_wide_dataset.py,
data/lending_club_loan/lending_club_dataset.py)."""
import os
import pickle

import numpy as np
import pytest

from fedml_b200.data import tabular_readers as T


def test_uci_stream_loader_susy_and_room_occupancy(tmp_path):
    rs = np.random.RandomState(0)
    n, clients = 240, [0, 1, 2, 3]
    centers = rs.randn(4, 18) * 6
    rows = []
    for i in range(n):
        k = i % 4
        rows.append([float(k % 2)] + list(centers[k] + 0.1 * rs.randn(18)))
    susy = tmp_path / "SUSY.csv"
    susy.write_text("\n".join(",".join(f"{v:.6f}" for v in r) for r in rows))
    streams = T.UCIStreamLoader("SUSY", str(susy), clients, n, beta=0.5).load_datastream()
    assert sorted(streams) == clients and all(len(v) == n // 4 for v in streams.values())
    assert streams[0][0]["x"].shape == (18,) and streams[0][0]["y"] in (0, 1)
    # the clustered half is non-IID: the first samples of a client come from one generating centre
    for c in clients:
        first = np.stack([s["x"] for s in streams[c][:20]])
        assert np.linalg.norm(first - first.mean(0), axis=1).max() < 3.0
    # beta = 0: round-robin fill only
    flat = T.UCIStreamLoader("SUSY", str(susy), clients, n, beta=0.0).load_datastream()
    assert [len(flat[c]) for c in clients] == [60] * 4 and np.allclose(flat[0][0]["x"], rows[0][1:], atol=1e-5)
    ro = tmp_path / "ro.csv"
    ro.write_text('"id","date","Temperature","Humidity","Light","CO2","HumidityRatio","Occupancy"\n' + "\n".join(
        f'"{i}","2015-02-04 17:51:00",{23 + i % 3},{27.2},{426 + i},{721.25},{0.0047},{i % 2}' for i in range(40)))
    s2 = T.UCIStreamLoader("RO", str(ro), [0, 1], 40, beta=0.25).load_datastream()
    assert sum(len(v) for v in s2.values()) <= 40 and s2[0][0]["x"].shape == (5,) and set(s["y"] for v in s2.values() for s in v) == {0, 1}


def test_nus_wide_vertical_split_readers(tmp_path):
    pd = pytest.importorskip("pandas")
    rs = np.random.RandomState(1)
    n = 50
    root = tmp_path / "NUS_WIDE"
    (root / "Groundtruth" / "AllLabels").mkdir(parents=True)
    (root / "Groundtruth" / "TrainTestLabels").mkdir(parents=True)
    (root / "Low_Level_Features").mkdir()
    (root / "NUS_WID_Tags").mkdir()
    lab = {"sky": (np.arange(n) % 3 == 0), "water": (np.arange(n) % 3 == 1), "cat": (np.arange(n) % 10 == 2)}
    for k, v in lab.items():
        (root / "Groundtruth" / "AllLabels" / f"Labels_{k}.txt").write_text("\n".join(str(int(b)) for b in v))
        (root / "Groundtruth" / "TrainTestLabels" / f"Labels_{k}_Train.txt").write_text("\n".join(str(int(b)) for b in v))
    for name, width in (("Train_Normalized_CH.dat", 6), ("Train_Normalized_EDH.dat", 4)):
        (root / "Low_Level_Features" / name).write_text("\n".join(" ".join(f"{x:.4f}" for x in rs.rand(width)) + " " for _ in range(n)))
    (root / "NUS_WID_Tags" / "Train_Tags1k.dat").write_text("\n".join("\t".join(str(int(x)) for x in rs.rand(8) > 0.5) + "\t" for _ in range(n)))
    assert set(T.nus_wide_top_k_labels(str(root), 2)) == {"sky", "water"}
    (xa, xb, y), (xat, xbt, yt) = T.nus_wide_load_two_party_data(str(root), ["sky", "water"], neg_label=0)
    keep = int((lab["sky"] ^ lab["water"]).sum())                 # rows with exactly one of the two concepts
    assert xa.shape[0] + xat.shape[0] == keep and xa.shape[1] == 10 and xb.shape[1] == 8 and xa.shape[0] == int(0.8 * keep)
    assert set(np.unique(np.concatenate([y, yt]))) == {0, 1} and abs(float(np.concatenate([xa, xat]).mean())) < 1e-6
    tr3, te3 = T.nus_wide_load_three_party_data(str(root), ["sky", "water"])
    assert [a.shape[1] for a in tr3] == [10, 4, 4, 1] and set(np.unique(tr3[3])) <= {-1, 1}


def test_lending_club_vertical_split_readers(tmp_path):
    pd = pytest.importorskip("pandas")
    rs = np.random.RandomState(2)
    n = 60
    cols = [c for g in T.LOAN_GROUPS.values() for c in g]
    df = pd.DataFrame({c: rs.rand(n) * 100 for c in cols if c != "annual_inc_comp"})
    df["grade"] = rs.choice(list("ABCDEFG"), n)
    df["emp_length"] = rs.choice(["< 1 year", "3 years", "10+ years", None], n)
    df["home_ownership"] = rs.choice(["RENT", "OWN", "MORTGAGE"], n)
    df["verification_status"] = rs.choice(["Verified", "Not Verified"], n)
    df["verification_status_joint"] = rs.choice(["Verified", None], n)
    df["annual_inc"], df["annual_inc_joint"] = rs.rand(n) * 1e5, rs.rand(n) * 2e5
    df["term"] = rs.choice([" 36 months", " 60 months"], n)
    df["initial_list_status"] = rs.choice(["w", "f"], n)
    df["purpose"] = rs.choice(["credit_card", "car", "small_business"], n)
    df["application_type"] = rs.choice(["Individual", "Joint App"], n)
    df["disbursement_method"] = "Cash"
    df["loan_status"] = rs.choice(["Fully Paid", "Charged Off", "Current", "Late (31-120 days)"], n)
    df["issue_d"] = ["Mar-2018" if i % 4 else "Mar-2017" for i in range(n)]
    d = str(tmp_path) + os.sep
    df.to_csv(d + "loan.csv", index=False)
    (xa, xb, y), (xat, xbt, yt) = T.loan_load_two_party_data(d)
    n18 = sum(1 for i in range(n) if i % 4)
    assert xa.shape == (int(0.8 * n18), 15) and xb.shape[1] == 68 and xat.shape[0] == n18 - int(0.8 * n18)
    assert set(np.unique(np.concatenate([y, yt]))) <= {0, 1} and os.path.exists(d + "processed_loan.csv")
    bad = df[df["issue_d"] == "Mar-2018"]["loan_status"].isin(["Charged Off", "Late (31-120 days)"]).sum()
    assert int(np.concatenate([y, yt]).sum()) == int(bad)
    tr3, te3 = T.loan_load_three_party_data(d)                      # second call reads the cached processed table
    assert [a.shape[1] for a in tr3] == [15, 35, 33, 1]


def test_vertical_fl_runs_on_lending_club_files(tmp_path):
    """The SP vertical-FL runner picks up a loan.csv in ``data_cache_dir`` (three parties: 15 / 35 / 33 features)."""
    pd = pytest.importorskip("pandas")
    import fedml
    from fedml_b200.arguments import Arguments
    from fedml_b200.simulation.sp.vertical_fl import VerticalFLAPI
    rs = np.random.RandomState(3)
    n = 200
    cols = [c for g in T.LOAN_GROUPS.values() for c in g]
    df = pd.DataFrame({c: rs.randn(n) for c in cols if c != "annual_inc_comp"})
    signal = df["int_rate"] + df["dti"] - df["total_pymnt"]
    df["loan_status"] = np.where(signal > 0, "Charged Off", "Fully Paid")
    df["annual_inc"], df["annual_inc_joint"] = rs.rand(n) * 1e5, rs.rand(n) * 1e5
    df["verification_status_joint"] = None
    df["issue_d"] = "Jun-2018"
    df.to_csv(tmp_path / "loan.csv", index=False)
    args = fedml.init(Arguments.from_dict(dict(training_type="simulation", federated_optimizer="classical_vertical", dataset="lending_club_loan",
                                               data_cache_dir=str(tmp_path), comm_round=30, epochs=1, batch_size=32, learning_rate=0.1, log_level="WARNING")))
    api = VerticalFLAPI(args)
    assert [x.shape[1] for x in api.xs] == [15, 35, 33] and api.y.shape[0] == 160
    api.train()
    assert api.history[-1]["test_acc"] > 0.65                      # 40 test rows, 83 noisy features


def test_image_dataset_helpers_partition_truncated_sets_and_loaders(tmp_path, monkeypatch):
    """The helper surface of data/cifar10/data_loader.py + datasets.py (same for cifar100 / cinic10)."""
    import torch
    from fedml_b200.data import dataset_utils as U
    monkeypatch.setattr(U, "SYNTHETIC_TRAIN_SIZE", 1500)
    x, y, xt, yt, idx_map, counts = U.partition_data("cifar10", None, "hetero", 6, 0.5)
    assert x.shape == (1500, 32, 32, 3) and x.dtype == np.uint8 and xt.shape[0] == 300
    assert sorted(np.concatenate([np.asarray(v) for v in idx_map.values()]).tolist()) == list(range(1500))
    assert min(len(v) for v in idx_map.values()) >= 10
    assert counts == U.record_net_data_stats(y, idx_map) and sum(sum(c.values()) for c in counts.values()) == 1500
    again = U.partition_data("cifar10", None, "hetero", 6, 0.5)[4]                   # seeded like the reference: reproducible
    assert all(np.array_equal(idx_map[k], again[k]) for k in idx_map)
    homo = U.partition_data("cifar10", None, "homo", 6, 0.5)[4]
    assert sorted(len(v) for v in homo.values()) == [250] * 6
    # frozen split files round-trip (the "hetero-fix" mode), also when written as JSON
    m, d = U.write_net_dataidx_map(str(tmp_path / "net_dataidx_map.txt"), idx_map), U.write_data_distribution(str(tmp_path / "distribution.txt"), counts)
    fixed = U.partition_data("cifar10", None, "hetero-fix", 6, 0.5, dataidx_map_file=m, distribution_file=d)
    assert {k: list(map(int, v)) for k, v in idx_map.items()} == fixed[4] and fixed[5] == counts
    import json
    (tmp_path / "as_json.txt").write_text(json.dumps({str(k): [int(i) for i in v] for k, v in idx_map.items()}, indent=1))
    assert U.read_net_dataidx_map(str(tmp_path / "as_json.txt")) == fixed[4]
    with pytest.raises(ValueError):
        U.partition_data("cifar10", None, "hetero-fix", 6, 0.5)
    # truncated data set + transforms
    ds = U.CIFAR10_truncated(None, dataidxs=idx_map[2], train=True, transform=U.data_transforms("cifar10")[0])
    assert len(ds) == len(idx_map[2]) and np.array_equal(ds.target, y[np.asarray(idx_map[2])])
    img, label = ds[0]
    assert img.shape == (3, 32, 32) and img.dtype == torch.float32 and label == int(ds.target[0])
    ds.truncate_channel([0])
    assert ds.data[0, :, :, 1:].sum() == 0 and x[np.asarray(idx_map[2])[0], :, :, 1:].sum() > 0        # the source is untouched
    g = torch.Generator().manual_seed(1)
    cut = U.Cutout(16, g)(torch.ones(3, 32, 32))
    zeros = int((cut[0] == 0).sum())
    assert 64 <= zeros <= 256 and torch.equal(cut[0], cut[1])                                             # clipped at the border
    plain = U.data_transforms("cifar100")[1](x[0])
    want = (torch.from_numpy(x[0]).permute(2, 0, 1).float() / 255 - torch.tensor(U.S.CIFAR100_MEAN).view(3, 1, 1)) / torch.tensor(U.S.CIFAR100_STD).view(3, 1, 1)
    torch.testing.assert_close(plain, want)
    # loaders: drop_last on both sides, the client's samples only
    tr, te = U.get_dataloader("cifar10", None, 32, 64, idx_map[0])
    assert len(tr) == len(idx_map[0]) // 32 and len(te) == 300 // 64
    xb, yb = next(iter(tr))
    assert xb.shape == (32, 3, 32, 32) and yb.dtype == torch.int64
    tr2, te2 = U.get_dataloader_test_CIFAR10(None, 16, 16, idx_map[1], list(range(40)))
    assert len(te2.dataset) == 40 and len(tr2.dataset) == len(idx_map[1])
    assert U.CIFAR100_truncated(None, train=False).data.shape[0] == 300 and int(U.CIFAR100_truncated(None).target.max()) > 10


def test_per_module_loader_functions_through_the_reference_import_paths(monkeypatch):
    """``fedml.data.<set>.data_loader.get_dataloader`` keeps the reference's calling convention: the dataset argument is ignored,
    the fifth argument is a list of sample indices for the partitioned sets and a client index for the federated ones."""
    import torch
    from fedml_b200.data import dataset_utils as U
    monkeypatch.setattr(U, "SYNTHETIC_TRAIN_SIZE", 1200)
    from fedml.data.cifar100.data_loader import get_dataloader as c100_loader, partition_data as c100_partition
    from fedml.data.FederatedEMNIST.data_loader import DEFAULT_BATCH_SIZE, get_dataloader as femnist_loader
    from fedml.data.fed_shakespeare.data_loader import DEFAULT_TRAIN_FILE, get_dataloader as shakespeare_loader
    from fedml.data.MNIST.data_loader import batch_data
    assert DEFAULT_BATCH_SIZE == 20 and DEFAULT_TRAIN_FILE == "shakespeare_train.h5"
    _, y, _, _, idx_map, counts = c100_partition("ignored", None, "homo", 4, 0.5)
    assert int(y.max()) > 10 and len(idx_map) == 4 and len(counts[0]) > 10
    tr, te = c100_loader("ignored", None, 50, 50, idx_map[1])
    assert len(tr.dataset) == 300 and next(iter(tr))[0].shape == (50, 3, 32, 32)
    tr, te = femnist_loader("ignored", None, 16, 16, 3)
    fed = U._federated("femnist", None)
    assert len(tr.dataset) == fed.client_size(3) and next(iter(tr))[0].shape[1:] == fed.x.shape[1:]
    tr_all, _ = femnist_loader("ignored", None, 16, 16)
    assert len(tr_all.dataset) == len(fed.x)
    tr, _ = shakespeare_loader("ignored", None, 4, 4, 0)
    xb, yb = next(iter(tr))
    assert xb.shape == (4, 80) and yb.shape == (4, 80) and xb.dtype == torch.int64
    batches = batch_data({"x": [[0.0] * 784] * 5, "y": [1, 2, 3, 4, 5]}, 2)
    assert [b[0].shape[0] for b in batches] == [2, 2, 1] and batches[0][0].dtype == torch.float32 and batches[0][1].dtype == torch.int64


def test_lazy_folder_datasets_imagenet_landmarks(tmp_path):
    """File-backed data sets of the large image sets (reference data/ImageNet/datasets.py, data/Landmarks/datasets.py)."""
    import torch
    from PIL import Image
    from fedml.data.ImageNet.datasets import ImageNet, ImageNet_truncated, find_classes, has_file_allowed_extension, make_dataset
    from fedml.data.Landmarks.datasets import Landmarks
    from fedml_b200.data import folder_datasets as FD
    rs = np.random.RandomState(0)
    counts = {"n01": 3, "n02": 2, "n03": 4}
    for split, mult in (("train", 1), ("val", 1)):
        for wnid, n in counts.items():
            d = tmp_path / "imagenet" / split / wnid
            d.mkdir(parents=True)
            for i in range(n * mult):
                Image.fromarray(rs.randint(0, 255, (40 + i, 50, 3), dtype=np.uint8)).save(d / f"img_{i}.JPEG")
            (d / "notes.txt").write_text("not an image")
    root = str(tmp_path / "imagenet")
    classes, class_to_idx = find_classes(os.path.join(root, "train"))
    assert classes == ["n01", "n02", "n03"] and has_file_allowed_extension("a.JPEG", FD.IMG_EXTENSIONS) and not has_file_allowed_extension("a.txt", FD.IMG_EXTENSIONS)
    images, num, ranges = make_dataset(os.path.join(root, "train"), class_to_idx, FD.IMG_EXTENSIONS)
    assert len(images) == 9 and num == {0: 3, 1: 2, 2: 4} and ranges == {0: (0, 3), 1: (3, 5), 2: (5, 9)}
    full = ImageNet(root, train=True)
    assert len(full) == 9 and full.get_data_local_num_dict() == num and full.get_net_dataidx_map() == ranges
    assert len(ImageNet(root, dataidxs=2)) == 4 and len(ImageNet(root, dataidxs=[0, 1])) == 5
    view = ImageNet_truncated(full, [1, 2], full.get_net_dataidx_map(), transform=FD.image_transforms(32)[0])
    x, y = view[0]
    assert len(view) == 6 and x.shape == (3, 32, 32) and x.dtype == torch.float32 and y == 1
    tr, te = FD.get_dataloader_ImageNet(root, 4, 4, dataidxs=[0, 2], image_size=32)
    xb, yb = next(iter(tr))
    assert xb.shape == (4, 3, 32, 32) and set(yb.tolist()) <= {0, 2} and len(te.dataset) == 9
    ev = FD.image_transforms(32)[1](Image.fromarray(np.full((80, 120, 3), 128, dtype=np.uint8)))
    assert ev.shape == (3, 32, 32) and float(ev.std(dim=(1, 2)).max()) < 1e-5                      # resize + centre crop of a flat image stays flat
    # Landmarks: mapping csv -> per-user ranges -> lazy data set
    (tmp_path / "lm" / "images").mkdir(parents=True)
    rows = [("7", "a", 0), ("7", "b", 1), ("3", "c", 1), ("7", "d", 2), ("3", "e", 0)]
    for _, image_id, _ in rows:
        Image.fromarray(rs.randint(0, 255, (30, 30, 3), dtype=np.uint8)).save(tmp_path / "lm" / "images" / f"{image_id}.jpg")
    (tmp_path / "lm" / "train.csv").write_text("user_id,image_id,class\n" + "".join(f"{u},{i},{c}\n" for u, i, c in rows))
    files, per_user, ranges = FD.get_mapping_per_user(str(tmp_path / "lm" / "train.csv"))
    assert [r["image_id"] for r in files] == ["a", "b", "d", "c", "e"] and per_user == {7: 3, 3: 2} and ranges == {7: (0, 3), 3: (3, 5)}
    ds = Landmarks(str(tmp_path / "lm" / "images"), files, dataidxs=ranges[3])
    img, label = ds[0]
    assert len(ds) == 2 and label == 1 and img.shape == (3, 30, 30)
    tr, te = FD.get_dataloader_Landmarks(str(tmp_path / "lm" / "images"), files, files, 2, 2, dataidxs=ranges[7], image_size=16)
    assert len(tr.dataset) == 3 and len(te.dataset) == 5 and next(iter(tr))[0].shape == (2, 3, 16, 16)
    (tmp_path / "lm" / "bad.csv").write_text("user,image\n1,2\n")
    with pytest.raises(ValueError, match="user_id, image_id and class"):
        FD.get_mapping_per_user(str(tmp_path / "lm" / "bad.csv"))
    # generic folder view with explicit indices
    sub = FD.ImageFolderTruncated(os.path.join(root, "train"), dataidxs=[0, 3, 8])
    assert len(sub) == 3 and sub.get_train_labels.tolist() == [0, 1, 2]


def test_fed_cifar100_tff_preprocessing():
    import torch
    from fedml.data.fed_cifar100.utils import cifar100_transform, preprocess_cifar_img
    rs = np.random.RandomState(1)
    img = torch.from_numpy(rs.randint(0, 256, (5, 32, 32, 3)).astype(np.uint8))
    test = preprocess_cifar_img(img, train=False)
    assert test.shape == (5, 3, 24, 24) and test.dtype == torch.float32
    x0 = img[0].double() / 255
    want = ((x0 - x0.mean()) / x0.std()).permute(2, 0, 1)[:, 4:28, 4:28].float()
    torch.testing.assert_close(test[0], want)
    g = torch.Generator().manual_seed(0)
    train = preprocess_cifar_img(img, train=True, generator=g)
    assert train.shape == (5, 3, 24, 24) and not torch.equal(train, test)
    # every training image is some 24x24 window of the standardised image, possibly mirrored
    full = ((x0 - x0.mean()) / x0.std()).permute(2, 0, 1).float()
    windows = [full[:, t:t + 24, l:l + 24] for t in range(9) for l in range(9)]
    assert any(torch.allclose(train[0], w_, atol=1e-5) or torch.allclose(train[0], w_.flip(-1), atol=1e-5) for w_ in windows)
    t = cifar100_transform(0.5, 0.25, train=False, crop_size=(8, 8))
    assert torch.equal(t(torch.full((3, 16, 16), 0.75)), torch.ones(3, 8, 8))


def test_edge_case_sets_from_their_original_files(tmp_path):
    from fedml_b200.data import edge_case as E
    rs = np.random.RandomState(0)
    d = tmp_path / "southwest_cifar10"
    d.mkdir()
    train, test = rs.randint(0, 255, (150, 32, 32, 3)).astype(np.uint8), rs.randint(0, 255, (30, 32, 32, 3)).astype(np.uint8)
    pickle.dump(train, open(d / "southwest_images_new_train.pkl", "wb"))
    pickle.dump(test, open(d / "southwest_images_new_test.pkl", "wb"))
    got = E.get_southwest_dataset(str(tmp_path), "edge-case")
    assert np.array_equal(got[0], train) and np.array_equal(got[1], test)
    assert E.get_southwest_dataset(str(tmp_path), "normal-case") is None
    with pytest.raises(NotImplementedError):
        E.get_southwest_dataset(str(tmp_path), "weird")
    cx, cy = rs.randint(0, 255, (1000, 32, 32, 3)).astype(np.uint8), rs.randint(0, 9, 1000)
    x, y, xt, yt = E.load_edge_case_arrays("cifar10", str(tmp_path), cx, cy, "southwest", "edge-case", seed=1)
    assert x.shape == (500, 32, 32, 3) and (y[400:] == 9).all() and (y[:400] < 9).all() and xt.shape[0] == 30 and (yt == 9).all()
    # the 100 poisoned rows are photos from the file
    flat = {p.tobytes() for p in train}
    assert all(r.tobytes() in flat for r in x[400:])
    # ARDIS csv
    a = tmp_path / "ARDIS"
    a.mkdir()
    for split, n in (("train", 40), ("test", 12)):
        imgs = rs.randint(0, 255, (n, 784)).astype(np.float32)
        labs = np.eye(10)[rs.randint(0, 10, n)]
        labs[:5] = np.eye(10)[7]
        np.savetxt(a / f"ARDIS_{split}_2828.csv", imgs)
        np.savetxt(a / f"ARDIS_{split}_labels.csv", labs)
    sevens = E.get_ardis_dataset(str(tmp_path), "train")
    assert sevens.shape[1:] == (28, 28) and sevens.dtype == np.uint8 and len(sevens) >= 5
    mx, my = rs.rand(600, 784).astype(np.float32), rs.randint(0, 10, 600)
    x, y, xt, yt = E.load_edge_case_arrays("mnist", str(tmp_path), mx, my, "ardis")
    assert x.shape == (400 + len(sevens), 784) and x.dtype == np.float32 and float(x.max()) <= 1.0 and (y[400:] == 1).all() and (yt == 1).all()
    assert E.load_edge_case_arrays("cifar10", str(tmp_path / "nowhere"), cx, cy) is None

    # through the public loader: the attacker's batches come from the files, the targeted test set is the photos labelled 9
    import fedml_b200
    from fedml_b200.arguments import Arguments
    a = Arguments.from_dict(dict(dataset="cifar10", data_cache_dir=str(tmp_path), batch_size=50, client_num_in_total=4, client_num_per_round=2,
                                 synthetic_train_size=800, poison_type="southwest", attack_case="edge-case", log_level="WARNING"))
    poisoned, clean, targeted, n = fedml_b200.data.load_poisoned_dataset_from_edge_case_examples(a)
    assert n == 500 and sum(len(y_) for _, y_ in poisoned) == 500 and sum(len(y_) for _, y_ in targeted) == 30
    assert all((y_ == 9).all() for _, y_ in targeted) and poisoned[0][0].shape[1:] == (3, 32, 32)
