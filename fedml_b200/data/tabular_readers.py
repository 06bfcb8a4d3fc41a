"""Readers for the tabular / feature-partitioned datasets of the reference, from the original files.

* UCI SUSY and Room-Occupancy streams for decentralized online learning (reference data/UCI/data_loader_for_susy_and_ro.py):
  the first ``beta`` fraction of the stream is clustered with k-means, cluster i going to client i (the "adversarial",
  non-IID part), the rest is dealt out so that every client ends with ``sample_num_in_total / n_clients`` samples.
* NUS-WIDE two- / three-party vertical split (data/NUS_WIDE/nus_wide_dataset.py): party A = 634 low-level image features,
  party B = 1k tag features (three-party: the tags cut in two), label = first selected concept vs the others.
* Lending-Club loan table (data/lending_club_loan/lending_club_dataset.py): 2018 loans, good / bad target, 83 standardised
  features grouped into qualification + loan (party A), debt + repayment (B) and multi-account + bad-behaviour (B or C).
All return numpy arrays in the reference's ``[Xa, Xb, (Xc,) y]`` train / test list form (80 / 20 split in file order).
"""
from __future__ import annotations

import csv
import os
from typing import Dict, List, Sequence

import numpy as np


def _standardize(x: np.ndarray) -> np.ndarray:
    x = np.asarray(x, dtype=np.float64)
    mu, sd = x.mean(axis=0), x.std(axis=0)
    return (x - mu) / np.where(sd == 0, 1.0, sd)


def _split_80_20(arrays: Sequence[np.ndarray]):
    n_train = int(0.8 * arrays[0].shape[0])
    return [a[:n_train] for a in arrays], [a[n_train:] for a in arrays]


# ------------------------------------------------------------------------------------------------ UCI streams
def _kmeans_labels(x: np.ndarray, k: int, seed: int = 0, iters: int = 50) -> np.ndarray:
    try:
        from sklearn.cluster import KMeans
        return KMeans(n_clusters=k, n_init=10, random_state=seed).fit(x).labels_
    except ImportError:                                   # plain Lloyd iterations
        rs = np.random.RandomState(seed)
        cent = x[rs.choice(len(x), k, replace=False)].astype(np.float64)
        lab = np.zeros(len(x), dtype=np.int64)
        for _ in range(iters):
            d = ((x[:, None, :] - cent[None, :, :]) ** 2).sum(-1)
            new = d.argmin(1)
            if np.array_equal(new, lab) and _:
                break
            lab = new
            for j in range(k):
                if (lab == j).any():
                    cent[j] = x[lab == j].mean(0)
        return lab


class UCIStreamLoader:
    """``load_datastream()`` -> ``{client_id: [{"x": features, "y": label}, ...]}``.

    ``data_name``: "SUSY" (label in column 0, 18 features) or "RO" (room occupancy: columns 2..-2 are features, the last one
    the label).  ``beta`` in [0, 1]: fraction of the first ``sample_num_in_total`` rows that is assigned by k-means cluster."""

    def __init__(self, data_name: str, data_path: str, client_list: Sequence[int], sample_num_in_total: int, beta: float):
        self.data_name, self.data_path = data_name, data_path
        self.client_list, self.total, self.beta = list(client_list), int(sample_num_in_total), float(beta)
        self.x: List[np.ndarray] = []
        self.y: List[int] = []
        self.streams: Dict[int, List[dict]] = {}

    def _read(self):
        with open(self.data_path, newline="") as fh:
            for i, row in enumerate(csv.reader(fh)):
                if i >= self.total:
                    break
                try:
                    if self.data_name == "SUSY":
                        feat, lab = row[1:], row[0]
                    else:
                        feat, lab = row[2:-1], row[-1]
                    self.x.append(np.asarray(feat, dtype=np.float32))
                    self.y.append(int(float(lab)))
                except ValueError:                        # header line of the room-occupancy files
                    continue

    def load_datastream(self) -> Dict[int, List[dict]]:
        self._read()
        n_clients = len(self.client_list)
        self.streams = {c: [] for c in self.client_list}
        n_adv = int(self.total * self.beta) if self.beta > 0 else 0
        n_adv = min(n_adv, len(self.x))
        if n_adv:
            labels = _kmeans_labels(np.stack(self.x[:n_adv]), n_clients)
            for i, cl in enumerate(labels):
                self.streams[self.client_list[int(cl)]].append({"x": self.x[i], "y": self.y[i]})
        per_client = self.total // n_clients
        pool = [{"x": self.x[i], "y": self.y[i]} for i in range(n_adv, len(self.x))]
        for c in self.client_list:                        # clusters larger than a client's share give their surplus back
            if len(self.streams[c]) > per_client:
                pool.extend(self.streams[c][per_client:])
                self.streams[c] = self.streams[c][:per_client]
        it = iter(pool)
        for c in self.client_list:
            while len(self.streams[c]) < per_client:
                nxt = next(it, None)
                if nxt is None:
                    return self.streams
                self.streams[c].append(nxt)
        return self.streams


# ------------------------------------------------------------------------------------------------ NUS-WIDE
def nus_wide_top_k_labels(data_dir: str, top_k: int = 5) -> List[str]:
    import pandas as pd
    d = os.path.join(data_dir, "Groundtruth", "AllLabels")
    counts = {}
    for fn in os.listdir(d):
        path = os.path.join(d, fn)
        if os.path.isfile(path):
            counts[os.path.splitext(fn)[0].split("_")[-1]] = int((pd.read_csv(path, header=None)[0] == 1).sum())
    return [k for k, _ in sorted(counts.items(), key=lambda kv: kv[1], reverse=True)[:top_k]]


def _nus_wide_raw(data_dir: str, selected_labels: Sequence[str], n_samples: int, split: str):
    import pandas as pd
    lab_dir = os.path.join(data_dir, "Groundtruth", "TrainTestLabels")
    labels = pd.concat([pd.read_csv(os.path.join(lab_dir, f"Labels_{lb}_{split}.txt"), header=None).rename(columns={0: lb})
                        for lb in selected_labels], axis=1)
    chosen = labels[labels.sum(axis=1) == 1] if len(selected_labels) > 1 else labels     # exactly one of the concepts
    feat_dir = os.path.join(data_dir, "Low_Level_Features")
    parts = [pd.read_csv(os.path.join(feat_dir, fn), header=None, sep=" ").dropna(axis=1)
             for fn in sorted(os.listdir(feat_dir)) if fn.startswith(f"{split}_Normalized")]
    xa = pd.concat(parts, axis=1).loc[chosen.index].values
    tags = pd.read_csv(os.path.join(data_dir, "NUS_WID_Tags", f"{split}_Tags1k.dat"), header=None, sep="\t").dropna(axis=1)
    xb = tags.loc[chosen.index].values
    y = chosen.values
    if n_samples != -1:
        xa, xb, y = xa[:n_samples], xb[:n_samples], y[:n_samples]
    return xa, xb, y


def nus_wide_load_two_party_data(data_dir, selected_labels, neg_label=-1, n_samples=-1):
    xa, xb, y = _nus_wide_raw(data_dir, selected_labels, n_samples, "Train")
    xa, xb = _standardize(xa), _standardize(xb)
    y = np.where(y[:, 0] == 1, 1, neg_label).reshape(-1, 1)
    return _split_80_20([xa, xb, y])


def nus_wide_load_three_party_data(data_dir, selected_labels, neg_label=-1, n_samples=-1):
    xa, xb, y = _nus_wide_raw(data_dir, selected_labels, n_samples, "Train")
    half = xb.shape[1] // 2
    xa, xb1, xb2 = _standardize(xa), _standardize(xb[:, :half]), _standardize(xb[:, half:])
    y = np.where(y[:, 0] == 1, 1, neg_label).reshape(-1, 1)
    return _split_80_20([xa, xb1, xb2, y])


# ------------------------------------------------------------------------------------------------ Lending Club
LOAN_GROUPS = {
    "qualification": "grade emp_length home_ownership annual_inc_comp verification_status total_rev_hi_lim tot_hi_cred_lim "
                     "total_bc_limit total_il_high_credit_limit",
    "loan": "loan_amnt term initial_list_status purpose application_type disbursement_method",
    "debt": "int_rate installment revol_bal revol_util out_prncp recoveries dti dti_joint tot_coll_amt mths_since_rcnt_il "
            "total_bal_il il_util max_bal_bc all_util bc_util total_bal_ex_mort revol_bal_joint mo_sin_old_il_acct "
            "mo_sin_old_rev_tl_op mo_sin_rcnt_rev_tl_op mort_acc num_rev_tl_bal_gt_0 percent_bc_gt_75",
    "repayment": "num_sats num_bc_sats pct_tl_nvr_dlq bc_open_to_buy last_pymnt_amnt total_pymnt total_pymnt_inv "
                 "total_rec_prncp total_rec_int total_rec_late_fee tot_cur_bal avg_cur_bal",
    "multi_acc": "num_il_tl num_op_rev_tl num_rev_accts num_actv_rev_tl num_tl_op_past_12m open_rv_12m open_rv_24m open_acc_6m "
                 "open_act_il open_il_12m open_il_24m total_acc inq_last_6mths open_acc inq_fi inq_last_12m acc_open_past_24mths",
    "mal_behavior": "num_tl_120dpd_2m num_tl_30dpd num_tl_90g_dpd_24m pub_rec_bankruptcies mths_since_recent_revol_delinq "
                    "num_accts_ever_120_pd mths_since_recent_bc_dlq chargeoff_within_12_mths collections_12_mths_ex_med "
                    "mths_since_last_major_derog acc_now_delinq pub_rec mths_since_last_delinq delinq_2yrs delinq_amnt tax_liens",
}
LOAN_GROUPS = {k: v.split() for k, v in LOAN_GROUPS.items()}
_BAD_STATUS = {"Charged Off", "Default", "Does not meet the credit policy. Status:Charged Off", "In Grace Period",
               "Late (16-30 days)", "Late (31-120 days)"}
_CATEGORY_CODES = {
    "grade": dict(zip("ABCDEFG", range(6, -1, -1))),
    "emp_length": {"< 1 year": 1, "1 year": 2, "2 years": 2, "3 years": 2, "4 years": 3, "5 years": 3, "6 years": 3,
                   "7 years": 4, "8 years": 4, "9 years": 4, "10+ years": 5},
    "home_ownership": {"RENT": 0, "MORTGAGE": 1, "OWN": 2, "ANY": 3, "NONE": 3, "OTHER": 3},
    "verification_status": {"Not Verified": 0, "Source Verified": 1, "Verified": 2},
    "term": {" 36 months": 0, " 60 months": 1},
    "initial_list_status": {"w": 0, "f": 1},
    "purpose": {"debt_consolidation": 0, "credit_card": 0, "small_business": 1, "educational": 2},     # everything else: 3
    "application_type": {"Individual": 0, "Joint App": 1},
    "disbursement_method": {"Cash": 0, "DirectPay": 1},
}


def loan_prepare(csv_path: str, year: int = 2018):
    """loan.csv -> DataFrame of the ``year`` loans with the target, the composite annual income and coded categoricals."""
    import pandas as pd
    df = pd.read_csv(csv_path, low_memory=False)
    df["target"] = df["loan_status"].isin(_BAD_STATUS).astype(int)                    # 1 = bad loan
    joint = df["verification_status"] == df.get("verification_status_joint")
    df["annual_inc_comp"] = np.where(joint, df.get("annual_inc_joint"), df["annual_inc"])
    df["issue_year"] = pd.to_datetime(df["issue_d"]).dt.year
    for col, codes in _CATEGORY_CODES.items():
        if col in df:
            default = 3 if col == "purpose" else (0 if col == "emp_length" else np.nan)
            df[col] = df[col].map(lambda v, c=codes, d=default: c.get(v, d))
    return df[df["issue_year"] == year]


def loan_process(df):
    import pandas as pd
    cols = [c for g in LOAN_GROUPS.values() for c in g]
    feats = df[cols].fillna(-99)
    out = pd.DataFrame(_standardize(feats.values), columns=cols, index=df.index)
    out["target"] = df["target"].values
    return out


def loan_load_processed(data_dir: str):
    import pandas as pd
    cached = os.path.join(data_dir, "processed_loan.csv")
    if os.path.exists(cached):
        return pd.read_csv(cached, low_memory=False)
    df = loan_process(loan_prepare(os.path.join(data_dir, "loan.csv")))
    df.to_csv(cached, index=False)
    return df


def _loan_parties(data_dir, groups: Sequence[Sequence[str]]):
    df = loan_load_processed(data_dir)
    xs = [df[[c for g in gs for c in LOAN_GROUPS[g]]].values for gs in groups]
    return _split_80_20(xs + [df["target"].values.reshape(-1, 1)])


def loan_load_two_party_data(data_dir):
    return _loan_parties(data_dir, [("qualification", "loan"), ("debt", "repayment", "multi_acc", "mal_behavior")])


def loan_load_three_party_data(data_dir):
    return _loan_parties(data_dir, [("qualification", "loan"), ("debt", "repayment"), ("multi_acc", "mal_behavior")])


# ------------------------------------------------------------------------------------------------ entry point for the VFL runners
def load_vertical_dataset(args):
    """``(party_dims, (xs, y), (xs_test, y_test))`` as torch tensors for the vertical-FL runners: the real Lending-Club /
    NUS-WIDE files when ``args.data_cache_dir`` holds them (``dataset: lending_club_loan | NUS_WIDE``, ``vfl_party_num`` 2 or 3,
    ``nus_wide_labels`` optional), else the synthetic task of the same form (no network here to download anything)."""
    import torch
    from .synthetic import make_vertical_tabular
    name = str(getattr(args, "dataset", "lending_club")).lower()
    root = str(getattr(args, "data_cache_dir", "") or "")
    parties = int(getattr(args, "vfl_party_num", 0) or 0)
    loaded = None
    if root and "lending" in name and (os.path.exists(os.path.join(root, "loan.csv")) or os.path.exists(os.path.join(root, "processed_loan.csv"))):
        d = root if root.endswith(os.sep) else root + os.sep
        loaded = loan_load_two_party_data(d) if parties == 2 else loan_load_three_party_data(d)
    elif root and "nus" in name and os.path.isdir(os.path.join(root, "Groundtruth", "TrainTestLabels")):
        labels = list(getattr(args, "nus_wide_labels", None) or nus_wide_top_k_labels(root, 2))
        n = int(getattr(args, "vfl_samples", -1) or -1)
        loaded = (nus_wide_load_two_party_data(root, labels, 0, n) if parties == 2
                  else nus_wide_load_three_party_data(root, labels, 0, n))
    if loaded is None:
        dims = tuple(getattr(args, "vfl_party_dims", (32, 24, 24)))
        (xs, y), (xt, yt) = make_vertical_tabular(int(getattr(args, "vfl_samples", 4000)), dims, int(getattr(args, "random_seed", 0)), name)
        return dims, (xs, y), (xt, yt)
    train, test = loaded
    to_x = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))                       # noqa: E731
    to_y = lambda a: torch.from_numpy((np.asarray(a).reshape(-1) > 0).astype(np.float32))              # noqa: E731
    xs, xt = [to_x(a) for a in train[:-1]], [to_x(a) for a in test[:-1]]
    return tuple(int(x.shape[1]) for x in xs), (xs, to_y(train[-1])), (xt, to_y(test[-1]))
