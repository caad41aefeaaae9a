"""CPU oracle for Kmix -- TEST INFRASTRUCTURE ONLY, never imported by the product.

Restates ``Kmix`` (/root/reference/src/augmentations/augmentations.py:119-189) with the bank selection
written out explicitly.  Pinned: ``tests/test_oracle_kmix.py`` replays the reference's own class (importable
here: torch + numpy only) on a synthetic centroid file and demands identical picks and outputs, and
``tests/golden/kmix_reference.npz`` holds picks/outputs generated from the reference by
``tests/golden/make_golden.py``.

Randomness, as in the reference: ``np.random.random()`` for alpha, then ``np.random.randint(n)`` with
``n = len(candidates)`` once the bank holds 128 entries, ``n = len(bank)`` before.
"""
from __future__ import annotations

import numpy as np
import torch

F32_EPS = torch.finfo(torch.float32).eps


def log_mixup_exp(xa, xb, alpha):
    """augmentations.py:8-12."""
    xa = xa.exp()
    xb = xb.exp()
    x = alpha * xa + (1. - alpha) * xb
    return torch.log(x + F32_EPS)


def band_means(x):
    """``x.squeeze(0).T.mean(dim=0)`` (:148, :154): per-band mean over time of a [1, F, T] clip."""
    return x.squeeze(0).T.mean(dim=0)


def cluster_ids(x, centroids_unit):
    """(cluster as a bank entry, cluster as the incoming clip) of one [1, F, T] clip."""
    avg = band_means(x)
    unit = avg / avg.norm(dim=-1, keepdim=True)
    mem = int(torch.argmin(torch.cdist(unit.unsqueeze(0), centroids_unit, p=2), dim=1)[0])
    point = int(torch.argmin(torch.cdist(avg.unsqueeze(0), centroids_unit, p=2), dim=1)[0])
    return mem, point


def far_order(centroids_unit):
    """:152 -- for every centroid, all centroids sorted by DEcreasing distance (torch.topk order)."""
    k = centroids_unit.shape[0]
    return torch.topk(torch.cdist(centroids_unit, centroids_unit, p=2), k=k, dim=1).indices.numpy()


def candidates(bank_clusters, point_cluster, order):
    """:155-164 -- bank positions (FIFO order) whose cluster is the farthest non-empty one from the clip's,
    at most the first 128."""
    bank_clusters = np.asarray(bank_clusters)
    for tgt in order[point_cluster]:
        idx = np.nonzero(bank_clusters == tgt)[0]
        if idx.size:
            return idx[:128]
    return np.zeros(0, np.int64)


class KmixOracle:
    def __init__(self, centroids, ratio=0.4, n_memory=2048, use_log_mixup=True, top_k=1, np_rng=np.random):
        c = torch.as_tensor(centroids, dtype=torch.float32)
        self.centroids_unit = c / c.norm(dim=-1, keepdim=True)         # :146
        self.order = far_order(self.centroids_unit)
        self.ratio, self.n, self.use_log_mixup, self.top_k = ratio, n_memory, use_log_mixup, top_k
        self.np_rng = np_rng
        self.bank, self.bank_cluster = [], []
        self.records = []            # (alpha, bank position or -1, number of candidates)

    def __call__(self, x):
        x = torch.as_tensor(x, dtype=torch.float32)
        alpha = self.ratio * self.np_rng.random()
        mem, point = cluster_ids(x, self.centroids_unit)
        if self.bank:
            if len(self.bank) >= 128:
                cand = candidates(self.bank_cluster, point, self.order)
                j = int(cand[self.np_rng.randint(len(cand))])
                n_cand = len(cand)
            else:
                j = int(self.np_rng.randint(len(self.bank)))
                n_cand = len(self.bank)
            z = self.bank[j]
            mixed = log_mixup_exp(x, z, 1. - alpha) if self.use_log_mixup else alpha * z + (1. - alpha) * x
        else:
            j, n_cand, mixed = -1, 0, x
        self.bank = (self.bank + [x])[-self.n:]
        self.bank_cluster = (self.bank_cluster + [mem])[-self.n:]
        self.records.append((float(alpha), j, n_cand))
        return mixed.to(torch.float)
