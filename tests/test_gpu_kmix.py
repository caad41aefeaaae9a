"""GPU parity of the Kmix drop-in against the fixture generated from the reference's own class
(tests/golden/make_golden_kmix.py): bit-exact bank picks and candidate counts, values within the path tolerance."""
import os

import numpy as np
import pytest
import torch

from oracle import kmix as okmix

pytestmark = pytest.mark.gpu


def _close(got, want):
    assert (np.abs(got - want) <= 1e-5 + 1e-4 * np.abs(want)).all(), np.abs(got - want).max()


def test_cluster_ids_match_oracle(golden_dir):
    import audiossl_b200 as A
    g = np.load(os.path.join(golden_dir, "kmix_reference.npz"))
    mod = A.Kmix(ratio=0.4, n_memory=200, log_mixup_exp=True, top_k=1, centroids=g["centroids"])
    x = torch.from_numpy(g["x"][:, 0]).cuda()
    mem, point = mod._clusters(x)
    unit = torch.from_numpy(g["centroids"])
    unit = unit / unit.norm(dim=-1, keepdim=True)
    want = np.array([okmix.cluster_ids(torch.from_numpy(c), unit) for c in g["x"]])
    assert np.array_equal(mem, want[:, 0]) and np.array_equal(point, want[:, 1])
    assert len(set(mem.tolist())) > 1


@pytest.mark.parametrize("chunk", [1, 50, 330])
def test_kmix_replays_reference(golden_dir, chunk):
    """Clip by clip, in batches that straddle the 128-entry switch and the FIFO roll-over, and as one batch."""
    import audiossl_b200 as A
    g = np.load(os.path.join(golden_dir, "kmix_reference.npz"))
    x = torch.from_numpy(g["x"])
    np.random.seed(int(g["seed"]))
    mod = A.Kmix(ratio=0.4, n_memory=int(g["n_memory"]), log_mixup_exp=True, top_k=1, centroids=g["centroids"])
    outs, picks, ncand, alphas = [], [], [], []
    for lo in range(0, x.shape[0], chunk):
        part = x[lo:lo + chunk]
        out = mod(part[0] if chunk == 1 else part)
        outs.append(out.unsqueeze(0) if chunk == 1 else out)
        picks.append(mod.last_params.bank_index[:, 0]); ncand.append(mod.last_candidates); alphas.append(mod.last_params.alpha[:, 0])
    got = torch.cat(outs).cpu().numpy()
    assert np.array_equal(np.concatenate(picks), g["pick"])
    assert np.array_equal(np.concatenate(ncand), g["n_candidates"])
    assert np.array_equal(np.concatenate(alphas), g["alpha"])
    assert got.shape == g["mixed"].shape
    _close(got, g["mixed"])
    assert mod.memory_bank_size == int(g["n_memory"])


def test_kmix_rejects_what_the_reference_cannot_run():
    import audiossl_b200 as A
    c = np.ones((3, 8), np.float32)
    with pytest.raises(ValueError):
        A.Kmix(top_k=None, centroids=c)
    with pytest.raises(ValueError):
        A.Kmix(top_k=200, centroids=c)
    with pytest.raises(ValueError):
        A.Kmix(top_k=1)
    with pytest.raises(ValueError):
        A.Kmix(top_k=1, centroids=c)(torch.zeros(1, 9, 5))
