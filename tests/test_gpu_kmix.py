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


def test_augmentation_module_chains_kmix_after_the_crop(golden_dir, tmp_path):
    """config['pretrain']['augmentations']['Kmix'] with a centroid file: applied to the views like the reference's
    nn.Sequential(MixupBYOLA, RandomResizeCrop, Kmix), in the order view 1, view 2 of each clip."""
    import random
    import audiossl_b200 as A
    g = np.load(os.path.join(golden_dir, "kmix_reference.npz"))
    path = str(tmp_path / "centroids.pt")
    torch.save(torch.from_numpy(g["centroids"]), path)
    aug = {"MixupBYOLA": {"ratio": 0.4, "log_mixup_exp": True},
           "RandomResizeCrop": {"virtual_crop_scale": [1.0, 1.5], "freq_crop_scale": [0.6, 1.5], "time_crop_scale": [0.6, 1.5]}}
    cfg = {"pretrain": {"normalization": "mean_var", "augmentations": dict(aug)}}
    cfg_k = {"pretrain": {"normalization": "mean_var",
                          "augmentations": dict(aug, Kmix={"ratio": 0.3, "log_mixup_exp": True, "top_k": 1, "centroid_path": path})}}
    x = torch.from_numpy(g["x"][:150]).cuda()                      # [150, 1, 16, 20]: crosses the 128-entry switch
    np.random.seed(5); random.seed(5)
    plain = A.AugmentationModule(cfg, 1000)
    a1, a2 = plain(x)
    km = A.Kmix(ratio=0.3, log_mixup_exp=True, top_k=1, centroid_path=path)
    want = km(torch.stack((a1, a2), 1).reshape(-1, *a1.shape[1:])).reshape(a1.shape[0], 2, *a1.shape[1:])
    np.random.seed(5); random.seed(5)
    chained = A.AugmentationModule(cfg_k, 1000)
    b1, b2 = chained(x)
    assert torch.equal(b1, want[:, 0]) and torch.equal(b2, want[:, 1])
    assert not torch.equal(b1, a1)
    with pytest.warns(UserWarning):
        A.AugmentationModule({"pretrain": {"normalization": "mean_var", "augmentations": dict(
            aug, Kmix={"ratio": 0.3, "log_mixup_exp": True, "top_k": 1, "centroid_path": None})}}, 10)
