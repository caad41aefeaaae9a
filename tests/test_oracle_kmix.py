"""Pin ``oracle.kmix`` to the reference's Kmix (augmentations.py:119-189): golden fixture generated from the
reference's own class (tests/golden/make_golden_kmix.py) and, when /root/reference is mounted, a live replay."""
import os
import tempfile

import numpy as np
import torch

from oracle import kmix


def test_kmix_matches_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "kmix_reference.npz"))
    np.random.seed(int(g["seed"]))
    orc = kmix.KmixOracle(g["centroids"], ratio=0.4, n_memory=int(g["n_memory"]), use_log_mixup=True, top_k=1)
    got = torch.stack([orc(torch.from_numpy(c)) for c in g["x"]]).numpy()
    assert np.array_equal(got, g["mixed"])
    rec = np.array(orc.records)
    assert np.array_equal(rec[:, 1].astype(np.int32), g["pick"])
    assert np.array_equal(rec[:, 2].astype(np.int32), g["n_candidates"])
    assert g["pick"][0] == -1 and (g["pick"][1:] >= 0).all()
    assert (g["n_candidates"][129:] <= 128).all() and g["n_candidates"][129:].min() >= 1


def test_candidates_are_the_farthest_non_empty_cluster():
    c = torch.tensor([[1.0, 0.0], [0.0, 1.0], [-1.0, 0.0]])
    order = kmix.far_order(c / c.norm(dim=-1, keepdim=True))
    assert order[0][0] == 2 and order[0][-1] == 0               # farthest first, itself last
    bank = [0, 1, 1, 0, 1]
    assert kmix.candidates(bank, 0, order).tolist() == [1, 2, 4]   # cluster 2 is empty -> next farthest, FIFO order
    assert kmix.candidates([0, 0], 0, order).tolist() == [0, 1]    # only its own cluster is populated
    assert len(kmix.candidates([1] * 300, 0, order)) == 128        # l[:128]


def test_kmix_live_replay_linear_mix(reference_aug):
    """Same seed through the reference class and the oracle, linear mix branch (log_mixup_exp=False)."""
    import importlib
    ref = importlib.import_module("src.augmentations.augmentations")
    g = torch.Generator().manual_seed(5)
    centroids = torch.randn(4, 12, generator=g)
    x = torch.randn(180, 1, 12, 9, generator=g) + centroids[torch.randint(0, 4, (180,), generator=g)][:, None, :, None]
    with tempfile.NamedTemporaryFile(suffix=".pt", delete=False) as fh:
        torch.save(centroids, fh.name)
    try:
        np.random.seed(3)
        mod = ref.Kmix(ratio=0.3, n_memory=150, log_mixup_exp=False, top_k=4, centroid_path=fh.name)
        want = torch.stack([mod(c.clone()) for c in x])
    finally:
        os.unlink(fh.name)
    np.random.seed(3)
    orc = kmix.KmixOracle(centroids, ratio=0.3, n_memory=150, use_log_mixup=False, top_k=4)
    got = torch.stack([orc(c) for c in x])
    assert torch.equal(got, want)
