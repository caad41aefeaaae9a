"""Generate ``kmix_reference.npz`` from the reference's OWN ``Kmix`` class
(/root/reference/src/augmentations/augmentations.py:119-189).  Run HERE (the build container):

    python tests/golden/make_golden_kmix.py

The reference ships no centroid file (every config has ``centroid_path: None``), so a synthetic centroid
tensor is written to a temporary file and loaded through the class's own ``torch.load``.  The fixture holds
the centroids, the input clips, the mixed outputs and -- recorded by replaying the same seed through
``oracle.kmix.KmixOracle``, asserted bit-identical to the reference's outputs first -- alpha, the bank
position picked and the number of candidates per clip.  The bank is small (n_memory 200) but larger than 128,
so both branches of Kmix.forward (:172-176) and the FIFO roll-over are exercised.
"""
import importlib
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle import kmix  # noqa: E402


def make_inputs(seed=77, n_clips=330, f=16, t=20, k=6):
    g = torch.Generator().manual_seed(seed)
    centroids = torch.randn(k, f, generator=g) * 1.5 + 0.5
    # clips scattered around the centroids so every cluster fills up, plus per-clip offsets
    which = torch.randint(0, k, (n_clips,), generator=g)
    x = torch.randn(n_clips, 1, f, t, generator=g) * 0.8 + centroids[which][:, None, :, None] * 1.2
    return centroids.to(torch.float32), x.to(torch.float32)


def main():
    ref = importlib.import_module("src.augmentations.augmentations")
    centroids, x = make_inputs()
    with tempfile.NamedTemporaryFile(suffix=".pt", delete=False) as fh:
        torch.save(centroids, fh.name)
        path = fh.name
    try:
        np.random.seed(31)
        mod = ref.Kmix(ratio=0.4, n_memory=200, log_mixup_exp=True, top_k=1, centroid_path=path)
        want = torch.stack([mod(c.clone()) for c in x]).numpy()
    finally:
        os.unlink(path)
    np.random.seed(31)
    orc = kmix.KmixOracle(centroids, ratio=0.4, n_memory=200, use_log_mixup=True, top_k=1)
    got = torch.stack([orc(c) for c in x]).numpy()
    assert np.array_equal(got, want), "oracle != reference"
    rec = np.array(orc.records, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "kmix_reference.npz"), centroids=centroids.numpy(), x=x.numpy(), mixed=want,
                        alpha=rec[:, 0].astype(np.float32), pick=rec[:, 1].astype(np.int32),
                        n_candidates=rec[:, 2].astype(np.int32), seed=31, n_memory=200)
    print("kmix_reference.npz: picks", int((rec[:, 1] >= 0).sum()), "candidate-branch clips", int((np.arange(len(rec)) > 128).sum()))


if __name__ == "__main__":
    main()
