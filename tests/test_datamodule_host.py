"""Host logic of the device-side batch assembly, no GPU needed: epoch order equals the
reference's ``DataLoader(shuffle=True, drop_last=True)`` under the same torch seed, packing metadata."""
import numpy as np
import torch
from torch.utils.data import DataLoader

import audiossl_b200 as A


class _Indices(torch.utils.data.Dataset):
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return i


def test_epoch_order_equals_torch_dataloader():
    n, bs = 37, 8
    for seed in (0, 31):
        torch.manual_seed(seed)
        dl = DataLoader(_Indices(n), shuffle=True, batch_size=bs, drop_last=True, num_workers=0)
        want = [[b.tolist() for b in dl] for _ in range(2)]                 # two epochs on one generator stream
        torch.manual_seed(seed)
        loader = A.DeviceBatchLoader(A.WaveformDataset([np.zeros(4, np.float32)] * n), pipeline=None, batch_size=bs)
        got = [loader.batches() for _ in range(2)]
        assert got == want and got[0] != got[1] and len(loader) == n // bs


def test_sequential_and_keep_last():
    loader = A.DeviceBatchLoader(A.WaveformDataset([np.zeros(3, np.float32)] * 10), None, batch_size=4, shuffle=False,
                                 drop_last=False)
    assert loader.batches() == [[0, 1, 2, 3], [4, 5, 6, 7], [8, 9]]


def test_packed_offsets_and_dataset_views():
    p = A.PackedClips(raw=torch.zeros(10), lengths=np.array([3, 0, 7], np.int64))
    assert p.offsets.tolist() == [0, 3, 3] and len(p) == 3
    ds = A.WaveformDataset(["a", "b"], load=lambda path: torch.ones(5, dtype=torch.float64) * (1 if path == "a" else 2))
    assert ds[1].dtype == np.float32 and ds[1].tolist() == [2.0] * 5 and len(ds) == 2


def test_no_gpu_no_batches():
    if torch.cuda.is_available():
        return
    loader = A.DeviceBatchLoader(A.WaveformDataset([np.zeros(3, np.float32)] * 4), None, batch_size=2)
    try:
        next(iter(loader))
    except RuntimeError as e:
        assert "CUDA" in str(e)
    else:
        raise AssertionError("the loader must not produce batches without a GPU")
