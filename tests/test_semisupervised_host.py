"""CPU: host logic of the semi-supervised data path -- `validate_data_split` known answers, ratio-preserving
split, the two-stream loader's cycling / subsampling semantics (reference
src/scvi/dataloaders/_data_splitting.py:24-92,342-586, _semi_dataloader.py, _concat_dataloader.py)."""
import numpy as np
import pytest
import torch

from scvi_tools_b200.semisupervised import (SemiSupervisedDataLoader, SemiSupervisedDataSplitter,
                                            validate_data_split)


def test_validate_data_split_known_answers():
    # worked from the reference function by hand (ceil / floor arithmetic and the small-last-batch rule)
    assert validate_data_split(1000, 0.9) == (900, 100)
    assert validate_data_split(1000, 0.9, 0.05) == (900, 50)
    assert validate_data_split(1001, 0.9, None, 128) == (901, 100)          # 901 % 128 = 5: untouched
    assert validate_data_split(1139, 0.9, None, 128) == (1024, 115)         # ceil(1025.1) = 1026, tail 2 moves to val
    assert validate_data_split(1139, 0.9, None, 128, drop_last=True) == (1026, 113)
    with pytest.warns(UserWarning):                                          # explicit train_size: warn, keep
        assert validate_data_split(1139, 0.9, None, 128, train_size_is_none=False) == (1026, 113)
    for bad in ((10, 1.5), (10, 0.0)):
        with pytest.raises(ValueError):
            validate_data_split(*bad)
    with pytest.raises(ValueError):
        validate_data_split(10, 0.9, 0.5)
    with pytest.raises(ValueError):
        validate_data_split(10, 0.9, 1.0)


def _labels(n=1000, n_labels=5, frac_unlabelled=0.7, seed=0):
    rng = np.random.default_rng(seed)
    lab = rng.integers(0, n_labels, size=n)
    lab[rng.random(n) < frac_unlabelled] = n_labels
    return lab


def test_splitter_preserves_ratio_and_partitions_the_cells():
    lab = _labels()
    s = SemiSupervisedDataSplitter(lab, 5, train_size=0.8, validation_size=0.1, batch_size=64, seed=3)
    allidx = np.concatenate([s.train_idx, s.val_idx, s.test_idx])
    assert np.array_equal(np.sort(allidx), np.arange(1000))
    n_lab = int((lab != 5).sum())
    for idx, frac in ((s.train_idx, 0.8), (s.val_idx, 0.1)):
        got = int((lab[idx] != 5).sum())
        assert abs(got - frac * n_lab) <= 1.0
    # the shuffles are the reference's: RandomState(seed).choice(idx, len(idx), replace=False) per label group
    li = np.argwhere(lab != 5).ravel()
    perm = np.random.RandomState(seed=3).choice(li, len(li), replace=False)
    n_val = int(np.floor(len(li) * 0.1))
    assert np.array_equal(s.val_idx[:n_val], perm[:n_val])
    # unshuffled: sequential order
    s2 = SemiSupervisedDataSplitter(lab, 5, train_size=0.5, shuffle_set_split=False, batch_size=64)
    assert np.array_equal(s2.val_idx[: len(li) - int(np.ceil(0.5 * len(li)))], li[: len(li) - int(np.ceil(0.5 * len(li)))])


class _FakeData:
    class _X:
        device = torch.device("cpu")

    X = _X()

    def minibatch(self, idx):
        return {"idx": idx}


def test_loader_cycles_the_smaller_stream_and_subsamples_labels():
    lab = _labels(n=1000, frac_unlabelled=0.9, seed=1)
    dl = SemiSupervisedDataLoader(_FakeData(), lab, 5, shuffle=True, batch_size=100, seed=0)
    n_lab = int((lab != 5).sum())
    assert len(dl) == 10 and len(dl.labelled_idx) == n_lab
    pairs = dl.index_pairs()
    assert len(pairs) == 10
    full = torch.cat([f for f, _ in pairs])
    assert np.array_equal(np.sort(full.numpy()), np.arange(1000))            # the full stream covers every cell once
    n_lab_batches = -(-n_lab // 100)
    for i, (_, l) in enumerate(pairs):                                       # the labelled stream repeats its first pass
        assert torch.equal(l, pairs[i % n_lab_batches][1])
        assert (lab[l.numpy()] != 5).all()
    # per-label subsampling: at most k cells per label, redrawn on resample
    dl2 = SemiSupervisedDataLoader(_FakeData(), lab, 5, shuffle=False, batch_size=50, n_samples_per_label=7, seed=0)
    first = dl2.labelled_idx.copy()
    counts = np.bincount(lab[first], minlength=5)
    assert (counts <= 7).all() and counts.sum() == len(first)
    dl2.resample_labels()
    assert not np.array_equal(first, dl2.labelled_idx)
    # no labelled cell at all: labelled side is None
    dl3 = SemiSupervisedDataLoader(_FakeData(), np.full(30, 5), 5, batch_size=8)
    assert all(l is None for _, l in dl3.index_pairs()) and len(dl3) == 4
    # drop_last
    dl4 = SemiSupervisedDataLoader(_FakeData(), lab, 5, batch_size=300, drop_last=True)
    assert len(dl4) == 3 and all(f.numel() == 300 for f, _ in dl4.index_pairs())
