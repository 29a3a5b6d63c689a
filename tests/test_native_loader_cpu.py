"""Native batch loader (csrc/loader.cpp, utils/native_loader.py): sampler semantics of the reference's batchers
(examples/mnist.lua:31-40 'permutation', examples/cifar10.lua:53-54 'label-uniform'), partitioning, normalisation,
and independence of the batch stream from the number of worker threads."""
import struct

import pytest
import torch

from torch_distlearn_b200.utils.native_loader import NativeBatcher, read_cifar10_bin


def _dataset(n=203, shape=(3, 4, 4), classes=5, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (n,) + shape, generator=g, dtype=torch.uint8)
    y = torch.randint(0, classes, (n,), generator=g)
    x.view(n, -1)[:, 0] = (torch.arange(n) % 256).to(torch.uint8)      # sample id (mod 256) in the first byte
    x.view(n, -1)[:, 1] = (torch.arange(n) // 256).to(torch.uint8)
    return x, y


def _ids(xb):
    flat = xb.reshape(xb.shape[0], -1)
    return (flat[:, 0] + 256 * flat[:, 1]).round().long()


def test_permutation_visits_every_sample_once_per_epoch_and_reshuffles():
    x, y = _dataset()
    nb = NativeBatcher(x, y, batch_size=16, sampler="permutation", seed=3, threads=3, depth=4)
    try:
        assert len(nb) == 203 and nb.numBatches() == 13
        epochs = []
        for _ in range(2):
            seen, sizes = [], []
            for _ in range(nb.numBatches()):
                xb, yb, _no = nb.next()
                ids = _ids(xb)
                assert torch.equal(yb, y[ids])                           # labels travel with their samples
                seen.append(ids.clone())
                sizes.append(len(ids))
            assert sizes == [16] * 12 + [11]                             # short last batch
            seen = torch.cat(seen)
            assert sorted(seen.tolist()) == list(range(203))
            epochs.append(seen)
        assert not torch.equal(epochs[0], epochs[1])
    finally:
        nb.close()


@pytest.mark.parametrize("sampler", ["permutation", "label-uniform"])
def test_batch_stream_does_not_depend_on_the_thread_count(sampler):
    x, y = _dataset(n=500)
    streams = []
    for threads, depth in ((1, 2), (4, 6)):
        nb = NativeBatcher(x, y, batch_size=32, sampler=sampler, seed=11, threads=threads, depth=depth, hold=1)
        try:
            streams.append([tuple(t.clone() for t in nb.next()[:2]) for _ in range(40)])
        finally:
            nb.close()
    for (xa, ya), (xb, yb) in zip(*streams):
        assert torch.equal(xa, xb) and torch.equal(ya, yb)
    other = NativeBatcher(x, y, batch_size=32, sampler=sampler, seed=12)
    try:
        assert not torch.equal(other.next()[0], streams[0][0][0])
    finally:
        other.close()


def test_label_uniform_balances_classes_of_a_skewed_dataset():
    g = torch.Generator().manual_seed(1)
    y = torch.cat([torch.zeros(900, dtype=torch.long), torch.randint(1, 4, (100,), generator=g)])   # class 0 dominates, class 4 absent
    x = torch.randn(1000, 6, generator=g)
    nb = NativeBatcher(x, y, batch_size=100, sampler="label-uniform", num_classes=5, seed=5)
    try:
        counts = torch.zeros(5)
        for _ in range(60):
            _xb, yb, _ = nb.next()
            counts += torch.bincount(yb, minlength=5).float()
        assert counts[4] == 0
        freq = counts[:4] / counts.sum()
        assert float((freq - 0.25).abs().max()) < 0.03, freq
    finally:
        nb.close()


def test_partitions_are_strided_and_disjoint():
    x, y = _dataset(n=101)
    seen = []
    for part in (1, 2, 3):
        nb = NativeBatcher(x, y, batch_size=8, partition=part, partitions=3, seed=2)
        try:
            ids = torch.cat([_ids(nb.next()[0]) for _ in range(nb.numBatches())])
            assert sorted(ids.tolist()) == list(range(part - 1, 101, 3))   # uneven partition sizes: 34, 34, 33
            seen.append(ids)
        finally:
            nb.close()
    assert sorted(torch.cat(seen).tolist()) == list(range(101))
    with pytest.raises(RuntimeError):
        NativeBatcher(x[:2], y[:2], batch_size=2, partition=3, partitions=3)   # empty partition


def test_normalisation_matches_torch():
    x, y = _dataset(n=64, shape=(3, 5, 5))
    mean, std = [0.49, 0.48, 0.45], [0.25, 0.24, 0.26]
    nb = NativeBatcher(x, y, batch_size=64, scale=1 / 255, mean=mean, std=std, seed=0)
    try:
        xb, yb, _ = nb.next()
        raw = x.float().view(64, -1)
        ids = (xb.reshape(64, -1)[:, 0] * std[0] + mean[0]) * 255 + 256 * ((xb.reshape(64, -1)[:, 1] * std[0] + mean[0]) * 255).round()
        ids = ids.round().long()
        want = (x[ids].float() / 255 - torch.tensor(mean).view(1, 3, 1, 1)) / torch.tensor(std).view(1, 3, 1, 1)
        torch.testing.assert_close(xb, want, rtol=1e-5, atol=1e-5)
        assert raw.shape[1] == 75
    finally:
        nb.close()
    # float32 data, no normalisation: plain copies
    xf = torch.randn(40, 7)
    nb = NativeBatcher(xf, torch.zeros(40, dtype=torch.long), batch_size=40, seed=1)
    try:
        xb, _, _ = nb.next()
        assert sorted(map(tuple, xb.tolist())) == sorted(map(tuple, xf.tolist()))
    finally:
        nb.close()


def test_reference_style_batcher_and_slot_lifetime():
    x, y = _dataset(n=64)
    nb = NativeBatcher(x, y, batch_size=8, seed=4, depth=4, hold=2)
    try:
        getBatch, numBatches = nb.sampledBatcher(device="cpu")
        assert numBatches() == 8
        b = getBatch()
        assert b.batchSize == 8 and b.input.shape == (8, 3, 4, 4) and b.target.dtype == torch.int64
        # views of a slot stay intact while `hold` batches are outstanding
        xa, _, _ = nb.next()
        keep = xa.clone()
        nb.next()
        assert torch.equal(xa, keep)
        for _ in range(20):
            nb.next()
        assert nb.samples_done >= 23 * 8
    finally:
        nb.close()
    with pytest.raises(RuntimeError):
        nb.next()


def test_cifar_binary_reader(tmp_path):
    g = torch.Generator().manual_seed(9)
    px = torch.randint(0, 256, (7, 3072), generator=g, dtype=torch.uint8)
    lab = torch.randint(0, 10, (7,), generator=g)
    path = tmp_path / "data_batch_1.bin"
    with open(path, "wb") as f:
        for i in range(7):
            f.write(struct.pack("B", int(lab[i])))
            f.write(bytes(px[i].tolist()))
    x, y = read_cifar10_bin([path, path], max_records=5)
    assert x.shape == (10, 3, 32, 32) and torch.equal(y[:5], lab[:5]) and torch.equal(x[5:].reshape(5, -1), px[:5])
    with pytest.raises(OSError):
        read_cifar10_bin([tmp_path / "missing.bin"])
