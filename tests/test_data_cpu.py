"""Data pipeline and driver host logic on CPU: the circles-squares generator against a
loop restatement of bin/make-circles-squares.py, the within-class matchings of bin/make-mnist-svhn-idx.py /
bin/make_trimodal.py, the paired-batch sampler's rank sharding, and the driver's config / scheduler helpers."""
import json
import os

import numpy as np
import torch

from mmvae_b200 import data as D


def _script_loop(dataset_size, n_repeat, size_image, rs):
    """bin/make-circles-squares.py:18-45 as written (scalar draws, python lists)."""
    thick = 0.25
    x = np.linspace(-1, 1, size_image)
    circle = lambda X, Y, r: (X**2 + Y**2 <= (r + thick / 2)**2) * (X**2 + Y**2 >= (r - thick / 2)**2)      # noqa: E731
    square_line = lambda X, Y, r: (np.abs(X) + np.abs(Y) <= (r + thick / 2)) * (np.abs(X) + np.abs(Y) >= r - thick / 2)  # noqa: E731
    squares, circles, labels, r_s, r_c = [], [], [], [], []
    for _ in range(dataset_size):
        for _ in range(n_repeat):
            X, Y = np.meshgrid(x, x)
            r_c.extend([rs.uniform(0.3, 0.9) for _ in range(2)])
            r_s.extend([rs.uniform(0.3, 0.9) for _ in range(2)])
            squares.extend([np.abs(X) + np.abs(Y) <= r_s[-2], square_line(X, Y, r_s[-1])])
            circles.extend([X**2 + Y**2 <= r_c[-2]**2, circle(X, Y, r_c[-1])])
            labels.extend([1, 0])
    return np.array(squares), np.array(circles), np.array(labels), np.array(r_s), np.array(r_c)


def test_circles_squares_matches_script_loop():
    sq, ci, lab, rsq, rci = _script_loop(7, 3, 32, np.random.RandomState(5))
    out = D.make_circles_squares(dataset_size=7, n_repeat=3, size_image=32, seed=5, test_size=0.3)
    n = len(lab)
    assert out["squares_train"].shape[0] + out["squares_test"].shape[0] == n
    assert out["squares_test"].shape[0] == int(np.ceil(0.3 * n))
    assert out["squares_train"].shape[1:] == (1, 32, 32) and out["squares_train"].dtype == torch.float32
    # undo the split through the radii (unique with probability 1) and compare every image bit for bit
    for split in ("train", "test"):
        for j in range(out[f"labels_{split}"].shape[0]):
            i = int(np.argmin(np.abs(rci - out[f"r_circles_{split}"][j].item())))
            assert rsq[i] == out[f"r_squares_{split}"][j].item()
            assert lab[i] == out[f"labels_{split}"][j].item()
            assert np.array_equal(sq[i].astype(np.float32), out[f"squares_{split}"][j, 0].numpy())
            assert np.array_equal(ci[i].astype(np.float32), out[f"circles_{split}"][j, 0].numpy())
    assert set(out["squares_train"].unique().tolist()) <= {0.0, 1.0}


def test_save_circles_squares_file_names(tmp_path):
    D.save_circles_squares(str(tmp_path), dataset_size=3, n_repeat=2, seed=1)
    want = {f"{a}_{s}.pt" for a in ("squares", "circles", "labels", "r_squares", "r_circles") for s in ("train", "test")}
    assert want == set(os.listdir(tmp_path))                           # bin/make-circles-squares.py:68-77


def test_rand_match_on_idx_pairs_within_class():
    g = torch.Generator().manual_seed(0)
    labels = [torch.randint(0, 10, (500,), generator=g), torch.randint(0, 10, (700,), generator=g),
              torch.randint(0, 10, (400,), generator=g)]
    for M in (2, 3):
        idx = D.rand_match_on_idx(labels[:M], max_d=30, dm=4, generator=g)
        assert len({len(i) for i in idx}) == 1
        cls = [labels[m][idx[m]] for m in range(M)]
        for m in range(1, M):
            assert torch.equal(cls[0], cls[m])                         # every pair shares its class
        # per class: dm blocks of n = min(counts, max_d), each block a permutation of the same first-n datapoints
        for c in range(10):
            n = min(min(int((labels[m] == c).sum()) for m in range(M)), 30)
            sel = cls[0] == c
            assert int(sel.sum()) == 4 * n
            for m in range(M):
                blocks = idx[m][sel].reshape(4, n)
                first = labels[m].sort(stable=True)[1][labels[m].sort(stable=True)[0] == c][:n]
                for b in blocks:
                    assert torch.equal(b.sort()[0], first.sort()[0])
                assert not all(torch.equal(blocks[0], b) for b in blocks[1:])


def test_make_pair_indices_files_and_loader_splits(tmp_path):
    train_b, test_b = D.synthetic_bases(n_train=600, n_test=200, seed=3)
    tr, te = D.make_pair_indices(str(tmp_path), [l for _, l in train_b], [l for _, l in test_b], max_d=100, dm=2)
    assert {"train-ms-mnist-idx.pt", "train-ms-svhn-idx.pt", "test-ms-mnist-idx.pt", "test-ms-svhn-idx.pt"} == \
        set(os.listdir(tmp_path))                                      # bin/make-mnist-svhn-idx.py:40-47
    (tr2, te2) = D.load_pair_indices(str(tmp_path), 2)
    assert all(torch.equal(a, b) for a, b in zip(tr, tr2)) and all(torch.equal(a, b) for a, b in zip(te, te2))
    train, test, val = D.mnist_svhn_like_loaders(train_b, test_b, tr, te, batch_size=64, device="cpu")
    n = len(tr[0])
    assert val.dataset_len == min(10000, n // 10) and train.dataset_len == n - val.dataset_len   # dataloaders.py:276-280
    assert test.dataset_len == len(te[0])
    x = next(iter(train))
    assert x[0].shape == (64, 1, 28, 28) and x[1].shape == (64, 3, 32, 32) and x[0].dtype == torch.float32
    assert 0.0 <= float(x[1].min()) and float(x[1].max()) <= 1.0       # ToTensor range
    # pairs keep their class: recover the labels through the image -> index mapping of the sampler
    sel = torch.arange(10)
    li = [train_b[m][1][train.idx[m][sel]] for m in range(2)]
    assert torch.equal(li[0], li[1])


def test_paired_batches_shard_over_ranks():
    bases = [torch.arange(50, dtype=torch.float32).reshape(50, 1, 1, 1), torch.arange(50, dtype=torch.float32).reshape(50, 1, 1, 1) + 100]
    idx = [torch.randperm(50, generator=torch.Generator().manual_seed(1)) for _ in range(2)]
    seen = []
    for rank in range(3):
        pb = D.PairedBatches(bases, idx, batch_size=4, shuffle=True, device="cpu", seed=9, rank=rank, world=3)
        assert len(pb) == 4                                            # 48 usable pairs / 3 ranks / batch 4
        got = [x for x in pb]
        assert len(got) == 4 and all(x[0].shape[0] == 4 for x in got)
        seen.append(torch.cat([x[0].reshape(-1) for x in got]))
    allv = torch.cat(seen)
    assert allv.numel() == 48 and allv.unique().numel() == 48         # disjoint shards of one global permutation


def test_driver_helpers(tmp_path):
    from mmvae_b200 import drivers
    cfg = {"model": "mnist_svhn", "obj": "dreg", "looser": True, "K": 30, "latent_dim": 20, "dist": "laplace",
           "batch_size": 128, "learning_rate": 1e-3, "epochs": 2}
    p = tmp_path / "c.json"
    p.write_text(json.dumps(cfg))
    args = drivers.load_config(str(p))
    assert drivers.objective_name(args) == "m_dreg_looser" and args.seed == 1              # main.py:134-137
    args.obj = "elbo"
    assert drivers.objective_name(args) == "m_elbo"
    m, s = drivers.get_mean_std({"a": [1.0, 2.0, 3.0]})
    assert float(m["a"]) == 2.0 and abs(float(s["std_a"]) - 1.0) < 1e-6                    # utils.py:253-255

    class T:
        lr = 1e-3

        def set_lr(self, lr):
            self.lr = lr
    t = T()
    ours = drivers.ReduceLROnPlateau(t)
    w = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.Adam([w], lr=1e-3)
    ref = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, "min")
    rs = np.random.RandomState(0)
    for i in range(80):
        v = 10.0 - min(i, 20) * 0.1 + 0.0001 * rs.rand()
        ours.step(v)
        ref.step(v)
        assert abs(t.lr - opt.param_groups[0]["lr"]) < 1e-15, i
