"""Offline data pipeline on either side of the hot path.

* `make_circles_squares` — the paired toy dataset of bin/make-circles-squares.py:9-77 (disc <-> filled diamond,
  ring <-> diamond outline, radii ~ U(0.3, 0.9) on a 32x32 grid over [-1,1]^2), same tensors and file names.
* `rand_match_on_idx` / `make_pair_indices` — the within-class random matchings that pair MNIST with SVHN (and
  FashionMNIST): bin/make-mnist-svhn-idx.py:4-47, bin/make_trimodal.py:6-60 (`dm` permutations per class, at most
  `max_d` datapoints per class), saved under the reference's file names (`train-ms-mnist-idx.pt`, ...).
* `PairedBatches` — what `MNIST_SVHN_DL.getDataLoaders` (dataloaders.py:239-290) assembles from those files
  (ResampleDataset(base_m, i -> idx_m[i]), the seed-42 shuffle, the seed-42 train/val split), but resident on the
  GPU: the base images live in HBM once (uint8 or fp32), an epoch is a permutation of pair indices, and a batch is M
  gathers — no DataLoader workers, no per-step host work.  `synthetic_bases` makes label-carrying stand-ins of the
  MNIST / SVHN / FashionMNIST shapes for boxes without the datasets (no network on the GPU box).

Nothing here is on the timed hot path; torch indexing ops are used as plumbing.
"""
import os

import numpy as np
import torch

# ------------------------------------------------------------------------------------------- circles / squares
MIN_RAYON, MAX_RAYON, THICKNESS = 0.3, 0.9, 0.25          # bin/make-circles-squares.py:11-12


def make_circles_squares(dataset_size=10000, n_repeat=10, size_image=32, seed=1, test_size=0.3):
    """bin/make-circles-squares.py:9-65 vectorised.  Per (i, repeat): one (disc, filled diamond) pair with label 1 and
    one (ring, diamond outline) pair with label 0, each with its own radius ~ U(0.3, 0.9) drawn in the script's order
    (r_circle_full, r_circle_ring, r_square_full, r_square_line).  Returns a dict of the ten tensors the script saves
    ({squares,circles,labels,r_squares,r_circles}_{train,test}); images are float {0,1} of shape (N,1,S,S).
    The split is a seeded permutation (the script's sklearn `train_test_split` draws differently: same sizes, not the
    same membership)."""
    rs = np.random.RandomState(seed)
    n = dataset_size * n_repeat
    r = rs.uniform(MIN_RAYON, MAX_RAYON, size=(n, 4))               # order of the script's two list comprehensions
    r_circ = np.stack([r[:, 0], r[:, 1]], 1).reshape(-1)            # (full disc, ring) interleaved
    r_sq = np.stack([r[:, 2], r[:, 3]], 1).reshape(-1)              # (filled diamond, outline) interleaved
    x = np.linspace(-1, 1, size_image)
    X, Y = np.meshgrid(x, x)
    d2 = (X ** 2 + Y ** 2)[None]                                    # (1,S,S)
    l1 = (np.abs(X) + np.abs(Y))[None]
    full = np.arange(2 * n) % 2 == 0
    rc, rq = r_circ[:, None, None], r_sq[:, None, None]
    h = THICKNESS / 2
    circles = np.where(full[:, None, None], d2 <= rc ** 2, (d2 <= (rc + h) ** 2) & (d2 >= (rc - h) ** 2))
    squares = np.where(full[:, None, None], l1 <= rq, (l1 <= rq + h) & (l1 >= rq - h))
    labels = full.astype(np.int64)
    perm = np.random.RandomState(seed + 1).permutation(2 * n)
    n_test = int(np.ceil(test_size * 2 * n))
    idx = {"test": perm[:n_test], "train": perm[n_test:]}
    out = {}
    for split, ii in idx.items():
        out[f"squares_{split}"] = torch.from_numpy(squares[ii].astype(np.float32)).unsqueeze(1)
        out[f"circles_{split}"] = torch.from_numpy(circles[ii].astype(np.float32)).unsqueeze(1)
        out[f"labels_{split}"] = torch.from_numpy(labels[ii])
        out[f"r_squares_{split}"] = torch.from_numpy(r_sq[ii])
        out[f"r_circles_{split}"] = torch.from_numpy(r_circ[ii])
    return out


def save_circles_squares(output_path, **kw):
    """Writes `<output_path>/{squares,circles,labels,r_squares,r_circles}_{train,test}.pt` (script lines 68-77)."""
    os.makedirs(output_path, exist_ok=True)
    data = make_circles_squares(**kw)
    for k, v in data.items():
        torch.save(v, os.path.join(output_path, k + ".pt"))
    return data


# ------------------------------------------------------------------------------------- within-class matchings
def rand_match_on_idx(labels, max_d=10000, dm=10, generator=None):
    """bin/make-mnist-svhn-idx.py:4-20 / bin/make_trimodal.py:6-20 for M >= 2 label vectors.

    labels: list of M 1-D integer tensors (the class of every datapoint of each base dataset).  For every class, the
    first n = min(count_m, max_d) datapoints of each modality (in label-sorted order, as the script's `sort()`) are
    matched `dm` times through independent random permutations.  -> list of M index tensors of equal length."""
    sorted_l, sorted_i = zip(*[torch.as_tensor(l).sort(stable=True) for l in labels])
    out = [[] for _ in labels]
    for c in sorted_l[0].unique():
        per = [si[sl == c] for sl, si in zip(sorted_l, sorted_i)]
        n = min(min(p.size(0) for p in per), max_d)
        for _ in range(dm):
            for m, p in enumerate(per):
                out[m].append(p[:n][torch.randperm(n, generator=generator)])
    return [torch.cat(o) for o in out]


MS_NAMES = ("mnist", "svhn")
MSF_NAMES = ("mnist", "svhn", "fashion")


def make_pair_indices(data_path, train_labels, test_labels, max_d=10000, dm=5, seed=1):
    """Writes `{train,test}-ms-<name>-idx.pt` (M=2) or `{train,test}-msf-<name>-idx.pt` (M=3) under data_path; the test
    matching is shuffled once more (make-mnist-svhn-idx.py:43-47)."""
    M = len(train_labels)
    tag, names = ("ms", MS_NAMES) if M == 2 else ("msf", MSF_NAMES)
    g = torch.Generator().manual_seed(seed)
    os.makedirs(data_path, exist_ok=True)
    tr = rand_match_on_idx(train_labels, max_d, dm, g)
    te = rand_match_on_idx(test_labels, max_d, dm, g)
    shuffle = torch.randperm(len(te[0]), generator=g)
    te = [t[shuffle] for t in te]
    for name, a, b in zip(names, tr, te):
        torch.save(a, os.path.join(data_path, f"train-{tag}-{name}-idx.pt"))
        torch.save(b, os.path.join(data_path, f"test-{tag}-{name}-idx.pt"))
    return tr, te


def load_pair_indices(data_path, M=2):
    tag, names = ("ms", MS_NAMES) if M == 2 else ("msf", MSF_NAMES)
    paths = [[os.path.join(data_path, f"{split}-{tag}-{n}-idx.pt") for n in names] for split in ("train", "test")]
    if not all(os.path.exists(p) for ps in paths for p in ps):
        raise RuntimeError("Generate transformed indices with mmvae_b200.data.make_pair_indices")   # dataloaders.py:249
    return [[torch.load(p) for p in ps] for ps in paths]


def synthetic_bases(shapes=((1, 28, 28), (3, 32, 32)), n_train=6000, n_test=1000, n_classes=10, seed=1):
    """Label-carrying stand-ins for the base datasets: per modality uint8 images (N,C,H,W) whose content depends on the
    class (a class-specific low-frequency pattern + noise), and the labels.  -> (train, test), each a list of
    (images_uint8, labels) per modality."""
    g = torch.Generator().manual_seed(seed)

    def one(n, shape, salt):
        labels = torch.randint(0, n_classes, (n,), generator=g)
        C, H, W = shape
        yy, xx = torch.meshgrid(torch.linspace(0, 1, H), torch.linspace(0, 1, W), indexing="ij")
        proto = torch.stack([0.5 + 0.5 * torch.sin(6.28 * ((c + 1 + salt) * xx / n_classes + (c + 2) * yy / 3))
                             for c in range(n_classes)])                       # (classes,H,W)
        img = proto[labels].unsqueeze(1).expand(n, C, H, W) * 0.8 + 0.2 * torch.rand(n, C, H, W, generator=g)
        return (img.clamp(0, 1) * 255).to(torch.uint8), labels

    train = [one(n_train, s, i) for i, s in enumerate(shapes)]
    test = [one(n_test, s, i) for i, s in enumerate(shapes)]
    return train, test


# ------------------------------------------------------------------------------------ GPU-resident paired batches
class PairedBatches:
    """Paired multimodal batches served from HBM.

    bases: list over modalities of image tensors (N_m,C,H,W), uint8 (ToTensor scaling 1/255 applied at gather time) or
    float in [0,1].  idx: list over modalities of equally long index tensors: pair i = (bases[m][idx[m][i]])_m, the
    ResampleDataset mapping of dataloaders.py:266-273.  Iterating yields `[x_0, ..., x_{M-1}]` fp32 device batches in
    the reference's `unpack_data` form (utils.py:108-127), ready for `objective(model, x, K, ...)`.
    """

    def __init__(self, bases, idx, batch_size, shuffle=True, device="cuda", drop_last=False, seed=0, rank=0, world=1):
        assert len(bases) == len(idx) and len({len(i) for i in idx}) == 1
        self.device = torch.device(device)
        self.bases = [b.to(self.device) for b in bases]
        self.idx = [i.to(self.device, dtype=torch.long) for i in idx]
        self.batch_size, self.shuffle, self.drop_last = batch_size, shuffle, drop_last
        self.rank, self.world = rank, world
        self.gen = torch.Generator().manual_seed(seed)           # same seed on every rank: one global permutation
        self.n_pairs = len(idx[0])
        self.n_used = (self.n_pairs // world) * world            # equal work per rank (collectives stay in step)

    @property
    def dataset_len(self):
        return self.n_pairs

    def __len__(self):
        per_rank = self.n_used // self.world
        return per_rank // self.batch_size if self.drop_last else -(-per_rank // self.batch_size)

    def _gather(self, sel):
        out = []
        for b, i in zip(self.bases, self.idx):
            x = b.index_select(0, i.index_select(0, sel))
            out.append(x.float().div_(255.0) if x.dtype == torch.uint8 else x.float())
        return out

    def __iter__(self):
        order = torch.randperm(self.n_pairs, generator=self.gen) if self.shuffle else torch.arange(self.n_pairs)
        order = order[:self.n_used][self.rank::self.world].to(self.device)   # batch sharding over ranks
        nb = len(self)
        for k in range(nb):
            yield self._gather(order[k * self.batch_size:(k + 1) * self.batch_size])


def mnist_svhn_like_loaders(train_bases, test_bases, train_idx, test_idx, batch_size, shuffle=True, device="cuda",
                            len_train=None, rank=0, world=1):
    """train / test / val `PairedBatches` with the splits of MNIST_SVHN_DL.getDataLoaders (dataloaders.py:258-290):
    the train matching is permuted with RandomState(42) and truncated to len_train; min(10000, len//10) pairs are held
    out for validation by a seed-42 `random_split`."""
    rd = torch.from_numpy(np.random.RandomState(seed=42).permutation(len(train_idx[0])))
    tr_idx = [i[rd] for i in train_idx]
    if len_train is not None:
        tr_idx = [i[:len_train] for i in tr_idx]
    n = len(tr_idx[0])
    len_val = min(10000, n // 10)
    split = torch.randperm(n, generator=torch.Generator().manual_seed(42))
    tr_sel, va_sel = split[:n - len_val], split[n - len_val:]
    imgs_tr = [b for b, _ in train_bases]
    imgs_te = [b for b, _ in test_bases]
    train = PairedBatches(imgs_tr, [i[tr_sel] for i in tr_idx], batch_size, shuffle, device, rank=rank, world=world)
    val = PairedBatches(imgs_tr, [i[va_sel] for i in tr_idx], batch_size, False, device, rank=rank, world=world)
    test = PairedBatches(imgs_te, test_idx, batch_size, False, device, rank=rank, world=world)
    return train, test, val
