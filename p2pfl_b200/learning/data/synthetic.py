"""Synthetic federated data modules.

There is no network in the target environment, so every entry point runs on
synthetic data of the named shape (BASELINE.json).  The partitioning mirrors
the reference's ``MnistFederatedDM`` (p2pfl/learning/pytorch/mnist_examples/
mnistfederated_dm.py:105-126): the ``sub_id``-th contiguous slice of a
``n_train`` / ``n_test`` sample pool, 90/10 train/val split, non-IID = sort by
label before slicing.

Samples are *learnable*: ``x = prototype[y] + noise`` with class prototypes
fixed by ``proto_seed`` (identical on every node) so that federated accuracy
is a meaningful signal in tests.

Besides the Lightning-style ``train/val/test_dataloader()`` the module exposes
``tensors(split)`` (whole split as two tensors) which the device-resident
learners use, and ``pinned_tensors(split)`` for the end-to-end H2D path.
"""

from __future__ import annotations

from typing import Dict, Tuple

import torch
from torch.utils.data import DataLoader, TensorDataset


class SyntheticFederatedDM:
    def __init__(
        self,
        sub_id: int = 0,
        number_sub: int = 1,
        batch_size: int = 32,
        num_workers: int = 0,
        val_percent: float = 0.1,
        iid: bool = True,
        shape: Tuple[int, ...] = (1, 28, 28),
        num_classes: int = 10,
        n_train: int = 60000,
        n_test: int = 10000,
        noise: float = 1.0,
        seed: int = 1234,
        proto_seed: int = 7,
    ) -> None:
        if not 0 <= sub_id < number_sub:
            raise ValueError(f"sub_id {sub_id} out of range for {number_sub} partitions")
        self.sub_id, self.number_sub = sub_id, number_sub
        self.batch_size, self.num_workers = batch_size, num_workers
        self.val_percent, self.iid = val_percent, iid
        self.shape, self.num_classes = tuple(shape), num_classes
        self.n_train_total, self.n_test_total = n_train, n_test
        self.noise, self.seed, self.proto_seed = noise, seed, proto_seed
        self._cache: Dict[str, Tuple[torch.Tensor, torch.Tensor]] = {}
        self._pinned: Dict[str, Tuple[torch.Tensor, torch.Tensor]] = {}

    # -------------------------------------------------------------- generate
    def _prototypes(self) -> torch.Tensor:
        g = torch.Generator().manual_seed(self.proto_seed)
        return torch.randn((self.num_classes, *self.shape), generator=g)

    def _make(self, n: int, salt: int) -> Tuple[torch.Tensor, torch.Tensor]:
        g = torch.Generator().manual_seed(self.seed * 1000003 + self.sub_id * 101 + salt)
        if self.iid:
            y = torch.randint(0, self.num_classes, (n,), generator=g)
        else:
            # label-sorted pool => this slice only sees a narrow label band
            lo = (self.sub_id * self.num_classes) // self.number_sub
            hi = max(lo + 1, ((self.sub_id + 1) * self.num_classes + self.number_sub - 1) // self.number_sub)
            y = torch.randint(lo, min(hi, self.num_classes), (n,), generator=g)
        x = self._prototypes()[y] + self.noise * torch.randn((n, *self.shape), generator=g)
        return x.contiguous(), y.contiguous()

    def _build(self) -> None:
        if self._cache:
            return
        n_slice = self.n_train_total // self.number_sub
        n_val = int(round(n_slice * self.val_percent))
        n_tr = n_slice - n_val
        n_te = self.n_test_total // self.number_sub
        self._cache["train"] = self._make(n_tr, 1)
        self._cache["val"] = self._make(n_val, 2)
        self._cache["test"] = self._make(n_te, 3)

    # ---------------------------------------------------------------- access
    def tensors(self, split: str) -> Tuple[torch.Tensor, torch.Tensor]:
        self._build()
        return self._cache[split]

    def pinned_tensors(self, split: str) -> Tuple[torch.Tensor, torch.Tensor]:
        """Page-locked copies (only meaningful when CUDA is present)."""
        if split not in self._pinned:
            x, y = self.tensors(split)
            if torch.cuda.is_available():
                x, y = x.pin_memory(), y.pin_memory()
            self._pinned[split] = (x, y)
        return self._pinned[split]

    def num_samples(self) -> Tuple[int, int]:
        self._build()
        return self._cache["train"][0].shape[0], self._cache["test"][0].shape[0]

    def _loader(self, split: str, shuffle: bool) -> DataLoader:
        x, y = self.tensors(split)
        return DataLoader(TensorDataset(x, y), batch_size=self.batch_size, shuffle=shuffle, num_workers=self.num_workers)

    def train_dataloader(self) -> DataLoader:
        return self._loader("train", True)

    def val_dataloader(self) -> DataLoader:
        return self._loader("val", False)

    def test_dataloader(self) -> DataLoader:
        return self._loader("test", False)


class MnistFederatedDM(SyntheticFederatedDM):
    """MNIST-shaped partition (1x28x28, 10 classes, 60 000 / 10 000 pool), same
    constructor as the reference's data module (mnistfederated_dm.py:56-72).
    The reference downloads MNIST; offline, the pool is synthetic."""

    def __init__(
        self,
        sub_id: int = 0,
        number_sub: int = 1,
        batch_size: int = 32,
        num_workers: int = 0,
        val_percent: float = 0.1,
        iid: bool = True,
        **kw,
    ) -> None:
        super().__init__(sub_id, number_sub, batch_size, num_workers, val_percent, iid, shape=(1, 28, 28), num_classes=10, **kw)


class Cifar10FederatedDM(SyntheticFederatedDM):
    """CIFAR-10-shaped partition (3x32x32, 10 classes, 50 000 / 10 000 pool)."""

    def __init__(self, sub_id: int = 0, number_sub: int = 1, batch_size: int = 32, **kw) -> None:
        kw.setdefault("n_train", 50000)
        kw.setdefault("n_test", 10000)
        super().__init__(sub_id, number_sub, batch_size, shape=(3, 32, 32), num_classes=10, **kw)


class TokenFederatedDM(SyntheticFederatedDM):
    """Synthetic token streams for the GPT-2 config: ``x`` = tokens[:-1],
    ``y`` = tokens[1:], generated by a fixed random bigram table so the
    next-token task is learnable."""

    def __init__(
        self,
        sub_id: int = 0,
        number_sub: int = 1,
        batch_size: int = 8,
        seq_len: int = 1024,
        vocab: int = 50257,
        n_train: int = 512,
        n_test: int = 64,
        **kw,
    ) -> None:
        super().__init__(sub_id, number_sub, batch_size, shape=(seq_len,), num_classes=vocab, n_train=n_train, n_test=n_test, **kw)
        self.seq_len, self.vocab = seq_len, vocab

    def _make(self, n: int, salt: int) -> Tuple[torch.Tensor, torch.Tensor]:
        g = torch.Generator().manual_seed(self.seed * 1000003 + self.sub_id * 101 + salt)
        gp = torch.Generator().manual_seed(self.proto_seed)
        succ = torch.randint(0, self.vocab, (self.vocab, 4), generator=gp)  # 4 likely successors per token
        toks = torch.empty((n, self.seq_len + 1), dtype=torch.long)
        toks[:, 0] = torch.randint(0, self.vocab, (n,), generator=g)
        pick = torch.randint(0, 4, (n, self.seq_len), generator=g)
        for t in range(self.seq_len):
            toks[:, t + 1] = succ[toks[:, t], pick[:, t]]
        return toks[:, :-1].contiguous(), toks[:, 1:].contiguous()
