"""Host tabulation of the random subsets sklearn's RANSACRegressor draws (reference:
point_cloud_processing/vel_estimator.py:92-99 constructs it with random_state=42).

``random_state=42`` is turned into a fresh ``np.random.RandomState(42)`` at the top of every ``fit``; each trial then
calls ``sample_without_replacement(n_points, min_samples, random_state=rng)`` once.  The indices therefore depend only on
(n_points, trial) and can be tabulated with numpy's RandomState itself - the very generator the reference uses, whose
stream numpy freezes across releases.  Which of its methods is called depends on the scikit-learn release
(utils/_random.pyx):

* ``"permutation"`` (scikit-learn as installed next to the reference here, 1.9.0): ``rng.permutation(n)[:k]`` when
  0.01 < k/n < 0.99, tracking selection when k/n <= 0.01, reservoir sampling when k/n >= 0.99;
* ``"reservoir"``: reservoir sampling for every k/n >= 0.01 (the rule of older releases; offered for users pinned to
  them, not checked against one here).
"""
from __future__ import annotations

import numpy as np


def subsets_for(n_points: int, min_samples: int = 10, max_trials: int = 20, seed: int = 42,
                policy: str = "permutation") -> np.ndarray:
    """-> int32 [max_trials, min_samples] point indices of every trial for a cloud of n_points points."""
    if n_points < min_samples:
        raise ValueError("`min_samples` may not be larger than number of samples: n_samples = %d." % n_points)
    rng = np.random.RandomState(seed)
    out = np.empty((max_trials, min_samples), dtype=np.int32)
    ratio = min_samples / n_points
    for t in range(max_trials):
        if policy == "permutation" and 0.01 < ratio < 0.99:
            out[t] = rng.permutation(n_points)[:min_samples]
        elif ratio < (0.2 if policy == "permutation" else 0.01):
            chosen = {}
            for i in range(min_samples):
                j = rng.randint(n_points)
                while j in chosen:
                    j = rng.randint(n_points)
                chosen[j] = i
            out[t] = list(chosen)
        else:
            res = np.arange(min_samples)
            for i in range(min_samples, n_points):
                j = rng.randint(0, i + 1)
                if j < min_samples:
                    res[j] = i
            out[t] = res
    return out


class SubsetTable:
    """Blocks of ``subsets_for`` keyed by point count, grown on demand, mirrored on the device as
    (blocks int32 [n_blocks, max_trials, min_samples], index int32 [max_n + 1])."""

    def __init__(self, min_samples: int = 10, max_trials: int = 20, seed: int = 42, policy: str = "permutation"):
        self.min_samples, self.max_trials, self.seed, self.policy = min_samples, max_trials, seed, policy
        self.blocks = {}
        self._device = None

    def ensure(self, counts) -> bool:
        """Tabulate every point count in ``counts`` (>= min_samples); True if anything was added."""
        added = False
        for n in np.unique(np.asarray(counts, dtype=np.int64)):
            n = int(n)
            if n >= self.min_samples and n not in self.blocks:
                self.blocks[n] = subsets_for(n, self.min_samples, self.max_trials, self.seed, self.policy)
                added = True
        if added:
            self._device = None
        return added

    def device_arrays(self, device):
        import torch
        if self._device is None or self._device[2] != str(device):
            keys = sorted(self.blocks)
            max_n = keys[-1] if keys else self.min_samples
            index = np.full(max_n + 1, -1, dtype=np.int32)
            data = np.zeros((max(1, len(keys)), self.max_trials, self.min_samples), dtype=np.int32)
            for b, n in enumerate(keys):
                index[n] = b
                data[b] = self.blocks[n]
            self._device = (torch.from_numpy(data).to(device), torch.from_numpy(index).to(device), str(device), max_n)
        return self._device[0], self._device[1], self._device[3]
