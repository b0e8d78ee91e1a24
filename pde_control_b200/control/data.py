"""Scene directory format + batch reading (the input side of ControlTraining).

Layout written by the reference's Scene (PF/data/fluidformat.py:35,58,140,183-186,260,296):
    <datapath>/sim_%06d/<fieldname>_%06d.npz      one np.savez_compressed array per field and frame
`Dataset.load(path, range)` mirrors Dataset.load(datapath, range) as used by src/control/control_training.py:89-93;
an in-memory dict {channel: array[sims, frames, ...]} is accepted everywhere a path is.
"""
import os
import re

import numpy as np


def write_sim(datapath, sim_index, fields):
    """fields: {name: array[frames, ...]} -> sim_%06d/<name>_%06d.npz (np.savez_compressed, like Scene.write)."""
    d = os.path.join(datapath, "sim_%06d" % sim_index)
    os.makedirs(d, exist_ok=True)
    for name, frames in fields.items():
        for t, arr in enumerate(frames):
            np.savez_compressed(os.path.join(d, "%s_%06d.npz" % (name, t)), np.asarray(arr))
    return d


class Dataset(object):
    """channel name -> array [sims, frames, ...] (float32)."""

    def __init__(self, channels):
        self.channels = {k: np.asarray(v, np.float32) for k, v in channels.items()}
        sizes = {v.shape[0] for v in self.channels.values()}
        assert len(sizes) <= 1, "all channels need the same number of simulations"
        self.size = sizes.pop() if sizes else 0

    @staticmethod
    def load(path_or_dict, indices=None):
        if isinstance(path_or_dict, Dataset):
            data = path_or_dict.channels
        elif isinstance(path_or_dict, dict):
            data = path_or_dict
        else:
            return Dataset._load_scenes(os.path.expanduser(path_or_dict), indices)
        if indices is not None:
            idx = list(indices)
            data = {k: np.asarray(v)[idx] for k, v in data.items()}
        return Dataset(data)

    @staticmethod
    def _load_scenes(path, indices):
        sims = sorted(int(m.group(1)) for m in (re.match(r"sim_(\d{6})$", d) for d in os.listdir(path)) if m)
        if indices is not None:
            wanted = set(indices)
            sims = [s for s in sims if s in wanted]
        channels = {}
        for s in sims:
            d = os.path.join(path, "sim_%06d" % s)
            per_field = {}
            for f in sorted(os.listdir(d)):
                m = re.match(r"(.+)_(\d{6})\.npz$", f)
                if m:
                    with np.load(os.path.join(d, f)) as z:
                        per_field.setdefault(m.group(1), []).append(z[z.files[0]])
            for name, frames in per_field.items():
                arr = np.stack(frames)
                if arr.ndim >= 2 and arr.shape[1] == 1:          # Scene arrays carry a batch dimension of 1
                    arr = arr[:, 0]
                channels.setdefault(name, []).append(arr)
        return Dataset({k: np.stack(v) for k, v in channels.items()})

    def shard(self, rank, world_size):
        """Contiguous shard of the simulations for data-parallel training (one process per GPU)."""
        per = self.size // world_size
        return Dataset({k: v[rank * per:(rank + 1) * per] for k, v in self.channels.items()})

    def batches(self, batch_size, shuffle_seed=None):
        """Endless iterator over index arrays of `batch_size` simulations (BatchReader semantics: wraps around)."""
        order = np.arange(self.size)
        rng = np.random.RandomState(shuffle_seed) if shuffle_seed is not None else None
        pos = 0
        while True:
            if pos + batch_size > self.size:
                pos = 0
                if rng is not None:
                    rng.shuffle(order)
            yield order[pos:pos + batch_size]
            pos += batch_size
