"""Batch planning and device-side batch assembly — the caller side of the training step.

Mirrors `dataset/audio_dataset.py` of the reference:
  * `BucketizeBatchSampler` (:19-143)  — length-bucketed batches under a token (sample) budget or a fixed batch size,
  * `DistributedBatchSampler` (:146-221) — whole batches dealt out to ranks,
  * `CollateFnAudio` (:312-363)        — crop every utterance of a batch to the shortest (or zero-pad to the longest).
The samplers are host-side integer work and reproduce the reference's batches *exactly* for the same torch RNG state
(the same `randperm` / `randint` calls in the same order).  Decoding is taken out of the step loop: a `WaveformPool`
holds every utterance back to back in HBM (16-bit PCM of LibriSpeech-960 is ~110 GB — it fits a 180 GB B200), and
`DeviceCollate` assembles a batch with one gather kernel (`dph_collate_crop`) instead of per-utterance CPU slicing,
`pad_sequence` and a host-to-device copy.
"""
from typing import Iterator, List, Optional

import torch

from . import ops


class BucketizeBatchSampler:
    """Same constructor, batches, `__iter__` and `__len__` as the reference's (audio_dataset.py:19-143)."""

    def __init__(self, lengths: List[int], num_buckets: int, min_len: int = 0, max_len: Optional[int] = None,
                 max_token_count: Optional[int] = None, batch_size: Optional[int] = None, shuffle: bool = True,
                 drop_last: bool = False) -> None:
        if max_len is None:
            max_len = max(lengths)
        if not (0 <= min_len <= max_len):
            raise AssertionError("``min_len`` should be non-negative and smaller than ``max_len``")
        if max_token_count is not None and batch_size is not None:
            raise AssertionError("The ``max_token_count`` and ``batch_size`` can't be both set.")
        if max_token_count is None and batch_size is None:
            raise AssertionError("One of ``max_token_count`` or ``batch_size`` must be set.")
        if max_token_count is not None:
            assert max_len <= max_token_count, \
                "The  ``max_token_count`` must be greater than or equal to the maximum value of ``lengths``."
        lens = torch.as_tensor(list(lengths), dtype=torch.int64)
        keep = torch.nonzero((lens >= min_len) & (lens <= max_len)).flatten()
        if keep.numel() == 0:
            raise AssertionError("``lengths`` cannot be empty after filtering.")
        order = torch.sort(lens[keep], stable=True).indices  # ties keep dataset order, like sorted(key=length)
        self.indices = keep[order].tolist()
        self.lengths = lens[keep][order].tolist()
        self.max_token_count = max_token_count
        self.batch_size = batch_size
        self.shuffle = shuffle
        self.drop_last = drop_last
        self.buckets = self._get_buckets(self.lengths, num_buckets, min_len, max_len)
        self._update_iter_list()

    @staticmethod
    def _get_buckets(lengths, num_buckets, min_len, max_len):
        boundaries = torch.linspace(min_len - 1, max_len + 1, num_buckets + 1)
        ids = torch.bucketize(torch.tensor(lengths), boundaries)
        # lengths are sorted, so every bucket is one contiguous run of positions
        uniq, counts = torch.unique_consecutive(ids, return_counts=True)
        buckets, pos = {}, 0
        for k, c in zip(uniq.tolist(), counts.tolist()):
            buckets[k] = torch.arange(pos, pos + c, dtype=torch.int)
            pos += c
        return buckets

    def _update_iter_list(self) -> None:
        if self.shuffle:
            for k in self.buckets:
                self.buckets[k] = self.buckets[k][torch.randperm(self.buckets[k].size(0))]
        self.iter_list = []
        budget = self.max_token_count if self.max_token_count else self.batch_size
        total, batch = 0, []
        for k in self.buckets:
            for index in self.buckets[k].tolist():
                cost = self.lengths[index] if self.max_token_count else 1
                if total + cost <= budget:
                    batch.append(self.indices[index])
                    total += cost
                else:
                    self.iter_list.append(batch)
                    batch = [self.indices[index]]
                    total = cost
        if len(batch) > 0 and (self.max_token_count or not self.drop_last):
            self.iter_list.append(batch)

    def __iter__(self) -> Iterator[List[int]]:
        return iter(self.iter_list)

    def __len__(self):
        if self.batch_size or (self.max_token_count and not self.shuffle):
            return len(self.iter_list)


class DistributedBatchSampler:
    """Whole batches of a `BucketizeBatchSampler` dealt to ranks (audio_dataset.py:146-221), including the reference's
    padding rule: `num_replicas - len % num_replicas` batches are repeated from the front, i.e. a full extra round when
    the count already divides."""

    def __init__(self, batch_sampler, num_replicas: Optional[int] = None, rank: Optional[int] = None,
                 shuffle: bool = True, seed: int = 0, drop_last: bool = False) -> None:
        import torch.distributed as dist

        self.batch_sampler = batch_sampler
        if num_replicas is None:
            if not dist.is_available():
                raise RuntimeError("Requires distributed package to be available")
            num_replicas = dist.get_world_size()
        if rank is None:
            if not dist.is_available():
                raise RuntimeError("Requires distributed package to be available")
            rank = dist.get_rank()
        self.num_replicas, self.rank = num_replicas, rank
        self.shuffle, self.epoch, self.seed, self.drop_last = shuffle, 0, seed, drop_last
        batches = batch_sampler.iter_list
        if shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            indices = [batches[i] for i in torch.randperm(len(batches), generator=g).tolist()]
        else:
            indices = batches  # (the reference extends the sampler's own list in this case; so does this)
        if self.drop_last:
            self.total_size = len(indices) - len(indices) % self.num_replicas
        else:
            padding_size = self.num_replicas - len(indices) % self.num_replicas
            indices += indices[:padding_size]
            self.total_size = len(indices)
        self.num_samples = self.total_size // self.num_replicas
        self.subset = indices[self.rank:self.total_size:self.num_replicas]
        assert len(self.subset) == self.num_samples

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def __iter__(self):
        return iter(self.subset)

    def __len__(self):
        return self.num_samples


class WaveformPool:
    """All utterances of a subset, decoded once, back to back in one device buffer (fp32 samples or int16 PCM)."""

    def __init__(self, waveforms, device, dtype=torch.float32):
        """`waveforms`: iterable of 1-D (time,) or (1, time) tensors, in dataset index order."""
        assert dtype in (torch.float32, torch.int16)
        flat, starts, lens, pos = [], [], [], 0
        for w in waveforms:
            w = w.reshape(-1)
            if dtype == torch.int16 and w.dtype != torch.int16:
                w = torch.clamp(torch.round(w * 32768.0), -32768, 32767).to(torch.int16)
            flat.append(w.to(dtype))
            starts.append(pos)
            lens.append(w.numel())
            pos += w.numel()
        self.data = torch.cat(flat).to(device)
        self.starts, self.lengths = starts, lens
        self.device = self.data.device

    def __len__(self):
        return len(self.starts)


class DeviceCollate:
    """`CollateFnAudio` (audio_dataset.py:312-363) over a `WaveformPool`: takes the sampler's list of dataset indices,
    returns `(waveforms (B, num_frames) fp32 on the device, lengths (B,) int64)`.  Random crop offsets come from
    `torch.randint(diff, size=(1,))` per over-long utterance in batch order — the reference's exact RNG usage
    (:303-309) — so a seeded run produces bit-identical batches."""

    def __init__(self, pool: WaveformPool, pad: bool = False, rand_crop: bool = True, quantum: int = 0) -> None:
        """`quantum` > 0 (extension, crop mode only): the common length is rounded DOWN to a multiple of `quantum`
        samples, i.e. every utterance is cropped by up to quantum - 1 samples more than the reference would — so that
        batch shapes recur and `NativeDistillStep.step` can replay captured graphs.  0 = the reference's exact lengths."""
        self.pool, self.pad, self.rand_crop, self.quantum = pool, pad, rand_crop, int(quantum)

    def plan(self, batch: List[int]):
        """Host side: (start offsets into the pool, lengths to copy, num_frames)."""
        lens = [self.pool.lengths[i] for i in batch]
        num_frames = max(lens) if self.pad else min(lens)
        if self.quantum > 0 and not self.pad and num_frames >= self.quantum:
            num_frames -= num_frames % self.quantum
        starts, take = [], []
        for i, n in zip(batch, lens):
            off = 0
            if n > num_frames and self.rand_crop:
                off = int(torch.randint(n - num_frames, size=(1,)))
            starts.append(self.pool.starts[i] + off)
            take.append(min(n, num_frames))
        return starts, take, num_frames

    def __call__(self, batch: List[int], out: Optional[torch.Tensor] = None):
        starts, take, T = self.plan(batch)
        dev = self.pool.device
        B = len(batch)
        start_d = torch.tensor(starts, dtype=torch.int64).to(dev, non_blocking=True)
        take_d = torch.tensor(take, dtype=torch.int32).to(dev, non_blocking=True)
        if out is None:
            out = torch.empty(B, T, dtype=torch.float32, device=dev)
        lengths = torch.empty(B, dtype=torch.int64, device=dev)
        ops.collate_crop(self.pool.data, start_d, take_d, out, lengths, B, T)
        return out, lengths


# ----------------------------------------------------------------------------------------------------------------
# Manifest-driven loading (audio_dataset.py:220-283): tsv list -> decoded utterances -> HBM-resident pool
# ----------------------------------------------------------------------------------------------------------------
def read_manifest(tsv_dir, subset):
    """`AudioDataset._get_lists` (audio_dataset.py:239-263): first line of `<tsv_dir>/<subset>.tsv` is the root
    directory, every following line `relative/path<TAB>num_samples`.  Returns (paths, indices, lengths) as lists."""
    import os

    f_list, ind_list, len_list = [], [], []
    with open(os.path.join(str(tsv_dir), "%s.tsv" % subset)) as f:
        root = f.readline().rstrip()
        for index, line in enumerate(f):
            path, nsample = line.split("\t")
            f_list.append("%s/%s" % (root, path))
            ind_list.append(index)
            len_list.append(int(nsample))
    return f_list, ind_list, len_list


def decode_audio(path):
    """One utterance as a 1-D tensor: int16 PCM (kept as such: half the HBM and copy bytes, converted to the
    reference's float scale x / 32768 inside `dph_collate_crop`) or float32.  PCM `.wav` through the standard library,
    `.npy` / raw `.pcm` directly, anything else (`.flac`, ...) through `torchaudio.load` when its backend is installed
    (`AudioDataset._load_audio`, audio_dataset.py:265-277).  Multi-channel files are refused like the reference's
    collate would mis-handle them (it reads channel 0 only)."""
    import numpy as np

    ext = path.rsplit(".", 1)[-1].lower()
    if ext == "wav":
        import wave

        with wave.open(path, "rb") as f:
            nch, width, n = f.getnchannels(), f.getsampwidth(), f.getnframes()
            raw = f.readframes(n)
        if width == 2:
            x = torch.from_numpy(np.frombuffer(raw, dtype="<i2").copy())
        elif width == 4:
            x = torch.from_numpy((np.frombuffer(raw, dtype="<i4") / 2147483648.0).astype(np.float32))
        else:
            raise ValueError("%s: unsupported PCM sample width %d" % (path, width))
        if nch != 1:
            x = x.view(-1, nch)[:, 0].contiguous()
        return x
    if ext == "npy":
        x = torch.from_numpy(np.load(path))
        return x.reshape(-1) if x.dtype in (torch.int16, torch.float32) else x.reshape(-1).float()
    if ext in ("pcm", "raw"):
        return torch.from_numpy(np.fromfile(path, dtype="<i2"))
    try:
        import torchaudio

        w, _ = torchaudio.load(path)
    except ImportError as e:
        raise RuntimeError("decoding %s needs a torchaudio backend (torchcodec) that is not installed; convert the "
                           "subset to PCM .wav / .npy or pass decode=<callable>" % path) from e
    return w[0].contiguous()


def load_pool(paths, lengths, device, keep=None, dtype=torch.int16, chunk_samples=1 << 24, workers=8,
              decode=decode_audio):
    """Decode the utterances `keep` (dataset indices; default all) ONCE and lay them back to back in one device
    buffer.  Decoding runs on a thread pool; samples go through two pinned staging buffers and are copied on a side
    stream, so decode of chunk i + 1 overlaps the host-to-device copy of chunk i.  Utterances the batch sampler
    filtered out (min_len / max_len) are never read.  Returns a `WaveformPool` indexed by dataset index."""
    from concurrent.futures import ThreadPoolExecutor

    device = torch.device(device)
    order = sorted(range(len(paths)) if keep is None else set(int(i) for i in keep))
    total = sum(int(lengths[i]) for i in order)
    pool = WaveformPool.__new__(WaveformPool)
    pool.data = torch.empty(max(total, 1), dtype=dtype, device=device)
    pool.starts = [-1] * len(paths)
    pool.lengths = [int(n) for n in lengths]
    pool.device = pool.data.device
    cuda = device.type == "cuda"
    chunk = max(int(chunk_samples), max([pool.lengths[i] for i in order] or [1]))
    stage = [torch.empty(chunk, dtype=dtype).pin_memory() if cuda else torch.empty(chunk, dtype=dtype) for _ in range(2)]
    free = [torch.cuda.Event() if cuda else None for _ in range(2)]
    used = [False, False]
    copy_stream = torch.cuda.Stream(device) if cuda else None
    state = {"buf": 0, "fill": 0, "pos": 0}

    def flush():
        b, n = state["buf"], state["fill"]
        if n:
            dst = pool.data[state["pos"]:state["pos"] + n]
            if cuda:
                with torch.cuda.stream(copy_stream):
                    dst.copy_(stage[b][:n], non_blocking=True)
                    free[b].record()
                used[b] = True
            else:
                dst.copy_(stage[b][:n])
            state["pos"] += n
        state["buf"], state["fill"] = 1 - b, 0
        if cuda and used[state["buf"]]:
            free[state["buf"]].synchronize()   # the copy that last read this staging buffer has finished

    def convert(x):
        x = x.reshape(-1)
        if dtype == torch.int16 and x.dtype != torch.int16:
            x = torch.clamp(torch.round(x.float() * 32768.0), -32768, 32767).to(torch.int16)
        elif dtype == torch.float32 and x.dtype == torch.int16:
            x = x.float() / 32768.0
        return x.to(dtype)

    with ThreadPoolExecutor(max_workers=max(1, workers)) as ex:
        for i, x in zip(order, ex.map(lambda j: convert(decode(paths[j])), order)):
            n = x.numel()
            if n != pool.lengths[i]:   # audio_dataset.py:276
                raise AssertionError("%s has %d samples, the manifest says %d" % (paths[i], n, pool.lengths[i]))
            if state["fill"] + n > chunk:
                flush()
            stage[state["buf"]][state["fill"]:state["fill"] + n].copy_(x)
            pool.starts[i] = state["pos"] + state["fill"]
            state["fill"] += n
    flush()
    if cuda:
        torch.cuda.current_stream(device).wait_stream(copy_stream)
    return pool


class DeviceLoader:
    """What `DataLoader(dataset, batch_sampler=..., collate_fn=CollateFnAudio(...))` is to the reference
    (lightning.py:306-342): iterating yields `(waveforms (B, T) fp32, lengths (B,) int64)` per batch of the sampler —
    already on the device, assembled by one gather kernel from the HBM-resident pool."""

    def __init__(self, pool: WaveformPool, batch_sampler, pad: bool = False, rand_crop: bool = True, quantum: int = 0):
        self.pool, self.batch_sampler = pool, batch_sampler
        self.collate = DeviceCollate(pool, pad=pad, rand_crop=rand_crop, quantum=quantum)

    def __len__(self):
        return len(self.batch_sampler)

    def __iter__(self):
        for batch in self.batch_sampler:
            yield self.collate(batch)
