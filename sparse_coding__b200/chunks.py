"""Input pipeline and output formats either side of the train step.

Reference behaviour being replaced:
  * `chunk = torch.load("{i}.pt").to(device="cpu", dtype=torch.float32)`, optional `chunk -= means`
    (big_sweep.py:400-407), then per step `dataset[batch_idxs].to(device)` - a host-side random gather of fp32
    rows plus a pageable H2D copy (big_sweep.py:170, basic_l1_sweep.py:81-103, cluster_runs.py:28-37);
  * `BatchSampler(RandomSampler(range(n)), batch_size, drop_last=False)` index batches;
  * `torch.save([(LearnedDict, hyperparams), ...])` checkpoints (big_sweep.py:245-268, :421-427).

Here the chunk is uploaded ONCE in its storage format (fp16 as activation_dataset.py:393-397 writes it:
half the PCIe and HBM bytes) and stays resident; `chunk[idxs]` gathers, converts and centres a batch on the
device (`sc_gather_rows`), and the sampler draws its permutation on the device, so a training step has no
host->device traffic at all.  `ResidentChunk` quacks like the fp32 tensor the reference loop expects
(`dataset[batch_idxs].to(device)`, `.shape`), so `ensemble_train_loop` keeps its signature.

The loop does NOT upload a host tensor behind the caller's back: for one pass over an fp32 host chunk the upload moves
the same bytes over PCIe as the per-step copies and is serial (measured: no gain).  Residency pays when the chunk is
uploaded in fp16 (half the bytes) and shared by every ensemble trained on it - which is what `train_on_chunks` does.
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _native
from ._native import ScGatherArgs, check, lib, ptr, stream_ptr


class ResidentChunk:
    """An activation chunk `[n, d]` (fp16 or fp32) resident in HBM; indexing returns fp32 batches on the device.

    means: None, a `[d]` tensor (e.g. the reference's saved `means.pt`), or True to compute the column means of
    this chunk on the device (`chunk.mean(dim=0)`, big_sweep.py:404).  Values are never modified in place - the
    subtraction happens in the gather, in fp32, exactly as `chunk.to(float32) - means` would give.
    """

    def __init__(self, data, device=None, means=None):
        if not isinstance(data, torch.Tensor) or data.dim() != 2:
            raise TypeError("ResidentChunk needs a 2-D tensor [n, d]")
        if data.dtype not in (torch.float16, torch.float32):
            raise TypeError(f"chunk dtype must be float16 or float32, got {data.dtype}")
        device = torch.device(device if device is not None else (data.device if data.is_cuda else "cuda"))
        if device.type != "cuda":
            raise NotImplementedError(f"ResidentChunk lives in HBM: device {device} is not a CUDA device (no CPU path)")
        vec = 8 if data.dtype == torch.float16 else 4
        if data.shape[1] % vec:
            raise ValueError(f"activation width must be a multiple of {vec} for {data.dtype} chunks, got {data.shape[1]}")
        self.data = data.to(device).contiguous()
        self.device = self.data.device
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.means = None
        if means is True:
            self.means = self.column_means()
        elif means is not None:
            self.means = torch.as_tensor(means, dtype=torch.float32).to(self.device).contiguous()
            if self.means.shape != (self.data.shape[1],):
                raise ValueError(f"means must have shape [{self.data.shape[1]}], got {tuple(self.means.shape)}")

    @classmethod
    def from_file(cls, path, device="cuda", means=None):
        """Reader for the reference's `{i}.pt` chunk files (one `[n, d]` tensor, big_sweep.py:400-401)."""
        t = torch.load(path, map_location="cpu")
        return cls(t, device=device, means=means)

    # -- tensor-like surface used by the reference loops ------------------------------------------------------
    @property
    def shape(self):
        return self.data.shape

    @property
    def dtype(self):
        return torch.float32  # what indexing returns

    @property
    def is_cuda(self):
        return True

    def size(self, dim=None):
        return self.data.size() if dim is None else self.data.size(dim)

    def __len__(self):
        return self.data.shape[0]

    def to(self, *args, **kwargs):
        """`dataset.to(device)` / `.to(dtype=torch.float32)` of the reference loops: already resident, already
        fp32 on read - returns self when the target is this device / float32, refuses anything else."""
        for a in list(args) + list(kwargs.values()):
            if isinstance(a, torch.dtype) and a != torch.float32:
                raise TypeError("a ResidentChunk reads as float32 only")
            if isinstance(a, (str, torch.device)) and torch.device(a).type != "cuda":
                raise NotImplementedError("a ResidentChunk stays in HBM (no CPU path)")
        return self

    def nbytes(self):
        return self.data.numel() * self.data.element_size()

    # -- device work ----------------------------------------------------------------------------------------
    def column_means(self):
        n, d = self.data.shape
        out = torch.empty(d, dtype=torch.float32, device=self.device)
        ws = torch.empty(lib().sc_column_mean_workspace_bytes(d), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            check(lib().sc_column_mean(ptr(self.data), self.data.element_size(), n, d, ptr(out), ptr(ws),
                                       stream_ptr(self.device)), "sc_column_mean")
        return out

    def _indices(self, idxs):
        n = self.data.shape[0]
        if isinstance(idxs, torch.Tensor) and idxs.is_cuda:
            if idxs.dtype != torch.int64:
                idxs = idxs.to(torch.int64)
            return idxs.contiguous(), False
        if isinstance(idxs, slice):
            idxs = range(*idxs.indices(n))
        if isinstance(idxs, (int, np.integer)):
            raise TypeError("ResidentChunk indexing takes a batch of row numbers (list, array or tensor)")
        host = torch.as_tensor(np.asarray(idxs) if not isinstance(idxs, torch.Tensor) else idxs)
        if host.dtype == torch.bool:
            raise TypeError("boolean masks are not supported; pass row numbers")
        host = host.to(torch.int64).reshape(-1)
        if host.numel() and (int(host.min()) < -n or int(host.max()) >= n):
            bad = int(host.max()) if int(host.max()) >= n else int(host.min())
            raise IndexError(f"index {bad} is out of bounds for dimension 0 with size {n}")
        return host.to(self.device, non_blocking=True), True

    def __getitem__(self, idxs):
        idx, checked = self._indices(idxs)
        n, d = self.data.shape
        out = torch.empty((idx.numel(), d), dtype=torch.float32, device=self.device)
        if idx.numel() == 0:
            return out
        a = ScGatherArgs()
        a.chunk, a.elem_bytes, a.n_rows, a.d = ptr(self.data), self.data.element_size(), n, d
        a.idx, a.B, a.mean, a.out = ptr(idx), idx.numel(), ptr(self.means), ptr(out)
        a.err_flag = None if checked else ptr(self._flag)
        with torch.cuda.device(self.device):
            a.stream = stream_ptr(self.device)
            check(lib().sc_gather_rows(C.byref(a)), "sc_gather_rows")
        return out

    def check(self):
        """Row numbers given as DEVICE tensors cannot be validated without a sync, so the kernel zero-fills the
        row and raises a flag; call this (one sync) where the reference would have raised IndexError."""
        if int(self._flag.item()) != 0:
            self._flag.zero_()
            raise IndexError(f"a device index batch held a row number outside [-{len(self)}, {len(self)})")


class DeviceBatchSampler:
    """`BatchSampler(RandomSampler(range(n)), batch_size, drop_last)` (cluster_runs.py:28-32,
    basic_l1_sweep.py:97-101) with the permutation drawn and kept on the device: yields int64 CUDA index
    tensors, every row exactly once per pass, last batch ragged unless drop_last."""

    def __init__(self, n, batch_size, drop_last=False, device="cuda", seed=None):
        if batch_size <= 0:
            raise ValueError("batch_size should be a positive integer value")
        self.n, self.batch_size, self.drop_last = int(n), int(batch_size), bool(drop_last)
        self.device = torch.device(device)
        self.generator = None
        if seed is not None:
            self.generator = torch.Generator(device=self.device)
            self.generator.manual_seed(int(seed))

    def __len__(self):
        return self.n // self.batch_size if self.drop_last else (self.n + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        perm = torch.randperm(self.n, device=self.device, generator=self.generator)
        for i in range(len(self)):
            yield perm[i * self.batch_size:(i + 1) * self.batch_size]


def chunk_sampler(dataset, batch_size, drop_last=False, seed=None):
    """The sampler a caller builds next to the chunk: on-device for a ResidentChunk, torch's own otherwise."""
    if isinstance(dataset, ResidentChunk):
        return DeviceBatchSampler(len(dataset), batch_size, drop_last, dataset.device, seed)
    return torch.utils.data.BatchSampler(torch.utils.data.RandomSampler(range(dataset.shape[0])),
                                         batch_size=batch_size, drop_last=drop_last)


def save_learned_dicts(learned_dicts, path):
    """Writer of the reference's checkpoint format (big_sweep.py:421-424, basic_l1_sweep.py:110-118): a pickled
    list of `(LearnedDict, hyperparams)`; the classes pickle under `autoencoders.learned_dict`, so the
    reference's plotting / interpretation scripts load the file unchanged."""
    for ld, hyper in learned_dicts:
        if not isinstance(hyper, dict):
            raise TypeError("each entry must be (LearnedDict, hyperparams dict)")
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    torch.save(learned_dicts, path)


def load_learned_dicts(path, map_location="cpu"):
    """Reader of the same files, including those the reference wrote (output_basic_test/*.pt)."""
    from . import install_reference_aliases
    if not install_reference_aliases():
        raise RuntimeError("another `autoencoders` package is imported; call install_reference_aliases(force=True) "
                           "to read reference pickles with this engine's classes")
    return torch.load(path, map_location=map_location, weights_only=False)


def train_on_chunks(ensembles, cfg, ensemble_hyperparams, buffer_hyperparams, batch_size=None, fista_dictionary_update=False,
                    progress=None):
    """The chunk loop of `sweep` (big_sweep.py:386-431) for one process / one GPU, without the subprocess
    dispatch and wandb: permuted chunk order (np.random.permutation after np.random.seed(0), :335, :389),
    `n_repetitions`, optional centring with the FIRST chunk's means saved to `means.pt` (:402-407), and
    `learned_dicts.pt` + `config.yaml` written to `_{i}` after chunks 8, 16, ..., 512 and the last (:421-427).

    ensembles: list of (ensemble, args, name) as an ensemble_init_func returns.  Returns the last learned_dicts.
    """
    import yaml
    from .big_sweep import ensemble_train_loop, unstacked_to_learned_dicts

    class _Counter:
        value = 0

    torch.manual_seed(0)
    np.random.seed(0)
    os.makedirs(cfg.output_folder, exist_ok=True)
    n_chunks = len([f for f in os.listdir(cfg.dataset_folder) if f.endswith(".pt")])
    chunk_order = np.random.permutation(n_chunks)
    if getattr(cfg, "n_repetitions", None) is not None:
        chunk_order = np.tile(chunk_order, cfg.n_repetitions)
    cfg.ensemble_hyperparams, cfg.buffer_hyperparams = ensemble_hyperparams, buffer_hyperparams
    means, learned_dicts = None, []
    for i, chunk_idx in enumerate(chunk_order):
        path = os.path.join(cfg.dataset_folder, f"{chunk_idx}.pt")
        device = ensembles[0][1]["device"]
        chunk = ResidentChunk.from_file(path, device=device)
        if getattr(cfg, "center_activations", False):
            if i == 0:
                means = chunk.column_means()
                torch.save(means.cpu(), os.path.join(cfg.output_folder, "means.pt"))
            chunk.means = means
        for ensemble, args, name in ensembles:
            bs = batch_size or args.get("batch_size", getattr(cfg, "batch_size", 256))
            sampler = chunk_sampler(chunk, bs)
            ensemble_train_loop(ensemble, cfg, args, name, sampler, chunk, progress or _Counter(), torch.float32,
                                fista_dictionary_update=fista_dictionary_update)
        learned_dicts = []
        for ensemble, args, _ in ensembles:
            learned_dicts.extend(unstacked_to_learned_dicts(ensemble, args, ensemble_hyperparams, buffer_hyperparams))
        del chunk
        if i == len(chunk_order) - 1 or (i + 1) in [2 ** j for j in range(3, 10)]:
            iter_folder = os.path.join(cfg.output_folder, f"_{i}")
            save_learned_dicts(learned_dicts, os.path.join(iter_folder, "learned_dicts.pt"))
            with open(os.path.join(iter_folder, "config.yaml"), "w") as f:
                plain = {k: v for k, v in (dict(cfg) if hasattr(cfg, "keys") else vars(cfg)).items()
                         if isinstance(v, (int, float, str, bool, list, type(None)))}
                yaml.dump(plain, f)
    return learned_dicts
