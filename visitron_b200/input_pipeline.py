"""The step before forward: masked-LM masking of a pre-training batch on the GPU.

Replaces ``PretrainDataset._mask_tokens`` (tasks/viewpoint_select/data_loader_pretrain.py:549-612), which runs per
SAMPLE on the host CPU (three ``torch.bernoulli`` passes, a ``randint`` and five boolean index assignments), with one
kernel launch over the whole [batch, text_len] block; same rule: 15 % of the non-special tokens are chosen, 80 % of them
become ``[MASK]``, 10 % a random word, 10 % stay; with ``masked_token_prediction`` every region-label position
(``token_classes != -1``) is forced to ``[MASK]`` with its MLM label ignored; attention = not ``[PAD]``.
CUDA tensors only; there is no CPU fallback.
"""
from __future__ import annotations

from typing import Iterable, Optional, Tuple

import torch

from . import lib


def special_lut(vocab_size: int, special_ids: Iterable[int], device) -> torch.Tensor:
    """uint8 [vocab]: 1 for ``tokenizer.all_special_ids``."""
    t = torch.zeros(vocab_size, dtype=torch.uint8)
    t[torch.as_tensor(sorted(set(int(i) for i in special_ids)), dtype=torch.long)] = 1
    return t.to(device)


def mask_tokens(inputs: torch.Tensor, token_classes: Optional[torch.Tensor], lut: torch.Tensor, mask_token_id: int, pad_token_id: int,
                seed: int, mlm_probability: float = 0.15) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """-> (inputs', labels, attention_mask) like the reference's ``_mask_tokens`` (the input tensor is left untouched)."""
    lib.require_device(inputs)
    if inputs.dtype != torch.int64 or (token_classes is not None and token_classes.dtype != torch.int64):
        raise lib.VtError("mask_tokens expects int64 token ids / classes")
    ids = inputs.contiguous()
    tc = None if token_classes is None else token_classes.contiguous()
    if tc is not None and tc.shape != ids.shape:
        raise ValueError("token_classes must have the shape of inputs")
    out, labels = torch.empty_like(ids), torch.empty_like(ids)
    att = torch.empty(ids.shape, dtype=torch.bool, device=ids.device)   # the kernel writes 0 / 1 bytes
    with torch.cuda.device(ids.device):
        lib.check(lib.init(ids.device.index).vt_mlm_mask(ids.data_ptr(), lib.ptr(tc), lut.data_ptr(), ids.numel(), lut.numel(), int(mask_token_id),
                                                         int(pad_token_id), float(mlm_probability), int(seed) & (2 ** 64 - 1), out.data_ptr(),
                                                         labels.data_ptr(), att.data_ptr(), lib.stream_ptr()), "vt_mlm_mask")
    return out, labels, att


def draws(n: int, vocab_size: int, seed: int, device) -> Tuple[torch.Tensor, torch.Tensor]:
    """The (u1, u2, u3) uniforms [3, n] and random words [n] ``mask_tokens`` uses for this seed (for parity tests)."""
    u = torch.empty((3, n), dtype=torch.float32, device=device)
    r = torch.empty(n, dtype=torch.int64, device=device)
    with torch.cuda.device(device):
        lib.check(lib.init(torch.device(device).index).vt_mlm_draws(n, vocab_size, int(seed) & (2 ** 64 - 1), u.data_ptr(), r.data_ptr(),
                                                                   lib.stream_ptr()), "vt_mlm_draws")
    return u, r


def pack_regions(feats: torch.Tensor, loc: torch.Tensor, offsets: torch.Tensor, regions: int,
                 attention_out: Optional[torch.Tensor] = None, attention_col: int = 0) -> Tuple[torch.Tensor, torch.Tensor]:
    """Ragged region rows -> ([B, regions, D], [B, regions, Dl]): keep the LAST ``regions`` rows of a sample that has more, zero-pad
    one that has fewer (data_loader_pretrain.py:656-690), one launch for both arrays.  ``offsets`` int64 [B + 1] on the device;
    with ``attention_out`` ([B, >= attention_col + regions] int64) columns [attention_col, attention_col + regions) receive 1 / 0."""
    lib.require_device(feats)
    if feats.dtype != torch.float32 or loc.dtype != torch.float32 or offsets.dtype != torch.int64:
        raise lib.VtError("pack_regions expects fp32 rows and int64 offsets")
    if feats.dim() != 2 or loc.dim() != 2 or feats.shape[0] != loc.shape[0]:
        raise ValueError("feats [rows, D] and loc [rows, Dl] must have the same number of rows")
    feats, loc, offsets = feats.contiguous(), loc.contiguous(), offsets.contiguous()
    B = offsets.numel() - 1
    out_f = torch.empty((B, regions, feats.shape[1]), dtype=torch.float32, device=feats.device)
    out_l = torch.empty((B, regions, loc.shape[1]), dtype=torch.float32, device=feats.device)
    att_ptr, ld = None, 0
    if attention_out is not None:
        if attention_out.dtype != torch.int64 or attention_out.stride(-1) != 1 or attention_out.shape[1] < attention_col + regions:
            raise lib.VtError("attention_out must be a row-major int64 [B, >= attention_col + regions] tensor")
        att_ptr, ld = attention_out.data_ptr() + 8 * attention_col, attention_out.stride(0)
    with torch.cuda.device(feats.device):
        lib.check(lib.init(feats.device.index).vt_pack_regions(lib.ptr(feats) or None, feats.shape[1], lib.ptr(loc) or None, loc.shape[1], offsets.data_ptr(), B,
                                                               regions, lib.ptr(out_f) or None, lib.ptr(out_l) or None, att_ptr, ld, lib.stream_ptr()),
                  "vt_pack_regions")
    return out_f, out_l


def collate_pretrain(token_ids: torch.Tensor, token_classes: torch.Tensor, feats: torch.Tensor, loc: torch.Tensor, offsets: torch.Tensor,
                     next_action: torch.Tensor, lut: torch.Tensor, mask_token_id: int, pad_token_id: int, seed: int, max_img_seq_length: int,
                     masked_token_prediction: bool = True, no_action_grounding: bool = False, mlm_probability: float = 0.15) -> dict:
    """A whole pre-training batch from its raw pieces, on the device, in two launches plus three small concatenations:
    what ``PretrainDataset._preprocess_item`` (data_loader_pretrain.py:636-712) does per sample on the host, with the same keys
    (``PreTrainOscar.forward(**batch)`` takes them as they are)."""
    B, T = token_ids.shape
    R = int(max_img_seq_length)
    ids, labels, att_text = mask_tokens(token_ids, token_classes if masked_token_prediction else None, lut, mask_token_id, pad_token_id, seed,
                                        mlm_probability)
    attention = torch.empty((B, T + R), dtype=torch.int64, device=ids.device)
    attention[:, :T] = att_text
    img, loc_out = pack_regions(feats, loc, offsets, R, attention_out=attention, attention_col=T)
    tail = torch.full((B, R), -1, dtype=torch.int64, device=ids.device)
    batch = dict(input_ids=ids, labels=torch.cat([labels, tail], 1), attention_mask=attention, img_feats=img, img_location_embeddings=loc_out,
                 token_labels=torch.cat([token_classes, tail], 1) if masked_token_prediction else None,
                 next_action=torch.full_like(next_action, -1) if no_action_grounding else next_action)
    return batch


class DevicePrefetcher:
    """Pinned host batches -> device, one batch ahead, on a copy stream (the loader side of vs/pretrain.py:166-176, where
    every tensor of a batch is moved with a blocking ``.to(device)`` at the top of the step).  Iterating yields device dictionaries whose
    H2D copy overlapped the previous step's compute; the consumer's stream waits on the copy's event and the tensors are tied to it with
    ``record_stream`` so the caching allocator does not hand their memory to the next copy early.  ``None`` values pass through."""

    def __init__(self, batches, device, pin: bool = True):
        self.it = iter(batches)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise lib.VtError("DevicePrefetcher moves batches to a CUDA device (there is no CPU path)")
        self.pin = pin
        self.stream = torch.cuda.Stream(self.device)
        self.slot = None
        self._fetch()

    def _fetch(self):
        try:
            host = next(self.it)
        except StopIteration:
            self.slot = None
            return
        if self.pin:
            host = {k: (v if v is None or not torch.is_tensor(v) or v.is_pinned() else v.pin_memory()) for k, v in host.items()}
        with torch.cuda.stream(self.stream):
            dev = {k: (v.to(self.device, non_blocking=True) if torch.is_tensor(v) else v) for k, v in host.items()}
            self.slot = (dev, self.stream.record_event(), host)   # the pinned source stays referenced until the copy has been consumed

    def __iter__(self):
        return self

    def __next__(self):
        if self.slot is None:
            raise StopIteration
        dev, ev, _ = self.slot
        self._fetch()                                   # the next batch's copy overlaps this step
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        for t in dev.values():
            if torch.is_tensor(t):
                t.record_stream(cur)
        return dev
