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
    att = torch.empty(ids.shape, dtype=torch.uint8, device=ids.device)
    with torch.cuda.device(ids.device):
        lib.check(lib.init(ids.device.index).vt_mlm_mask(ids.data_ptr(), lib.ptr(tc), lut.data_ptr(), ids.numel(), lut.numel(), int(mask_token_id),
                                                         int(pad_token_id), float(mlm_probability), int(seed) & (2 ** 64 - 1), out.data_ptr(),
                                                         labels.data_ptr(), att.data_ptr(), lib.stream_ptr()), "vt_mlm_mask")
    return out, labels, att.bool()


def draws(n: int, vocab_size: int, seed: int, device) -> Tuple[torch.Tensor, torch.Tensor]:
    """The (u1, u2, u3) uniforms [3, n] and random words [n] ``mask_tokens`` uses for this seed (for parity tests)."""
    u = torch.empty((3, n), dtype=torch.float32, device=device)
    r = torch.empty(n, dtype=torch.int64, device=device)
    with torch.cuda.device(device):
        lib.check(lib.init(torch.device(device).index).vt_mlm_draws(n, vocab_size, int(seed) & (2 ** 64 - 1), u.data_ptr(), r.data_ptr(),
                                                                   lib.stream_ptr()), "vt_mlm_draws")
    return u, r
