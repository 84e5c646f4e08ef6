"""utils.py of the reference (utils.py:20-37): top-k correctness of a batch of logits.

Same name, arguments and result keys as `topk_correct(logits, labels, mask=None, prefix='',
topk=(1, 5))`; each value is a float vector [batch] of 0/1 flags (times `mask`).  The ranking runs
as one warp-per-example kernel (sav_topk_correct) instead of argsort + vmapped isin; ties are
broken like the stable ascending `jnp.argsort` the reference uses (larger class index ranks higher).
"""
import ctypes as C

import torch

from sav import _abi
from sav.tape import Var


def topk_correct(logits, labels, mask=None, prefix='', topk=(1, 5)):
    t = logits.v if isinstance(logits, Var) else logits
    if not (t.is_cuda and t.dtype == torch.bfloat16 and t.dim() == 2 and t.stride(1) == 1):
        raise _abi.SavError("topk_correct: logits must be a CUDA bf16 [batch, classes] tensor")
    ks = tuple(int(k) for k in topk)
    if not 1 <= len(ks) <= 4:
        raise ValueError("topk_correct: between 1 and 4 values of k")
    B, classes = t.shape
    out = torch.empty((len(ks), B), dtype=torch.float32, device=t.device)
    pad = ks + (0,) * (4 - len(ks))
    lab = labels.to(device=t.device, dtype=torch.int32)
    _abi.check(_abi.lib().sav_topk_correct(
        C.c_void_p(t.data_ptr()), t.stride(0), C.c_void_p(lab.data_ptr()), B, classes, len(ks), *pad,
        C.c_void_p(out.data_ptr()), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    metrics = {}
    for i, k in enumerate(ks):
        correct = out[i]
        if mask is not None:
            correct = correct * mask.to(device=t.device, dtype=torch.float32)
        metrics[f'{prefix}top_{k}_acc'] = correct
    return metrics
