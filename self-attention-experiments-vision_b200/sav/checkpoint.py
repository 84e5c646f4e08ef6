"""Checkpoint save / restore in the Flax layout (train.py:123-127, simple_train.py:112-116):
`flax.training.checkpoints.save_checkpoint(ckpt_dir, state, step, keep=3)` writes
`<ckpt_dir>/checkpoint_<step>` = msgpack of the state dict with every array as
ExtType(1, packb((shape, dtype.name, bytes))) (flax.serialization) and keeps the newest `keep`.

Flax is not installable here, so byte-level compatibility is by construction from the published
format, not verified against the library (parity unpinned); the reference never restores
(SURVEY 5), restore_checkpoint is provided for resume.  State dict written:
{'step': int, 'params': <Flax-named tree, fp32>, 'opt_state': {'mu': tree, 'nu': tree, 'count'}}.
"""
import os
import re

import msgpack
import numpy as np

_EXT_NDARRAY = 1


def _pack_array(arr):
    arr = np.ascontiguousarray(arr)
    return msgpack.ExtType(_EXT_NDARRAY, msgpack.packb((list(arr.shape), arr.dtype.name, arr.tobytes("C")),
                                                       use_bin_type=True))


def _default(x):
    if isinstance(x, np.ndarray):
        return _pack_array(x)
    if isinstance(x, np.generic):
        return _pack_array(np.asarray(x))
    raise TypeError(f"cannot serialise {type(x)}")


def _ext_hook(code, data):
    if code == _EXT_NDARRAY:
        shape, dtype, buf = msgpack.unpackb(data, raw=False)
        return np.frombuffer(buf, dtype=np.dtype(dtype)).reshape(shape).copy()
    return msgpack.ExtType(code, data)


def to_bytes(state_dict):
    return msgpack.packb(state_dict, default=_default, use_bin_type=True, strict_types=True)


def from_bytes(data):
    return msgpack.unpackb(data, ext_hook=_ext_hook, raw=False, strict_map_key=False)


def _tree_to_numpy(tree):
    if isinstance(tree, dict):
        return {k: _tree_to_numpy(v) for k, v in tree.items()}
    return tree.detach().float().cpu().numpy()


def state_dict_of(train_state):
    """Host copy of a sav.train.TrainState (step, params, Adam moments) as nested numpy dicts."""
    arena = train_state.arena
    sd = {"step": int(train_state.step), "params": _tree_to_numpy(arena.tree())}
    if arena.m is not None:
        sd["opt_state"] = {"mu": _tree_to_numpy(arena._tree_of(arena.m)),
                           "nu": _tree_to_numpy(arena._tree_of(arena.v)),
                           "count": int(train_state.step)}
    return sd


def save_checkpoint(ckpt_dir, target, step, prefix="checkpoint_", keep=3, overwrite=False):
    """target: a TrainState or an already-built state dict.  Returns the file written."""
    sd = target if isinstance(target, dict) else state_dict_of(target)
    os.makedirs(ckpt_dir, exist_ok=True)
    path = os.path.join(ckpt_dir, f"{prefix}{step}")
    if os.path.exists(path) and not overwrite:
        raise FileExistsError(path)
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(to_bytes(sd))
    os.replace(tmp, path)
    steps = sorted(int(m.group(1)) for m in (re.fullmatch(re.escape(prefix) + r"(\d+)", n)
                                             for n in os.listdir(ckpt_dir)) if m)
    for old in steps[:-keep] if keep else []:
        os.remove(os.path.join(ckpt_dir, f"{prefix}{old}"))
    return path


def latest_checkpoint(ckpt_dir, prefix="checkpoint_"):
    if not os.path.isdir(ckpt_dir):
        return None
    steps = [int(m.group(1)) for m in (re.fullmatch(re.escape(prefix) + r"(\d+)", n)
                                       for n in os.listdir(ckpt_dir)) if m]
    return os.path.join(ckpt_dir, f"{prefix}{max(steps)}") if steps else None


def restore_checkpoint(ckpt_dir_or_file, train_state=None, prefix="checkpoint_"):
    """Returns the state dict; if a TrainState is given, loads params / moments / step into it."""
    path = ckpt_dir_or_file
    if os.path.isdir(path):
        path = latest_checkpoint(path, prefix)
        if path is None:
            return None
    with open(path, "rb") as f:
        sd = from_bytes(f.read())
    if train_state is not None:
        import torch
        arena = train_state.arena
        arena.load_tree(_map(sd["params"], torch.from_numpy))
        if "opt_state" in sd:
            arena.ensure_moments()
            arena.load_tree_into(arena.m, _map(sd["opt_state"]["mu"], torch.from_numpy))
            arena.load_tree_into(arena.v, _map(sd["opt_state"]["nu"], torch.from_numpy))
        train_state.step = int(sd["step"])
    return sd


def _map(tree, fn):
    if isinstance(tree, dict):
        return {k: _map(v, fn) for k, v in tree.items()}
    return fn(tree)
