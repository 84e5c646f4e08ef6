"""Checkpoint save / restore in the Flax layout (train.py:123-127, simple_train.py:112-116):
`flax.training.checkpoints.save_checkpoint(ckpt_dir, state, step, keep=3)` writes
`<ckpt_dir>/checkpoint_<step>` = `flax.serialization.to_bytes(state)`, msgpack of
`to_state_dict(TrainState)`, and keeps the newest `keep`.

Wire format (flax.serialization, restated from its published source; flax is not installable here
so byte compatibility is by construction -- parity unpinned -- and pinned against a golden byte
fixture built from that spec, tests/test_checkpoint_cpu.py):
  * ExtType 1 (ndarray)       = packb((shape, dtype.name, tobytes('C')), use_bin_type=True)
  * ExtType 2 (native complex) = packb((real, imag))
  * ExtType 3 (numpy scalar)   = the ndarray encoding of np.asarray(x); decodes to arr[()]
  * arrays over 2**30 bytes are split: {'__msgpack_chunked_array__': True, 'shape': [...],
    'chunks': {'0': ndarray, ...}} (decoded here; never produced at these model sizes)
State dict of the reference's TrainState (train.py:22-38): the params field holds the WHOLE
`model.init()` variables, so it is {'params': {'params': tree}}; opt_state is the optax.chain tuple
serialised by index: '0' clip_by_global_norm -> {}, '1' scale_by_adam -> {'count','mu','nu'} with
mu / nu shaped like params (same extra 'params' level), '2' additive_weight_decay -> {},
'3' scale_by_schedule -> {'count'}.  `step` and the counts are int32 scalars (0-d arrays after
jax.device_get, train.py:125).  The reference never restores; restore_checkpoint is
provided for resume and accepts reference-written files."""
import os
import re

import msgpack
import numpy as np

_EXT_NDARRAY = 1
_EXT_COMPLEX = 2
_EXT_NPSCALAR = 3


def _ndarray_to_bytes(arr):
    arr = np.ascontiguousarray(arr) if arr.ndim else np.asarray(arr)
    return msgpack.packb((list(arr.shape), arr.dtype.name, arr.tobytes("C")), use_bin_type=True)


def _ndarray_from_bytes(data):
    shape, dtype, buf = msgpack.unpackb(data, raw=True)
    dtype = dtype.decode() if isinstance(dtype, bytes) else dtype
    return np.frombuffer(buf, dtype=np.dtype(dtype)).reshape(shape).copy()


def _default(x):
    if isinstance(x, np.ndarray):
        return msgpack.ExtType(_EXT_NDARRAY, _ndarray_to_bytes(x))
    if isinstance(x, np.generic):
        return msgpack.ExtType(_EXT_NPSCALAR, _ndarray_to_bytes(np.asarray(x)))
    if isinstance(x, complex):
        return msgpack.ExtType(_EXT_COMPLEX, msgpack.packb((x.real, x.imag)))
    raise TypeError(f"cannot serialise {type(x)}")


def _ext_hook(code, data):
    if code == _EXT_NDARRAY:
        return _ndarray_from_bytes(data)
    if code == _EXT_COMPLEX:
        return complex(*msgpack.unpackb(data))
    if code == _EXT_NPSCALAR:
        return _ndarray_from_bytes(data)[()]
    return msgpack.ExtType(code, data)


def _unchunk(tree):
    if isinstance(tree, dict):
        if tree.get("__msgpack_chunked_array__"):
            chunks = tree["chunks"]
            flat = np.concatenate([np.asarray(chunks[str(i)]).reshape(-1) for i in range(len(chunks))])
            return flat.reshape(tree["shape"])
        return {k: _unchunk(v) for k, v in tree.items()}
    return tree


def to_bytes(state_dict):
    return msgpack.packb(state_dict, default=_default, use_bin_type=True, strict_types=True)


def from_bytes(data):
    return _unchunk(msgpack.unpackb(data, ext_hook=_ext_hook, raw=False, strict_map_key=False))


def _tree_to_numpy(tree):
    if isinstance(tree, dict):
        return {k: _tree_to_numpy(v) for k, v in tree.items()}
    return tree.detach().float().cpu().numpy()


def _i32(x):
    return np.asarray(int(x), dtype=np.int32)


def state_dict_of(train_state):
    """Host copy of a sav.train.TrainState in the reference TrainState's state-dict nesting."""
    arena = train_state.arena
    step = _i32(train_state.step)
    sd = {"step": step, "params": {"params": _tree_to_numpy(arena.tree())}}
    if getattr(train_state, "batch_stats", None):
        # the reference keeps the whole model.init() variables under `params` (train.py:31-37), so
        # the BatchNorm running averages sit beside the parameter tree
        sd["params"]["batch_stats"] = _tree_to_numpy(train_state.batch_stats)
    if arena.m is not None:
        sd["opt_state"] = {
            "0": {},
            "1": {"count": step.copy(),
                  "mu": {"params": _tree_to_numpy(arena._tree_of(arena.m))},
                  "nu": {"params": _tree_to_numpy(arena._tree_of(arena.v))}},
            "2": {},
            "3": {"count": step.copy()},
        }
    return sd


def _strip_params_level(tree):
    """The reference stores whole variables ({'params': tree}); older files of this repo stored the
    tree directly.  Return the parameter tree either way."""
    if isinstance(tree, dict) and "params" in tree and set(tree.keys()) <= {"params", "batch_stats"} \
            and isinstance(tree["params"], dict):
        return tree["params"]
    return tree


def _adam_state(opt_state):
    """(mu, nu, count) from either the optax.chain nesting ('0'..'3') or this repo's round-1 layout."""
    if "1" in opt_state and isinstance(opt_state["1"], dict) and "mu" in opt_state["1"]:
        o = opt_state["1"]
    else:
        o = opt_state
    return _strip_params_level(o["mu"]), _strip_params_level(o["nu"]), int(np.asarray(o["count"]))


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
        arena.load_tree(_map(_strip_params_level(sd["params"]), torch.from_numpy))
        saved_stats = sd["params"].get("batch_stats") if isinstance(sd["params"], dict) else None
        if saved_stats is not None and getattr(train_state, "batch_stats", None):
            _copy_into(train_state.batch_stats, saved_stats)
        if "opt_state" in sd:
            mu, nu, _count = _adam_state(sd["opt_state"])
            arena.ensure_moments()
            arena.load_tree_into(arena.m, _map(mu, torch.from_numpy))
            arena.load_tree_into(arena.v, _map(nu, torch.from_numpy))
        step = int(np.asarray(sd["step"]))
        if hasattr(train_state, "set_step"):
            train_state.set_step(step)
        else:
            train_state.step = step
    return sd


def _copy_into(dst_tree, src_tree):
    """In-place load of a nested dict of numpy arrays into same-named device tensors."""
    import torch
    for k, v in dst_tree.items():
        if isinstance(v, dict):
            _copy_into(v, src_tree[k])
        else:
            v.copy_(torch.from_numpy(np.ascontiguousarray(src_tree[k])).to(v.dtype))


def _map(tree, fn):
    if isinstance(tree, dict):
        return {k: _map(v, fn) for k, v in tree.items()}
    return fn(tree)
