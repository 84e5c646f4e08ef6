"""Checkpoint file format (train.py:123-127 -> flax.training.checkpoints / flax.serialization):
msgpack with ndarray ExtType(1, packb((shape, dtype.name, bytes))), files `checkpoint_<step>`,
newest `keep` retained.  Host-only logic; the device round trip is in the GPU tests."""
import os

import msgpack
import numpy as np

from sav import checkpoint as C


def _state(step):
    rng = np.random.default_rng(step)
    return {"step": step,
            "params": {"Dense_0": {"kernel": rng.standard_normal((4, 3)).astype(np.float32),
                                   "bias": np.zeros(3, np.float32)},
                       "cls": rng.standard_normal((1, 1, 4)).astype(np.float32)}}


def test_roundtrip_and_wire_format():
    sd = _state(7)
    blob = C.to_bytes(sd)
    back = C.from_bytes(blob)
    assert back["step"] == 7
    np.testing.assert_array_equal(back["params"]["Dense_0"]["kernel"], sd["params"]["Dense_0"]["kernel"])
    assert back["params"]["cls"].shape == (1, 1, 4)
    # wire format: arrays are ExtType 1 wrapping packb((shape, dtype name, raw bytes))
    raw = msgpack.unpackb(blob, raw=False, strict_map_key=False)
    ext = raw["params"]["Dense_0"]["bias"]
    assert isinstance(ext, msgpack.ExtType) and ext.code == 1
    shape, dtype, buf = msgpack.unpackb(ext.data, raw=False)
    assert shape == [3] and dtype == "float32" and len(buf) == 12


def test_keep_newest_three(tmp_path):
    for step in (10, 20, 30, 40, 50):
        C.save_checkpoint(str(tmp_path), _state(step), step, keep=3)
    assert sorted(os.listdir(tmp_path)) == ["checkpoint_30", "checkpoint_40", "checkpoint_50"]
    assert C.latest_checkpoint(str(tmp_path)).endswith("checkpoint_50")
    assert C.restore_checkpoint(str(tmp_path))["step"] == 50
