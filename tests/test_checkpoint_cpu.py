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


def _reference_shaped_state():
    """What flax.serialization.to_state_dict gives for the reference's TrainState after
    jax.device_get (train.py:22-38,123-127): whole variables under params, optax.chain by index."""
    rng = np.random.default_rng(0)
    tree = lambda: {"params": {"Dense_0": {"kernel": rng.standard_normal((4, 3)).astype(np.float32),
                                           "bias": rng.standard_normal(3).astype(np.float32)}}}
    return {"step": np.asarray(5, np.int32), "params": tree(),
            "opt_state": {"0": {}, "1": {"count": np.asarray(5, np.int32), "mu": tree(), "nu": tree()},
                          "2": {}, "3": {"count": np.asarray(5, np.int32)}}}


def test_golden_bytes_from_the_published_flax_spec():
    """Byte-level pin, hand-assembled from the flax.serialization format: a 0-d int32 `step` and a
    [2] float32 leaf.  fixmap(2) | 'a' -> ext8(type 1) of packb(([2], 'float32', 8 raw bytes)) |
    'step' -> ext8(type 1) of packb(([], 'int32', 4 raw bytes))."""
    leaf = np.asarray([1.0, -2.0], np.float32)
    inner_a = b"\x93" + b"\x91\x02" + b"\xa7float32" + b"\xc4\x08" + leaf.tobytes()
    inner_s = b"\x93" + b"\x90" + b"\xa5int32" + b"\xc4\x04" + np.asarray(7, np.int32).tobytes()
    golden = (b"\x82" + b"\xa1a" + b"\xc7" + bytes([len(inner_a)]) + b"\x01" + inner_a
              + b"\xa4step" + b"\xc7" + bytes([len(inner_s)]) + b"\x01" + inner_s)
    assert C.to_bytes({"a": leaf, "step": np.asarray(7, np.int32)}) == golden
    back = C.from_bytes(golden)
    assert int(back["step"]) == 7 and back["a"].tolist() == [1.0, -2.0]


def test_reference_shaped_dict_roundtrip_and_ext_types():
    sd = _reference_shaped_state()
    back = C.from_bytes(C.to_bytes(sd))
    assert set(back["opt_state"]) == {"0", "1", "2", "3"} and back["opt_state"]["0"] == {}
    assert int(back["opt_state"]["3"]["count"]) == 5
    np.testing.assert_array_equal(back["params"]["params"]["Dense_0"]["kernel"],
                                  sd["params"]["params"]["Dense_0"]["kernel"])
    mu, nu, count = C._adam_state(back["opt_state"])
    assert count == 5 and set(mu) == {"Dense_0"} and set(nu) == {"Dense_0"}
    assert set(C._strip_params_level(back["params"])) == {"Dense_0"}
    # ExtType 3 (numpy scalar, how flax writes np.generic) and ExtType 2 (native complex) decode
    blob = C.to_bytes({"s": np.float32(2.5), "i": np.int32(9), "c": complex(1.0, -3.0)})
    raw = msgpack.unpackb(blob, raw=False, strict_map_key=False)
    assert raw["s"].code == 3 and raw["c"].code == 2
    got = C.from_bytes(blob)
    assert got["s"] == np.float32(2.5) and got["i"] == 9 and got["c"] == complex(1.0, -3.0)
    # chunked arrays (flax splits leaves over 2**30 bytes) are reassembled
    big = {"__msgpack_chunked_array__": True, "shape": [2, 3],
           "chunks": {"0": np.arange(4, dtype=np.float32), "1": np.arange(4, 6, dtype=np.float32)}}
    assert C.from_bytes(C.to_bytes({"w": big}))["w"].tolist() == [[0, 1, 2], [3, 4, 5]]


def test_state_dict_keeps_batch_stats_beside_the_parameters():
    """The reference TrainState stores the whole model.init() variables under `params` (train.py:31-37):
    for a BatchNorm model that is {'params': ..., 'batch_stats': ...}.  Both nestings restore."""
    import numpy as np
    from sav import checkpoint as C
    tree = {"Dense_0": {"kernel": np.ones((2, 3), np.float32)}}
    stats = {"BatchNorm_0": {"mean": np.zeros(3, np.float32), "var": np.ones(3, np.float32)}}
    assert C._strip_params_level({"params": tree}) is tree
    assert C._strip_params_level({"params": tree, "batch_stats": stats}) is tree
    assert C._strip_params_level(tree) is tree                      # round-1 files: the tree itself
    sd = {"step": np.asarray(3, np.int32), "params": {"params": tree, "batch_stats": stats}}
    back = C.from_bytes(C.to_bytes(sd))
    assert set(back["params"]) == {"params", "batch_stats"}
    np.testing.assert_array_equal(back["params"]["batch_stats"]["BatchNorm_0"]["var"], stats["BatchNorm_0"]["var"])
