"""TensorFlow tensor-bundle reader (nerf_tex_b200/tf_checkpoint.py).  No TensorFlow and no TF-written file exists
here, so these are known-answer tests of the primitives (CRC-32C, varints, snappy, protobuf fields) and round trips
through the module's own writer — self-consistency with the published format, not proof of interoperability."""

import numpy as np
import pytest

from nerf_tex_b200 import tf_checkpoint as T
from oracle import nerftex_oracle as O


def test_crc32c_known_answers():
    assert T.crc32c(b"123456789") == 0xE3069283                     # the CRC-32C check value (iSCSI, RFC 3720)
    assert T.crc32c(b"") == 0
    assert T.crc32c(bytes(32)) == 0x8A9136AA                        # RFC 3720 B.4: 32 zero bytes
    assert T.crc32c(bytes([0xFF] * 32)) == 0x62A8AB43               # RFC 3720 B.4: 32 bytes of 0xFF
    assert T.crc32c(bytes(range(32))) == 0x46DD794E                 # RFC 3720 B.4: ascending bytes
    c = T.crc32c(b"foo")
    assert T.masked_crc(b"foo") == (((c >> 15) | (c << 17)) + 0xa282ead8) & 0xFFFFFFFF and T.masked_crc(b"foo") != c


def test_varints_and_proto_fields():
    for v in (0, 1, 127, 128, 300, 2 ** 32 - 1, 2 ** 63):
        enc = T._put_varint(v)
        assert T._varint(enc + b"\xAA", 0) == (v, len(enc))
    assert T._put_varint(300) == b"\xAC\x02"                        # protobuf documentation example
    e = T.Entry(1, (72, 256), 0, 123456, 72 * 256 * 4, 0xDEADBEEF)
    assert T._parse_entry(T._encode_entry(e)) == e
    assert T._parse_entry(T._encode_entry(T.Entry(9, (), 0, 0, 8, None))) == T.Entry(9, (), 0, 0, 8, None)


def test_snappy_decoder():
    # "abcdabcdabcdabcd!" as: length 17, literal "abcd", copy(offset 4, len 12) [1-byte-offset form needs len<=11,
    # so use the 2-byte-offset form], literal "!"
    stream = bytes([17]) + bytes([(4 - 1) << 2]) + b"abcd" + bytes([((12 - 1) << 2) | 2, 4, 0]) + bytes([0 << 2]) + b"!"
    assert T._snappy_decompress(stream) == b"abcdabcdabcdabcd!"
    # 1-byte-offset copy: len 7 (4..11), offset 3
    stream = bytes([10]) + bytes([(3 - 1) << 2]) + b"xyz" + bytes([((7 - 4) << 2) | 1 | (0 << 5), 3])
    assert T._snappy_decompress(stream) == b"xyzxyzxyzx"
    with pytest.raises(ValueError):
        T._snappy_decompress(bytes([5]) + bytes([((4 - 1) << 2) | 2, 9, 0]))


class FakeModel:
    """Layer table of FusedNerfTexMLP without a GPU."""

    def __init__(self, om):
        self.layer_names = [n for n, _, _ in om.layers]
        self.layers = [(0, 0, W.shape[0], W.shape[1]) for _, W, _ in om.layers]
        self.loaded = None

    def load_named_weights(self, named):
        self.loaded = named


def keras_keys(om, name, order):
    out = {}
    for i, lname in enumerate(order):
        W, b = om.named_weights()[lname]
        out[f"{name}/layer_with_weights-{i}/kernel{T.VALUE_SUFFIX}"] = W
        out[f"{name}/layer_with_weights-{i}/bias{T.VALUE_SUFFIX}"] = b
    return out


@pytest.mark.parametrize("block_size", [64, 4096])
def test_bundle_round_trip_and_layer_matching(tmp_path, block_size):
    om, om_f = O.ParamNerfOracle((1, 4), seed=1), O.ParamNerfOracle((1, 4), seed=2)
    om.randomize_biases(3, 0.1)
    creation = [n for n, _, _ in om.layers]                                      # trunk.., alpha, feature, colorhid0, colorhalf, rgb
    by_depth = [n for n in creation if n != "alpha" and n != "rgb"] + ["alpha", "rgb"]   # Keras' depth-sorted numbering
    tensors = {"step" + T.VALUE_SUFFIX: np.array(4321, np.int64), "_CHECKPOINTABLE_OBJECT_GRAPH": np.frombuffer(b"\x0a\x00", np.uint8)}
    tensors.update(keras_keys(om, "model", creation))
    tensors.update(keras_keys(om_f, "model_fine", by_depth))
    prefix = str(tmp_path / "checkpoints" / "ckpt-4")
    T.write_bundle(prefix, tensors, block_size=block_size, restart_interval=3)
    with open(tmp_path / "checkpoints" / "checkpoint", "w") as f:
        f.write('model_checkpoint_path: "ckpt-4"\nall_model_checkpoint_paths: "ckpt-4"\n')
    assert T.latest_checkpoint(str(tmp_path / "checkpoints")) == prefix
    n_shards, index = T.read_index(prefix)
    assert n_shards == 1 and set(index) == set(tensors)
    for k, a in tensors.items():
        got = T.read_tensor(prefix, index[k], n_shards, verify=True)
        assert got.dtype == a.dtype and got.shape == a.shape and np.array_equal(got, a)
    assert T.read_step(prefix) == 4321
    for name, ref in (("model", om), ("model_fine", om_f)):                      # both numberings give the same layers
        fm = FakeModel(ref)
        T.load_model_weights(prefix, fm, name, verify=True)
        for lname, (W, b) in ref.named_weights().items():
            assert np.array_equal(fm.loaded[lname][0], W) and np.array_equal(fm.loaded[lname][1], b), lname


def test_corruption_is_detected(tmp_path):
    prefix = str(tmp_path / "ckpt-1")
    T.write_bundle(prefix, {"a": np.arange(10, dtype=np.float32), "b": np.ones((3, 2), np.float32)})
    raw = bytearray(open(prefix + ".index", "rb").read())
    raw[5] ^= 0x40
    open(prefix + ".index", "wb").write(bytes(raw))
    with pytest.raises(ValueError):
        T.read_index(prefix)
    T.write_bundle(prefix, {"a": np.arange(10, dtype=np.float32)})
    n, index = T.read_index(prefix)
    data = bytearray(open(prefix + ".data-00000-of-00001", "rb").read())
    data[3] ^= 1
    open(prefix + ".data-00000-of-00001", "wb").write(bytes(data))
    with pytest.raises(ValueError):
        T.read_tensor(prefix, index["a"], n, verify=True)
    with pytest.raises(ValueError):
        T.read_table(prefix + ".data-00000-of-00001")                            # not a table at all
    fm = FakeModel(O.ParamNerfOracle((1, 4), seed=1))
    with pytest.raises(ValueError):
        T.load_model_weights(prefix, fm, "model")
