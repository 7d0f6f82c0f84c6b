"""TF V2 checkpoint bundle reader/writer against the reference's shipped fixture."""
import os

import numpy as np
import pytest

from actor_critic_algs_on_tensorflow_b200 import tf_bundle as B

EXPECTED = {  # name: (shape, masked crc32c) -- Appendix A
    "global_actor/Variable": ((), 0x0EE58BFB), "global_actor/Variable_1": ((), 0xF2C9EA08),
    "global_actor/Variable_2": ((), 0x2BDAA581), "global_actor/first_layer/bias": ((128,), 0x2E100A59),
    "global_actor/first_layer/kernel": ((3, 128), 0xFEC743EF), "global_actor/log_std": ((1,), 0x5011CE3C),
    "global_actor/mu_layer/bias": ((1,), 0xE430F9E3), "global_actor/mu_layer/kernel": ((64, 1), 0x2816198E),
    "global_actor/second_layer/bias": ((128,), 0x80D0FADD), "global_actor/second_layer/kernel": ((128, 128), 0xD7D76A5F),
    "global_actor/third_layer/bias": ((64,), 0xB7311B99), "global_actor/third_layer/kernel": ((128, 64), 0x92549B9F),
    "global_critic/Variable": ((), 0x3255CB69), "global_critic/Variable_1": ((), 0x3255CB69),
    "global_critic/first_layer/bias": ((256,), 0xC0B0E0E2), "global_critic/first_layer/kernel": ((3, 256), 0x239AF8D3),
    "global_critic/second_layer/bias": ((128,), 0x7D5E040E), "global_critic/second_layer/kernel": ((256, 128), 0xE333CDF0),
    "global_critic/third_layer/bias": ((128,), 0x0F8998EB), "global_critic/third_layer/kernel": ((128, 128), 0x699424BE),
    "global_critic/value/bias": ((1,), 0x8527746F), "global_critic/value/kernel": ((128, 1), 0x8960032F)}


def test_reads_the_shipped_checkpoint(golden_dir):
    t = B.read_bundle(os.path.join(golden_dir, "model-Pendulum_a3c"))
    assert list(t) == sorted(EXPECTED)                   # byte-sorted key order, 22 tensors
    for k, (shape, crc) in EXPECTED.items():
        assert t[k].shape == shape and t[k].dtype == np.float32
        assert B.masked_crc32c(t[k].tobytes()) == crc
    assert t["global_actor/Variable"] == np.float32(0.005) and t["global_actor/Variable_1"] == np.float32(0.01)
    assert t["global_actor/Variable_2"] == np.float32(1.0) and t["global_critic/Variable"] == np.float32(0.001)
    np.testing.assert_allclose(t["global_actor/log_std"], [-0.03052692], rtol=1e-6)
    np.testing.assert_allclose(t["global_critic/value/bias"], [-0.0658476], rtol=1e-5)
    assert sum(v.size for k, v in t.items() if "Variable" not in k) == 75907


def test_write_is_byte_identical_to_tensorflow(golden_dir, tmp_path):
    src = os.path.join(golden_dir, "model-Pendulum_a3c")
    t = B.read_bundle(src)
    dst = str(tmp_path / "ckpt" / "model-Pendulum-600")
    B.write_bundle(dst, t)
    for ext in (".index", ".data-00000-of-00001"):
        assert open(dst + ext, "rb").read() == open(src + ext, "rb").read()


def test_round_trip_other_shapes_and_many_keys(tmp_path):
    rng = np.random.default_rng(0)
    t = {"Actor/logits/kernel": rng.normal(size=(64, 3)).astype(np.float32), "Actor/Variable": np.float32(0.5),
         "global_step": np.int64(1234), "x/empty": np.zeros((0, 4), np.float32)}
    for i in range(40):                                   # > 2 restart intervals, shared prefixes
        t["scope/layer_%02d/kernel" % i] = rng.normal(size=(i % 5 + 1, 3)).astype(np.float32)
    p = str(tmp_path / "m")
    B.write_bundle(p, t)
    back = B.read_bundle(p)
    assert list(back) == sorted(t, key=lambda s: s.encode())
    for k in t:
        np.testing.assert_array_equal(back[k], np.asarray(t[k]))
        assert back[k].dtype == np.asarray(t[k]).dtype


def test_corruption_is_detected(golden_dir, tmp_path):
    src = os.path.join(golden_dir, "model-Pendulum_a3c")
    p = str(tmp_path / "bad")
    data = bytearray(open(src + ".data-00000-of-00001", "rb").read())
    data[5000] ^= 0xFF
    open(p + ".data-00000-of-00001", "wb").write(bytes(data))
    open(p + ".index", "wb").write(open(src + ".index", "rb").read())
    with pytest.raises(ValueError, match="crc32c"):
        B.read_bundle(p)
    assert B.read_bundle(p, verify=False)["global_actor/log_std"].shape == (1,)
    idx = bytearray(open(src + ".index", "rb").read())
    idx[100] ^= 0x01
    open(p + ".index", "wb").write(bytes(idx))
    with pytest.raises(ValueError, match="checksum"):
        B.read_bundle(p)
    open(p + ".index", "wb").write(b"not a table" * 10)
    with pytest.raises(ValueError, match="magic"):
        B.read_bundle(p)
    with pytest.raises(FileNotFoundError):
        B.read_bundle(str(tmp_path / "missing"))


def test_crc32c_known_answers():
    assert B.crc32c(b"123456789") == 0xE3069283           # the standard CRC-32C check value
    assert B.crc32c(b"") == 0 and B.crc32c(bytes(32)) == 0x8A9136AA


def test_oracle_side_checkpoint_table_matches_the_bundle_reader(golden_dir):
    """oracle/demo_ckpt.py (offset table of SURVEY.md Appendix A, used by bench.py's CPU arms so that they do not import
    the product) and the product's TF V2 bundle reader give the same 22 tensors."""
    import os
    import numpy as np
    from oracle.demo_ckpt import read_demo_checkpoint
    from actor_critic_algs_on_tensorflow_b200.tf_bundle import read_bundle
    prefix = os.path.join(golden_dir, "model-Pendulum_a3c")
    a, b = read_demo_checkpoint(prefix), read_bundle(prefix)
    assert set(a) == set(b) and len(a) == 22
    for k in a:
        np.testing.assert_array_equal(a[k], b[k])
