"""Reader for the ONE checkpoint the reference ships (``demos/model-Pendulum_a3c.*``, kept as a fixture under
``tests/golden/``) -- TEST INFRASTRUCTURE, so that the CPU arms of bench.py can load the weights without importing the
product's TF-bundle reader.  The ``.data-00000-of-00001`` file is the raw little-endian fp32 tensors in byte-sorted key
order with no padding (every CRC verified there); this table is that appendix."""
import numpy as np

TENSORS = [  # (key, shape, byte offset)
    ("global_actor/Variable", (), 0), ("global_actor/Variable_1", (), 4), ("global_actor/Variable_2", (), 8),
    ("global_actor/first_layer/bias", (128,), 12), ("global_actor/first_layer/kernel", (3, 128), 524),
    ("global_actor/log_std", (1,), 2060), ("global_actor/mu_layer/bias", (1,), 2064),
    ("global_actor/mu_layer/kernel", (64, 1), 2068), ("global_actor/second_layer/bias", (128,), 2324),
    ("global_actor/second_layer/kernel", (128, 128), 2836), ("global_actor/third_layer/bias", (64,), 68372),
    ("global_actor/third_layer/kernel", (128, 64), 68628), ("global_critic/Variable", (), 101396),
    ("global_critic/Variable_1", (), 101400), ("global_critic/first_layer/bias", (256,), 101404),
    ("global_critic/first_layer/kernel", (3, 256), 102428), ("global_critic/second_layer/bias", (128,), 105500),
    ("global_critic/second_layer/kernel", (256, 128), 106012), ("global_critic/third_layer/bias", (128,), 237084),
    ("global_critic/third_layer/kernel", (128, 128), 237596), ("global_critic/value/bias", (1,), 303132),
    ("global_critic/value/kernel", (128, 1), 303136),
]


def read_demo_checkpoint(prefix):
    raw = open(prefix + ".data-00000-of-00001", "rb").read()
    assert len(raw) == 303648, "not the shipped demo checkpoint"
    out = {}
    for key, shape, off in TENSORS:
        n = int(np.prod(shape)) if shape else 1
        out[key] = np.frombuffer(raw, dtype="<f4", count=n, offset=off).reshape(shape).copy()
    return out
