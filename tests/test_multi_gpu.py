"""In-process multi-device driver (b200jpegMulti*): plan by compressed bytes, decode shares concurrently."""
import numpy as np
import pytest

import oracle


def test_plan_matches_python_sharding_without_gpu():
    """The C planner and nvimagecodec_b200.sharding implement the same deterministic LPT rule."""
    import ctypes

    from nvimagecodec_b200 import _lib, sharding

    L = _lib.load()
    import torch

    if not torch.cuda.is_available():
        pytest.skip("b200jpegMultiCreate needs a device; the Python planner is covered by test_capi/test_sharding_gloo")
    rng = np.random.default_rng(2)
    sizes = [int(x) for x in rng.integers(1000, 900000, size=777)]
    h = ctypes.c_void_p()
    ids = (ctypes.c_int * 3)(0, 0, 0)
    assert L.b200jpegMultiCreate(ctypes.byref(h), ids, 3, b"") == _lib.OK
    lens = (ctypes.c_size_t * len(sizes))(*sizes)
    dev = (ctypes.c_int * len(sizes))()
    assert L.b200jpegMultiPlan(h, len(sizes), lens, dev) == _lib.OK
    parts = sharding.partition_by_bytes(sizes, 3)
    mine = [[i for i in range(len(sizes)) if dev[i] == d] for d in range(3)]
    assert mine == parts
    L.b200jpegMultiDestroy(h)


@pytest.mark.gpu
def test_multi_device_decode_bit_exact():
    import torch

    from nvimagecodec_b200 import MultiGpuDecoder, _lib, corpus

    ngpu = torch.cuda.device_count()
    devices = list(range(ngpu)) if ngpu > 1 else [0, 0]  # two lanes on one GPU still exercise planner + threads
    rng = np.random.default_rng(4)
    srcs = []
    for i in range(40):
        h, w = int(rng.integers(30, 500)), int(rng.integers(30, 500))
        srcs.append(corpus.encode(corpus.synth(h, w, 1200 + i), int(rng.integers(60, 97)), ["420", "444", "422"][i % 3], int(rng.integers(0, 9))))
    srcs.append(corpus.encode(corpus.synth(1080, 1920, 5), 90, "420"))
    dec = MultiGpuDecoder(devices, options=":fancy_upsampling=1")
    imgs, st, dev = dec.decode(srcs, "I_RGB")
    assert st == [_lib.OK] * len(srcs)
    assert len(set(dev)) == len(set(range(len(devices))))  # every lane got work
    loads = [sum(len(s) for s, d in zip(srcs, dev) if d == k) for k in range(len(devices))]
    assert max(loads) <= 1.6 * (sum(loads) / len(loads)) + max(len(s) for s in srcs)
    for s, im in zip(srcs, imgs):
        assert np.array_equal(im.cpu().numpy(), oracle.decode(s, oracle.FMT_I_RGB, True))
    dec.close()
