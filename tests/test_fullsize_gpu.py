"""Pixel parity on the paths that produce the benchmark numbers: batches large enough for the pipeline to cut them
into chunks that rotate over its lanes, bitstreams in pinned memory (copied from where they lie) and in pageable
memory (staged by helper threads), the automatic subsequence-length choice at full batch sizes, and the full-size
shapes of BASELINE.json's configs.  Every image is compared with the reference's own libjpeg-turbo path
(oracle/_ref/libref_turbo.so, the reference's jpeg_mem.cpp compiled here) when it is built, else with the oracle;
references are computed once per distinct stream.  Bar: bit-exact
(test/extensions/common_ext_decoder_test.h:187-196 with eps lowered to 0)."""
import ctypes
import os
import subprocess
import zlib

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DRIVER = os.path.join(ROOT, "oracle", "_ref", "host_driver")
EXT_DIR = os.path.join(ROOT, "nvimagecodec_b200", "lib")


def _reference(data, fmt="I_RGB", fancy=True, roi=None):
    from oracle import turbo

    if turbo.available() and fmt in ("I_RGB", "I_Y"):
        return turbo.decode(data, fmt, fancy, roi)
    return oracle.decode(data, {"I_RGB": oracle.FMT_I_RGB, "I_Y": oracle.FMT_Y}[fmt], fancy=fancy, roi=roi)


@pytest.fixture(scope="module")
def c2_streams():
    """20 distinct streams of BASELINE config #2's shape (1920x1080 4:2:0 q90, no RST) and their reference pixels."""
    from nvimagecodec_b200 import corpus

    datas = [corpus.encode(corpus.synth(1080, 1920, 3000 + s), 90, "420") for s in range(20)]
    return datas, [_reference(d) for d in datas]


def _pinned_sources(datas):
    import torch

    keep = [torch.frombuffer(bytearray(d), dtype=torch.uint8).pin_memory() for d in datas]
    return keep, [t.numpy() for t in keep]


@pytest.mark.parametrize("to_host", [False, True])
def test_chunked_pipeline_from_pinned_bitstreams(decoder_cls, c2_streams, to_host):
    """160 x 1080p (78 MB) from pinned buffers through b200jpegDecodeBatch with the default six lanes: ~10 chunks of
    8 MB, every lane is reused, every bitstream is copied straight from the caller's buffer.  Device and host outputs."""
    import torch
    from nvimagecodec_b200 import _lib

    datas, refs = c2_streams
    order = [(7 * i + i // 20) % 20 for i in range(160)]
    keep, srcs = _pinned_sources([datas[k] for k in order])
    assert sum(len(datas[k]) for k in order) > 64 << 20
    dec = decoder_cls(0, options=":fancy_upsampling=1")
    with torch.cuda.device(0):
        batch = dec.make_batch(srcs, "I_RGB", to_host=to_host)
        batch.storage.fill_(0)
        for rep in range(2):  # the second call reuses every lane's arenas
            status = dec.run(batch)
            torch.cuda.synchronize()
            assert status == [_lib.OK] * len(srcs)
            launches = dec.stat(_lib.STAT_KERNEL_LAUNCHES)
            assert launches >= 9 * 8, "the batch was not cut into chunks"
            dev_refs = {}
            for k, img in zip(order, batch.images):
                if to_host:
                    assert np.array_equal(img.numpy(), refs[k])
                else:
                    if k not in dev_refs:
                        dev_refs[k] = torch.from_numpy(refs[k]).cuda()
                    assert torch.equal(img, dev_refs[k])
            batch.storage.fill_(0)
    dec.close()


def test_chunked_pipeline_from_pageable_bitstreams(decoder_cls, c2_streams):
    """The same from pageable memory (numpy arrays): 24 MB chunks gathered into the pinned staging block by the helper threads."""
    import torch
    from nvimagecodec_b200 import _lib

    datas, refs = c2_streams
    order = [(3 * i) % 20 for i in range(120)]
    srcs = [np.frombuffer(datas[k], dtype=np.uint8).copy() for k in order]
    dec = decoder_cls(0, options=":fancy_upsampling=1")
    outs = dec.decode(srcs, "I_RGB")
    torch.cuda.synchronize()
    assert dec.last_status == [_lib.OK] * len(srcs)
    assert dec.stat(_lib.STAT_KERNEL_LAUNCHES) >= 2 * 8
    dev_refs = {k: torch.from_numpy(refs[k]).cuda() for k in set(order)}
    for k, o in zip(order, outs):
        assert torch.equal(o, dev_refs[k])
    dec.close()


@pytest.mark.parametrize("n_images,expect_bits", [(110, 2048), (420, 4096)])
def test_automatic_subsequence_length_at_full_batch_sizes(decoder_cls, c2_streams, n_images, expect_bits):
    """Prepare / Execute / Finish on a whole batch (what the device-resident bench leg times): batches of >= 400 Mbit
    and >= 1600 Mbit make the host pick 2048- and 4096-bit subsequences by itself (decoder.cpp, prepare())."""
    import torch
    from nvimagecodec_b200 import _lib

    datas, refs = c2_streams
    order = [(11 * i + i // 20) % 20 for i in range(n_images)]
    bits = 8 * sum(len(datas[k]) for k in order)
    assert bits >= ((400 if expect_bits == 2048 else 1600) << 20)
    if expect_bits == 2048:
        assert bits < (1600 << 20)
    srcs = [datas[k] for k in order]
    dec = decoder_cls(0, options=":fancy_upsampling=1")
    with torch.cuda.device(0):
        batch = dec.make_batch(srcs, "I_RGB")
        dec.prepare(batch)
        dec.execute()
        status = dec.finish(batch)
        torch.cuda.synchronize()
    assert status == [_lib.OK] * len(srcs)
    n_sub = dec.stat(_lib.STAT_NUM_SUBSEQ)
    want = sum((len(datas[k]) * 8 + expect_bits - 1) // expect_bits for k in order)  # upper bound from the stuffed lengths
    assert 0.9 * want <= n_sub <= want + len(order), (n_sub, want)
    dev_refs = {k: torch.from_numpy(refs[k]).cuda() for k in set(order)}
    for k, img in zip(order, batch.images):
        assert torch.equal(img, dev_refs[k])
    dec.close()


def test_c4_shape_3840x2160_444_q95_no_rst(decoder_cls):
    """BASELINE config #4's real shape: one 6.5 MB scan without restart markers, self-synchronisation only; alone (512-bit
    subsequences), in a batch of eight (2048-bit), and cropped."""
    import torch
    from nvimagecodec_b200 import _lib, corpus

    datas = [corpus.encode(corpus.synth(2160, 3840, 5000 + s), 95, "444") for s in range(2)]
    refs = [_reference(d) for d in datas]
    dec = decoder_cls(0, options=":fancy_upsampling=1")
    out = dec.decode(datas[:1], "I_RGB")
    assert dec.last_status == [_lib.OK]
    assert np.array_equal(out[0].cpu().numpy(), refs[0])
    order = [0, 1, 1, 0, 1, 0, 0, 1]
    with torch.cuda.device(0):
        batch = dec.make_batch([datas[k] for k in order], "I_RGB")
        dec.prepare(batch)
        dec.execute()
        assert dec.finish(batch) == [_lib.OK] * len(order)
        torch.cuda.synchronize()
    dev_refs = [torch.from_numpy(r).cuda() for r in refs]
    for k, img in zip(order, batch.images):
        assert torch.equal(img, dev_refs[k])
    roi = (1000, 37, 2159, 3801)
    crop = dec.decode(datas[1:], "I_RGB", [roi])[0]
    assert np.array_equal(crop.cpu().numpy(), refs[1][roi[0]:roi[2], roi[1]:roi[3]])
    dec.close()


def test_c5_shape_1024_mixed_images_roi_and_gray(decoder_cls):
    """BASELINE config #5's shape at 1024 images: C3-distributed small images + 10 % 1080p, half of them with a ROI,
    30 % gray output, one call through the pipeline.  Statuses for all, pixels for a seeded subset (ROI == crop of the
    full decode, Y output == Y plane) and for every 1080p image."""
    import torch
    from nvimagecodec_b200 import _lib, corpus

    jpegs, reqs = corpus.config_c5(1024)
    fmts = [r["fmt"] for r in reqs]
    rois = [r["roi"] for r in reqs]
    dec = decoder_cls(0)  # box upsampling, the binding default
    outs = dec.decode(jpegs, fmts, rois)
    torch.cuda.synchronize()
    assert dec.last_status == [_lib.OK] * len(jpegs)
    rng = np.random.default_rng(5)
    pick = sorted(set(int(i) for i in rng.integers(0, len(jpegs), size=96)) | set(range(9, len(jpegs), 97)))
    cache = {}
    for i in pick:
        key = (id(jpegs[i]), fmts[i])
        if key not in cache:
            cache[key] = _reference(jpegs[i], fmts[i], fancy=False)
        full = cache[key]
        r = rois[i]
        want = full if r is None else full[r[0]:r[2], r[1]:r[3]]
        assert np.array_equal(outs[i].cpu().numpy(), want), i
    dec.close()


@pytest.mark.skipif(not os.path.exists(DRIVER), reason="reference core not built (oracle/build_ref.sh)")
def test_framework_batch_large_enough_to_split(tmp_path, c2_streams):
    """80 x 1080p (39 MB, pageable io_stream memory) through the unmodified core's nvimgcodecDecoderDecode: the
    extension's pipeline splits it, stages with the framework's executor threads, reuses the marker walks of canDecode,
    and takes every arena from the allocators the host passes in nvimgcodecExecutionParams_t."""
    datas, refs = c2_streams
    for k, d in enumerate(datas):
        (tmp_path / f"s{k}.jpg").write_bytes(d)
    order = [(9 * i + i // 20) % 20 for i in range(80)]
    lines = [f"{tmp_path / ('s%d.jpg' % k)} I_RGB 0 0 0 0 {'host' if i % 16 == 5 else 'device'}" for i, k in enumerate(order)]
    (tmp_path / "manifest.txt").write_text("\n".join(lines) + "\n")
    out = tmp_path / "out"
    out.mkdir()
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.path.dirname(DRIVER) + ":" + env.get("LD_LIBRARY_PATH", "")
    env["HOST_DRIVER_CRC"] = "1"
    env["HOST_DRIVER_COUNT_ALLOC"] = "1"
    r = subprocess.run([DRIVER, EXT_DIR, str(tmp_path / "manifest.txt"), str(out), ":fancy_upsampling=1", "2"], capture_output=True, text=True,
                       env=env, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    want = [zlib.crc32(np.ascontiguousarray(x).tobytes()) for x in refs]
    rows = [l.split() for l in (out / "status.txt").read_text().splitlines()]
    assert len(rows) == len(order)
    for k, row in zip(order, rows):
        assert int(row[0]) == 1, row
        assert int(row[3]) == want[k]
    alloc = [l for l in r.stdout.splitlines() if l.startswith("ALLOC")]
    assert alloc, r.stdout[-500:]
    counts = dict(kv.split("=") for kv in alloc[0].split()[1:])
    # every device / pinned block came from (and went back to) the host's allocators
    assert int(counts["device_malloc"]) >= 6 and int(counts["device_malloc"]) == int(counts["device_free"])
    assert int(counts["device_bytes"]) == int(counts["device_bytes_freed"])
    assert int(counts["pinned_malloc"]) >= 2 and int(counts["pinned_malloc"]) == int(counts["pinned_free"])


def test_caller_allocators_through_the_c_abi(decoder_cls):
    """b200jpegCreateEx: every arena is taken from the caller's callbacks (stream-ordered, size handed back on free)."""
    import torch
    from nvimagecodec_b200 import _lib, corpus

    live, calls = {}, {"dm": 0, "df": 0, "pm": 0, "pf": 0}
    rt = ctypes.CDLL("libcudart.so.12")
    rt.cudaMalloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t]
    rt.cudaFree.argtypes = [ctypes.c_void_p]
    rt.cudaMallocHost.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t]
    rt.cudaFreeHost.argtypes = [ctypes.c_void_p]

    def dmalloc(ctx, pptr, size, stream):
        p = ctypes.c_void_p()
        if rt.cudaMalloc(ctypes.byref(p), size) != 0:
            return 1
        pptr[0] = p.value
        live[p.value] = size
        calls["dm"] += 1
        return 0

    def dfree(ctx, ptr, size, stream):
        assert live.pop(ptr) == size
        calls["df"] += 1
        return 0 if rt.cudaFree(ptr) == 0 else 1

    def pmalloc(ctx, pptr, size, stream):
        p = ctypes.c_void_p()
        if rt.cudaMallocHost(ctypes.byref(p), size) != 0:
            return 1
        pptr[0] = p.value
        live[p.value] = size
        calls["pm"] += 1
        return 0

    def pfree(ctx, ptr, size, stream):
        assert live.pop(ptr) == size
        calls["pf"] += 1
        return 0 if rt.cudaFreeHost(ptr) == 0 else 1

    al = _lib.Allocators(None, _lib.ALLOC_FN(dmalloc), _lib.FREE_FN(dfree), None, _lib.ALLOC_FN(pmalloc), _lib.FREE_FN(pfree))
    srcs = [corpus.encode(corpus.synth(200, 300, 4100 + i), 85, ["420", "444", "422"][i % 3], i) for i in range(6)]
    dec = decoder_cls(0, options=":fancy_upsampling=1", allocators=al)
    for _ in range(2):
        outs = dec.decode(srcs, "I_RGB")
        torch.cuda.synchronize()
        assert dec.last_status == [_lib.OK] * len(srcs)
        for d, o in zip(srcs, outs):
            assert np.array_equal(o.cpu().numpy(), oracle.decode(d, oracle.FMT_I_RGB, True))
    assert calls["dm"] >= 5 and calls["pm"] >= 2
    dec.close()
    assert not live and calls["dm"] == calls["df"] and calls["pm"] == calls["pf"]
