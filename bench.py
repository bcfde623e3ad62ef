#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric (decoded MPix/s, images/s) on its config #2:
batch 256 x 1920x1080 baseline 4:2:0 q90 synthetic JPEG -> interleaved RGB, per B200.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # libjpeg-turbo on the host cores

A "step" is one pass of the hot path over one batch.  `value` times the kernels only
(bitstreams already resident in HBM, outputs left in HBM).  `e2e` is the same metric through the
reference-facing plugin call -- nvimgcodecDecoderDecode of the UNMODIFIED nvImageCodec core
(oracle/_ref/libnvimgcodec.so, built from the reference sources) with this repo's extension picked
up by priority -- with HOST buffers: pinned bitstreams in, pinned host pixels out, H2D and D2H
inside the timed region (tests/host/host_driver.cpp drives it; one driver process per rank).
`e2e_capi` is the same through the plain C ABI (b200jpegDecodeBatch, pinned bitstreams in, pinned
pixels out); the *_device_out variants leave the pixels in HBM (what a GPU consumer sees).
Weak scaling: every rank decodes its own 256-image shard of a 256*N batch that is partitioned
by compressed byte count (nvimagecodec_b200/sharding.py); no data-path collective.  `strong` is a
fixed heterogeneous 2048-image batch (C3 distribution + 10 % 1080p) sharded over the N ranks by the
same rule.  `nvjpeg_ext` is the reference's own CUDA decoder extension through the same framework call.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "c2": "C2: 256 x 1920x1080 baseline JPEG 4:2:0 q90 (no RST) -> I_RGB uint8",
    "c3": "C3: 256 ImageNet-shaped (~500x375, 15% spread) JPEGs, mixed 4:4:4/4:2:2/4:2:0, q75-95, with RST markers -> I_RGB",
    "c4": "C4: 64 x 3840x2160 baseline JPEG 4:4:4 q95, no restart markers -> I_RGB",
    "c5": "C5: 8192 mixed images (C3 distribution + 10% 1080p), 50% with ROI, 30% grayscale output",
}
WORKLOAD = WORKLOADS["c2"]
HEIGHT, WIDTH, PER_GPU = 1080, 1920, 256


def _gen_spec(spec):
    from nvimagecodec_b200 import corpus

    h, w, seed, q, samp, rst = spec
    return corpus.encode(corpus.synth(h, w, seed), q, samp, rst_interval=rst)


def _gen_progressive(seed):
    from nvimagecodec_b200 import corpus

    return corpus.encode(corpus.synth(HEIGHT, WIDTH, seed), 90, "420", progressive=True)


def _gen_one(seed):
    return _gen_spec((HEIGHT, WIDTH, seed, 90, "420", 0))


def _pmap(fn, items, workers):
    from concurrent.futures import ProcessPoolExecutor

    workers = max(1, min(workers, len(items)))
    if workers == 1:
        return [fn(x) for x in items]
    with ProcessPoolExecutor(max_workers=workers) as ex:
        return list(ex.map(fn, items, chunksize=max(1, len(items) // (workers * 4))))


def make_corpus(n, workers):
    """n distinct seeded 1080p JPEGs, generated in parallel on the host (outside any timed region)."""
    return _pmap(_gen_one, list(range(n)), workers)


def c3_specs(n, seed=1234):
    import numpy as np

    rng = np.random.default_rng(seed)
    specs = []
    for i in range(n):
        h = int(np.clip(rng.normal(375, 0.15 * 375), 64, 1024))
        w = int(np.clip(rng.normal(500, 0.15 * 500), 64, 1024))
        samp = ["444", "422", "420"][i % 3]
        q = [75, 85, 95][(i // 3) % 3]
        hmax = 2 if samp in ("422", "420") else 1
        row = (w + 8 * hmax - 1) // (8 * hmax)
        specs.append((h, w, seed * 100003 + i, q, samp, [row, 8, 32][(i // 9) % 3]))
    return specs


def make_workload(name, n, workers):
    """Returns (jpegs, formats, rois, output pixel count) for one GPU's share of the named config."""
    import numpy as np

    from nvimagecodec_b200 import jpeg_host

    if name == "c2":
        jp = make_corpus(n, workers)
        return jp, ["I_RGB"] * n, [None] * n
    if name == "c3":
        jp = _pmap(_gen_spec, c3_specs(n), workers)
        return jp, ["I_RGB"] * n, [None] * n
    if name == "c4":
        jp = _pmap(_gen_spec, [(2160, 3840, 5000 + s, 95, "444", 0) for s in range(n)], workers)
        return jp, ["I_RGB"] * n, [None] * n
    # c5: a pool of distinct streams cycled to n images, random in-bounds ROI on half, gray output on 30%
    small = _pmap(_gen_spec, c3_specs(min(n, 461), seed=77), workers)
    big = _pmap(_gen_spec, [(1080, 1920, 9000 + s, 90, "420", 0) for s in range(max(1, min(n, 512) // 10))], workers)
    rng = np.random.default_rng(77)
    jp, fmts, rois = [], [], []
    for i in range(n):
        d = big[(i // 10) % len(big)] if i % 10 == 9 else small[i % len(small)]
        info = jpeg_host.probe(d)
        roi = None
        if rng.uniform() < 0.5:
            rh, rw = max(1, int(info.height * rng.uniform(0.25, 0.75))), max(1, int(info.width * rng.uniform(0.25, 0.75)))
            y0, x0 = int(rng.integers(0, info.height - rh + 1)), int(rng.integers(0, info.width - rw + 1))
            roi = (y0, x0, y0 + rh, x0 + rw)
        jp.append(d)
        fmts.append("I_Y" if rng.uniform() < 0.3 else "I_RGB")
        rois.append(roi)
    return jp, fmts, rois


DRIVER = os.path.join(ROOT, "oracle", "_ref", "host_driver")
EXT_DIR = os.path.join(ROOT, "nvimagecodec_b200", "lib")
NVJPEG_EXT_DIR = os.path.join(ROOT, "oracle", "_ref", "extensions_nvjpeg")


def reference_pixels(data, fmt="I_RGB", fancy=True, roi=None):
    """Checker for the parity spot checks (never inside a timed region): the reference's own libjpeg-turbo path when it
    was built, else the oracle."""
    try:
        from oracle import turbo

        if turbo.available():
            return turbo.decode(data, fmt, fancy, roi)
    except Exception:
        pass
    import oracle

    return oracle.decode(data, {"I_RGB": oracle.FMT_I_RGB, "I_Y": oracle.FMT_Y}[fmt], fancy=fancy, roi=roi)


class PluginDriver:
    """The reference-facing call: tests/host/host_driver.cpp drives nvimgcodecDecoderDecode of the unmodified core
    (oracle/_ref/libnvimgcodec.so) with an extension directory of our choice, on pageable host bitstreams."""

    def __init__(self, jpegs, device_index):
        import tempfile

        self.n = len(jpegs)
        self.dir = tempfile.mkdtemp(prefix="b200bench_")
        self.device_index = device_index
        seen = {}
        self.paths = []
        for i, d in enumerate(jpegs):
            if id(d) not in seen:
                p = os.path.join(self.dir, f"{len(seen)}.jpg")
                with open(p, "wb") as f:
                    f.write(d)
                seen[id(d)] = p
            self.paths.append(seen[id(d)])

    @staticmethod
    def available():
        return os.path.exists(DRIVER) and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libnvimgcodec.so"))

    def run(self, where, options, repeat, ext_dir=EXT_DIR, backend="any", fmts=None, rois=None, crc_every=0, timeout=600, pinned_in=True, threads=0):
        man = os.path.join(self.dir, f"manifest_{where}.txt")
        with open(man, "w") as f:
            for i, p in enumerate(self.paths):
                r = rois[i] if rois and rois[i] else (0, 0, 0, 0)
                f.write(f"{p} {fmts[i] if fmts else 'I_RGB'} {r[0]} {r[1]} {r[2]} {r[3]} {where}\n")
        out = os.path.join(self.dir, "out_" + where)
        os.makedirs(out, exist_ok=True)
        env = dict(os.environ)
        env["LD_LIBRARY_PATH"] = os.path.dirname(DRIVER) + ":/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
        env["CUDA_VISIBLE_DEVICES"] = str(self.device_index)
        env["HOST_DRIVER_PINNED"] = "1"
        env.pop("HOST_DRIVER_PINNED_INPUT", None)
        if pinned_in:
            env["HOST_DRIVER_PINNED_INPUT"] = "1"
        if threads:
            env["HOST_DRIVER_THREADS"] = str(threads)
        env["HOST_DRIVER_CRC"] = str(crc_every if crc_every else 1 << 30)
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
            env.pop(k, None)
        try:
            r = subprocess.run([DRIVER, ext_dir, man, out, options, str(repeat), backend], capture_output=True, text=True, env=env, timeout=timeout)
        except Exception as e:
            return {"error": type(e).__name__}
        line = [l for l in r.stdout.splitlines() if l.startswith("BENCH")]
        if r.returncode != 0 or not line:
            return {"error": "rc=%d %s" % (r.returncode, (r.stdout + r.stderr)[-200:].replace("\n", " | "))}
        kv = dict(x.split("=") for x in line[0].split()[1:])
        res = {"images_per_s": float(kv["images_per_s"]), "mpix_per_s": float(kv["mpix_per_s"]), "ms_per_batch": float(kv["ms_per_batch"]),
               "seconds": float(kv["seconds"]), "ok": int(kv["ok"]), "images": int(kv["images"]), "repeat": int(kv["repeat"])}
        if crc_every:
            rows = [l.split() for l in open(os.path.join(out, "status.txt")).read().splitlines()]
            res["crc"] = {i: int(rows[i][3]) for i in range(0, len(rows), crc_every)}
        return res

    def close(self):
        import shutil

        shutil.rmtree(self.dir, ignore_errors=True)


def nvjpeg_hw_backend_status():
    """What NvJpegHwDecoderPlugin::isPlatformSupported() asks (extensions/nvjpeg/hw_decoder.cpp:282-292):
    nvjpegCreateEx(NVJPEG_BACKEND_HARDWARE).  Returns the nvjpegStatus_t (0 = the hardware engine is usable)."""
    import ctypes

    try:
        nv = ctypes.CDLL("libnvjpeg.so.12")
    except OSError:
        try:
            nv = ctypes.CDLL("/usr/local/cuda/lib64/libnvjpeg.so")
        except OSError:
            return None
    h = ctypes.c_void_p()
    st = nv.nvjpegCreateEx(3, None, None, 0, ctypes.byref(h))
    if st == 0:
        nv.nvjpegDestroy(h)
    return int(st)



def pinned_near_gpu(nbytes, device):
    """A pinned host buffer for the decoded pixels on the memory node the GPU reads and writes fastest.  On these (virtualised,
    two-node) boxes the node a pinned allocation lands on is a matter of which CPU the allocating thread happens to run on, and
    the far node halves the D2H rate (measured: 48-70 ms instead of 33 ms per 1.59 GB step).  So: allocate under each candidate
    CPU set, time a D2H copy, keep the fastest.  Returns (tensor, {"gbs": [...], "chosen": i})."""
    import torch

    cpus = sorted(os.sched_getaffinity(0))
    half = len(cpus) // 2
    candidates = [set(cpus)] if half < 2 else [set(cpus[:half]), set(cpus[half:])]
    probe_bytes = min(nbytes, 256 << 20)
    src = torch.empty(probe_bytes, dtype=torch.uint8, device=device)
    best, rates = None, []
    try:
        for cs in candidates:
            os.sched_setaffinity(0, cs)
            t = torch.empty(max(nbytes, 1), dtype=torch.uint8, pin_memory=True)
            t[:probe_bytes].copy_(src, non_blocking=True)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(3):
                t[:probe_bytes].copy_(src, non_blocking=True)
            torch.cuda.synchronize()
            gbs = 3 * probe_bytes / (time.perf_counter() - t0) / 1e9
            rates.append(round(gbs, 1))
            if best is None or gbs > best[0]:
                best = (gbs, t)
            del t
    finally:
        os.sched_setaffinity(0, set(cpus))
    return best[1], {"d2h_gbs_per_candidate_node": rates, "chosen": rates.index(max(rates))}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def cpu_reference_decode(jpegs, threads, passes):
    """libjpeg-turbo (ISLOW + fancy upsampling == libjpeg_turbo_ext defaults) on the host cores.
    Uses the reference-side shim under oracle/ when it is built, else cv2.imdecode (libjpeg-turbo 3.1.2)."""
    import numpy as np

    try:
        from oracle import turbo  # noqa: F401

        if turbo.available():
            return turbo.time_batch(jpegs, threads, passes)
    except Exception:
        pass
    import cv2
    from concurrent.futures import ThreadPoolExecutor

    cv2.setNumThreads(1)
    arrs = [np.frombuffer(j, np.uint8) for j in jpegs]

    def one(a):
        return cv2.imdecode(a, cv2.IMREAD_COLOR).shape[0]

    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(one, arrs[:threads]))  # warm-up
        t0 = time.perf_counter()
        for _ in range(passes):
            list(ex.map(one, arrs))
        dt = time.perf_counter() - t0
    return dt, "libjpeg-turbo 3.1.2 via cv2.imdecode (BGR out, ISLOW, fancy upsampling)"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = os.cpu_count() or 1
    n = PER_GPU  # the whole batch of the configuration, as the CUDA arm decodes per step
    jpegs = make_corpus(n, cores)
    mpix = n * HEIGHT * WIDTH / 1e6
    for _ in range(args.warmup):
        cpu_reference_decode(jpegs, cores, 1)
    t0 = time.perf_counter()
    how = ""
    for _ in range(args.steps):
        _, how = cpu_reference_decode(jpegs, cores, 1)
    dt = time.perf_counter() - t0
    val = mpix * args.steps / dt
    line = {
        "impl": "reference", "metric": "decoded_mpix_per_s", "value": val, "unit": "MPix/s", "images_per_s": n * args.steps / dt,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "images_per_gpu": n, "upsampling": "fancy",
                   "note": "host threads are started inside every timed call (oracle/ref_turbo_shim.cpp); at >= 100 ms per step that is noise"},
        "cpu_baseline": {"value": val, "unit": "MPix/s", "cores": cores, "kind": "reference", "sample": f"all {n} images per step; {how}"},
        "e2e": {"value": val, "unit": "MPix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=PER_GPU, help="images per GPU (256 = the configuration the metric is quoted on)")
    ap.add_argument("--fancy", type=int, default=1, help="upsampling mode; 1 = libjpeg_turbo_ext default, what the CPU arm uses")
    ap.add_argument("--options", default="", help="extra decoder options, e.g. ' :subseq_bits=4096' (space-separated) (experiments only)")
    ap.add_argument("--no-extras", action="store_true", help="skip the strong-scaling, nvjpeg_ext and other-config extras")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS), help="c2 is the configuration the metric is quoted on; "
                    "the others are the remaining BASELINE.json configs, for extra lines only")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist

    from nvimagecodec_b200 import Decoder, _lib, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cores = os.cpu_count() or 1
    # ---- corpus: 256 distinct seeded images (worker processes are forked before CUDA is touched);
    # the N-GPU global batch is N copies, LPT-sharded by compressed bytes
    if args.workload != "c2" and args.batch == PER_GPU:
        args.batch = {"c3": 256, "c4": 64, "c5": 8192 // max(world, 1)}[args.workload]
    base, base_fmt, base_roi = make_workload(args.workload, args.batch, max(1, cores // world))
    extra_corpora = {}  # generated here, before CUDA is touched (the generators fork worker processes)
    if args.workload == "c2" and not args.no_extras:
        extra_corpora["c5"] = make_workload("c5", 2048, max(1, cores // world))
        if world == 1:
            extra_corpora["c3"] = make_workload("c3", 256, cores)
            extra_corpora["c4"] = make_workload("c4", 32, cores)
            pj = _pmap(_gen_progressive, list(range(64)), cores)  # 64 distinct progressive 1080p streams, cycled to 256
            extra_corpora["progressive"] = ([pj[i % 64] for i in range(256)], ["I_RGB"] * 256, [None] * 256)
    assert torch.cuda.is_available(), "bench.py needs a GPU (the decoder has no CPU path)"
    torch.cuda.set_device(local)
    # No CPU pinning per rank: measured on the 2-GPU box, the NVML CPU set of a GPU put the rank's pinned pixel buffer on the far
    # memory node (D2H 48 ms instead of 33 ms per 1.59 GB step); on the 8-GPU box (one node) it changed nothing.
    numa = "not pinned"
    # stdout carries exactly one JSON line: whatever native libraries print meanwhile (NCCL's version banner under
    # NCCL_DEBUG=VERSION, for one) goes to stderr
    sys.stdout.flush()
    stdout_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    global_sizes = [len(base[i % args.batch]) for i in range(args.batch * world)]
    mine = sharding.shard_for_rank(global_sizes, rank, world)
    jpegs = [base[i % args.batch] for i in mine]
    fmts = [base_fmt[i % args.batch] for i in mine]
    rois = [base_roi[i % args.batch] for i in mine]
    n_img = len(jpegs)

    dec = Decoder(local, options=f":fancy_upsampling={args.fancy}{args.options}")
    pinned_in = [torch.frombuffer(bytearray(j), dtype=torch.uint8).pin_memory() for j in jpegs]
    srcs = [t.numpy() for t in pinned_in]
    stream = torch.cuda.current_stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident: Prepare once (H2D), then time Execute only
    dev_batch = dec.make_batch(srcs, fmts, rois, to_host=False)
    # MPix counts OUTPUT pixels (ROI pixels where a ROI is requested)
    mpix_step = sum((r[2] - r[0]) * (r[3] - r[1]) if r else inf.width * inf.height for inf, r in zip(dev_batch.infos, rois)) / 1e6
    planes_bytes = 0
    for inf, f in zip(dev_batch.infos, fmts):
        ncomp = 1 if f in ("I_Y", "P_Y") else inf.num_components
        planes_bytes += sum(inf.mcus_x * inf.h[c] * 8 * inf.mcus_y * inf.v[c] * 8 for c in range(ncomp))
    dec.prepare(dev_batch)
    for _ in range(args.warmup):
        dec.execute()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        dec.execute()
    e1.record(stream)
    barrier()
    ms_dev = e0.elapsed_time(e1)
    launches_per_step = dec.stat(_lib.STAT_KERNEL_LAUNCHES)
    status = dec.finish(dev_batch)
    assert all(s == _lib.OK for s in status), "decode failed inside the bench"
    algo_bytes = dec.stat(_lib.STAT_ALGO_BYTES)
    comp_bytes = dec.stat(_lib.STAT_COMPRESSED_BYTES)
    n_coefs = dec.stat(_lib.STAT_NUM_COEFS)
    out_bytes = dec.stat(_lib.STAT_OUTPUT_BYTES)

    # ---- per-kernel times (CUDA events on the launch stream, same loop, profiling switched on)
    dec.set_profiling(True)
    dec.prepare(dev_batch)
    stage_acc = {}
    for it in range(args.warmup + args.steps):
        dec.execute()
        torch.cuda.synchronize()
        if it >= args.warmup:
            for k, v in dec.stage_times().items():
                stage_acc[k] = stage_acc.get(k, 0.0) + v
    dec.finish(dev_batch)
    dec.set_profiling(False)
    stages = {k: v / args.steps for k, v in stage_acc.items()}

    # ---- e2e_capi: host pinned bitstreams in, host pinned pixels out, through b200jpegDecodeBatch
    parity_checked = 0
    check_idx = sorted(set(int(i) for i in __import__("numpy").linspace(0, n_img - 1, 8)))

    def check_pixels(batch, where):
        import numpy as np

        n_ok = 0
        for i in check_idx:
            ref = reference_pixels(jpegs[i], "I_Y" if fmts[i] in ("I_Y", "P_Y") else "I_RGB", bool(args.fancy), rois[i])
            got = batch.images[i]
            got = got.numpy() if not got.is_cuda else got.cpu().numpy()
            assert np.array_equal(got, ref), f"{where}: image {i} differs from the reference decode"
            n_ok += 1
        return n_ok

    probe_batch = dec.make_batch(srcs, fmts, rois, to_host=True, out_storage=torch.empty(0, dtype=torch.uint8), size_only=True)
    host_storage, host_placement = pinned_near_gpu(probe_batch, torch.device("cuda", local))
    host_batch = dec.make_batch(srcs, fmts, rois, to_host=True, out_storage=host_storage)
    for _ in range(args.warmup):
        dec.run(host_batch)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st = dec.run(host_batch)
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    assert all(s == _lib.OK for s in st)
    parity_checked += check_pixels(host_batch, "e2e_capi")
    # the same call with the pixels left in HBM (what nvImageCodec's GPU backends hand to a GPU consumer)
    dev_batch.storage.zero_()
    for _ in range(2):
        dec.run(dev_batch)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        dec.run(dev_batch)
    torch.cuda.synchronize()
    t_e2e_dev = time.perf_counter() - t0
    parity_checked += check_pixels(dev_batch, "e2e_capi_device_out")
    clocks = sampler.stop()

    # ---- e2e: the plugin call (nvimgcodecDecoderDecode of the unmodified core + this extension), one driver process per rank
    opts = f":fancy_upsampling={args.fancy}{args.options}"
    plugin = {"host": None, "device": None}
    t_plugin = [0.0, 0.0]
    have_driver = PluginDriver.available()
    drv = PluginDriver(jpegs, local) if have_driver else None
    if drv:
        rep_host = max(3, min(args.steps, int(1.0 / max(t_e2e / args.steps, 1e-3)) + 1))
        rep_dev = max(5, min(10 * args.steps, int(1.0 / max(t_e2e_dev / args.steps, 1e-4)) + 1))
        barrier()
        drv_threads = max(2, cores // world) if world > 1 else 0  # the ranks share the host cores
        plugin["host"] = drv.run("host", opts, rep_host, fmts=fmts, rois=rois, crc_every=max(1, n_img // 8), threads=drv_threads)
        barrier()
        plugin["device"] = drv.run("device", opts, rep_dev, fmts=fmts, rois=rois, crc_every=max(1, n_img // 8), threads=drv_threads)
        barrier()
        if world == 1 and not args.no_extras:  # the same two calls with the bitstreams in pageable memory (staged by the extension)
            plugin["host_pageable"] = drv.run("host", opts, rep_host, fmts=fmts, rois=rois, pinned_in=False)
            plugin["device_pageable"] = drv.run("device", opts, rep_dev, fmts=fmts, rois=rois, pinned_in=False)
        import zlib

        import numpy as np

        for leg in ("host", "device"):
            res = plugin[leg]
            assert res and "error" not in res and res["ok"] == n_img, f"plugin leg failed: {res}"
            for i, crc in res["crc"].items():
                ref = reference_pixels(jpegs[i], "I_Y" if fmts[i] in ("I_Y", "P_Y") else "I_RGB", bool(args.fancy), rois[i])
                assert zlib.crc32(np.ascontiguousarray(ref).tobytes()) == crc, f"plugin {leg}: image {i} differs from the reference decode"
                parity_checked += 1
        t_plugin = [plugin["host"]["seconds"] / plugin["host"]["repeat"], plugin["device"]["seconds"] / plugin["device"]["repeat"]]

    # ---- strong scaling extra: one fixed heterogeneous batch, sharded over the ranks by compressed bytes
    strong = None
    if args.workload == "c2" and not args.no_extras:
        n_strong = 2048
        sj, sf, sr = extra_corpora["c5"]
        assert len(sj) == n_strong
        sizes = [len(j) for j in sj]
        parts = sharding.partition_by_bytes(sizes, world)
        mine_s = parts[rank]
        s_srcs = [torch.frombuffer(bytearray(sj[i]), dtype=torch.uint8).pin_memory() for i in mine_s]
        s_np = [t.numpy() for t in s_srcs]
        sb = dec.make_batch(s_np, [sf[i] for i in mine_s], [sr[i] for i in mine_s], to_host=False)
        s_mpix = sum((r[2] - r[0]) * (r[3] - r[1]) if r else inf.width * inf.height for inf, r in zip(sb.infos, [sr[i] for i in mine_s])) / 1e6
        for _ in range(2):
            st = dec.run(sb)
        assert all(x == _lib.OK for x in st)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            dec.run(sb)
        torch.cuda.synchronize()
        t_strong = time.perf_counter() - t0
        strong = [t_strong * 1e3, s_mpix, float(len(mine_s)), sharding.imbalance(sizes, parts)]
        del sb, s_srcs

    # ---- max over ranks, whole-job aggregate
    times = torch.tensor([ms_dev, t_e2e * 1e3, t_e2e_dev * 1e3, t_plugin[0] * 1e3, t_plugin[1] * 1e3, strong[0] if strong else 0.0],
                         dtype=torch.float64, device="cuda")
    total_mpix = torch.tensor([mpix_step, float(n_img), strong[1] if strong else 0.0, strong[2] if strong else 0.0, float(parity_checked)],
                              dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_reduce(total_mpix, op=dist.ReduceOp.SUM)
    ms_dev, ms_e2e, ms_e2e_dev, ms_plug_host, ms_plug_dev, ms_strong = [float(x) for x in times.tolist()]
    job_mpix, job_imgs, strong_mpix, strong_imgs, parity_total = [float(x) for x in total_mpix.tolist()]

    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
        # algorithmic bytes per launch of each kernel (DESIGN.md "Kernels"), for the N=1 shard of rank 0
        kbytes = {
            "scan(K0)": 2 * comp_bytes, "huff_sync(K1)": comp_bytes, "verify_scan(K2)": 16 * dec.stat(_lib.STAT_NUM_SUBSEQ),
            "huff_write(K3)": comp_bytes + 2 * n_coefs, "dc_scan(K3b)": 8 * (n_coefs // 64), "idct(K4a)": 2 * n_coefs + planes_bytes,
            "color(K4b)": planes_bytes + out_bytes, "pixels(K4)": 2 * n_coefs + out_bytes,
        }
        top = max(stages, key=stages.get) if stages else None
        stage_list = [{"kernel": k, "ms": round(v, 4), "share": round(v / max(sum(stages.values()), 1e-9), 4),
                       "algo_gbs": round(kbytes.get(k, 0) / (v * 1e-3) / 1e9, 1) if v > 0 else None} for k, v in stages.items()]
        roofline = None
        # DRAM bytes per launch of each kernel from the committed `ncu --set full` capture of this workload
        traffic = {}
        try:
            import glob

            newest = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_dram_traffic.json")))[-1]
            with open(newest) as f:
                traffic = {k.replace("_kernel", ""): v["dram_bytes"] for k, v in json.load(f)["kernels"].items()}
        except Exception:
            pass
        ncu_name = {"scan(K0)": "scan_scatter", "huff_sync(K1)": "huff_sync", "verify_scan(K2)": "verify_scan", "huff_write(K3)": "huff_emit" if "huff_emit" in traffic else "huff_write", "dc_scan(K3b)": "dc_apply",
                    "idct(K4a)": "idct", "color(K4b)": "color", "pixels(K4)": "pixels"}
        if top:
            ach = kbytes[top] / (stages[top] * 1e-3) / 1e9
            roofline = {"kernel": top, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                        "traffic": traffic.get(ncu_name.get(top, "")) if args.workload == "c2" and n_img == 256 else None,
                        "peak_source": peak_src,
                        "note": "dominant kernel by CUDA-event time; the entropy kernels are issue-bound, see profiles/ for warp-issue efficiency"}
        step_s = ms_dev / args.steps * 1e-3
        pipe = algo_bytes / step_s / 1e9
        # bounded CPU sample on rank 0 only: 64 of the images, 2 passes, all host threads
        cpu = None
        if world == 1:
            sample = jpegs[:64] if args.workload != "c5" else jpegs[:512]
            dt, how = cpu_reference_decode(sample, cores, 2)
            sample_mpix = sum(i.width * i.height for i in dev_batch.infos[:len(sample)]) / 1e6  # the CPU arm decodes whole images to RGB
            cpu = {"value": 2 * sample_mpix / dt, "unit": "MPix/s", "cores": cores, "kind": "reference",
                   "sample": f"{len(sample)} of the batch's images x 2 passes on {cores} threads; {how}"}
        e2e_capi = {"value": job_mpix * args.steps / (ms_e2e * 1e-3), "unit": "MPix/s", "images_per_s": job_imgs * args.steps / (ms_e2e * 1e-3),
                    "h2d_bytes_per_step": host_batch.h2d_bytes, "d2h_bytes_per_step": host_batch.d2h_bytes,
                    "call": "Decoder.run(batch) == b200jpegDecodeBatch, pinned host bitstreams in, pinned host pixels out",
                    "pinned_buffer_placement": host_placement}
        e2e_capi_dev = {"value": job_mpix * args.steps / (ms_e2e_dev * 1e-3), "unit": "MPix/s",
                        "h2d_bytes_per_step": dev_batch.h2d_bytes, "d2h_bytes_per_step": 16 * n_img,
                        "note": "same call, pixels left in HBM (what a GPU consumer of nvImageCodec sees); statuses read back"}
        if drv:
            e2e = {"value": job_mpix / (ms_plug_host * 1e-3), "unit": "MPix/s", "images_per_s": job_imgs / (ms_plug_host * 1e-3),
                   "h2d_bytes_per_step": host_batch.h2d_bytes, "d2h_bytes_per_step": host_batch.d2h_bytes,
                   "call": "nvimgcodecDecoderDecode + future wait + stream sync on the unmodified reference core (oracle/_ref/libnvimgcodec.so) "
                           "with this extension picked by priority; pinned host bitstreams in, pinned host pixels out; one driver "
                           "process (tests/host/host_driver.cpp) per rank, wall clock inside the driver, max over ranks",
                   "steps_timed": plugin["host"]["repeat"]}
            e2e_dev = {"value": job_mpix / (ms_plug_dev * 1e-3), "unit": "MPix/s", "images_per_s": job_imgs / (ms_plug_dev * 1e-3),
                       "h2d_bytes_per_step": dev_batch.h2d_bytes, "d2h_bytes_per_step": 0,
                       "note": "same plugin call, device output buffers", "steps_timed": plugin["device"]["repeat"]}
            for key, leg, tgt in (("pageable_in", "host_pageable", e2e), ("pageable_in", "device_pageable", e2e_dev)):
                r = plugin.get(leg)
                if r and "error" not in r:
                    tgt[key] = {"value": r["mpix_per_s"], "unit": "MPix/s",
                                "note": "same call with the bitstreams in pageable memory: the extension stages them into pinned memory on host threads"}
        else:
            e2e = dict(e2e_capi, note="oracle/_ref/host_driver is not built here: the plugin leg falls back to the C ABI")
            e2e_dev = e2e_capi_dev
        line = {
            "metric": "decoded_mpix_per_s", "value": job_mpix * args.steps / (ms_dev * 1e-3), "unit": "MPix/s",
            "images_per_s": job_imgs * args.steps / (ms_dev * 1e-3),
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload], "images_per_gpu": n_img, "upsampling": "fancy" if args.fancy else "box",
                       "l2": "working set per step (bitstreams + %.1f GB coefficients + pixels) is larger than the 126 MB L2, no flush needed"
                             % (2 * n_coefs / 1e9),
                       "sharding": "LPT by compressed bytes, one process per GPU, no collective", "host_affinity": numa},
            "e2e": e2e, "e2e_device_out": e2e_dev, "e2e_capi": e2e_capi, "e2e_capi_device_out": e2e_capi_dev,
            "parity_checked": int(parity_total),
            "gpu_launches": int(launches_per_step * args.steps),
            "roofline": roofline,
            "roofline_pipeline": {"bound": "hbm", "achieved": pipe, "peak": peak, "unit": "GB/s", "frac": pipe / peak,
                                  "algorithmic_bytes_per_step": algo_bytes, "formula": "C + 4*N_coef + P (SURVEY.md 8d)"},
            "stages": stage_list,
            "cpu_baseline": cpu,
            "clocks": clocks,
        }
        if strong:
            line["strong"] = {"workload": "2048 mixed images (C3 distribution + 10 % 1080p, 50 % ROI, 30 % gray output), pinned bitstreams in, pixels left in HBM, "
                                          "b200jpegDecodeBatch per rank", "images": int(strong_imgs), "value": strong_mpix * args.steps / (ms_strong * 1e-3),
                              "unit": "MPix/s", "images_per_s": strong_imgs * args.steps / (ms_strong * 1e-3), "ms_per_step": ms_strong / args.steps,
                              "byte_imbalance_max_over_mean": strong[3]}
        # ---- the reference's own CUDA decoder extension through the same framework call (SURVEY.md 8d "GPU comparison point")
        if drv and world == 1 and not args.no_extras:
            nvj = {"nvjpeg_hw_backend_status": nvjpeg_hw_backend_status(),
                   "note": "nvjpeg_hw_decoder registers only where nvjpegCreateEx(NVJPEG_BACKEND_HARDWARE) returns 0 (hw_decoder.cpp:282-292)"}
            if os.path.exists(os.path.join(NVJPEG_EXT_DIR, "libnvjpeg_ext.so")):
                for key, o in (("nvjpeg_cuda_decoder", opts), ("nvjpeg_cuda_decoder_gpu_huffman", opts + " nvjpeg_cuda_decoder:hybrid_huffman_threshold=0")):
                    r = drv.run("device", o, 3, ext_dir=NVJPEG_EXT_DIR, backend="hybrid", fmts=fmts, rois=rois)
                    nvj[key] = r if "error" in r else {"mpix_per_s": r["mpix_per_s"], "images_per_s": r["images_per_s"], "ok": r["ok"],
                                                       "call": "same nvimgcodecDecoderDecode, pageable bitstreams in, device pixels out"}
                if nvj["nvjpeg_hw_backend_status"] == 0:
                    r = drv.run("device", opts, 3, ext_dir=NVJPEG_EXT_DIR, backend="hw", fmts=fmts, rois=rois)
                    nvj["nvjpeg_hw_decoder"] = r if "error" in r else {"mpix_per_s": r["mpix_per_s"], "images_per_s": r["images_per_s"], "ok": r["ok"]}
            else:
                nvj["unavailable"] = "oracle/_ref/extensions_nvjpeg/libnvjpeg_ext.so not built"
            line["nvjpeg_ext"] = nvj
        # ---- the other BASELINE configs, bounded (N=1 only): kernels-only and C-ABI pixels-in-HBM numbers, statuses checked
        if world == 1 and args.workload == "c2" and not args.no_extras:
            others = []
            wl_names = dict(WORKLOADS, progressive="n3: 256 x 1920x1080 PROGRESSIVE (SOF2, libjpeg's ten-scan script) 4:2:0 q90 -> I_RGB")
            for name, nb in (("c3", 256), ("c4", 32), ("c5", 2048), ("progressive", 256)):
                try:
                    oj, of, orr = extra_corpora[name]
                    assert len(oj) == nb
                    ob = dec.make_batch([torch.frombuffer(bytearray(j), dtype=torch.uint8).pin_memory().numpy() for j in oj], of, orr, to_host=False)
                    ompix = sum((r[2] - r[0]) * (r[3] - r[1]) if r else inf.width * inf.height for inf, r in zip(ob.infos, orr)) / 1e6
                    dec.prepare(ob)
                    for _ in range(3):
                        dec.execute()
                    torch.cuda.synchronize()
                    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    f0.record(stream)
                    for _ in range(5):
                        dec.execute()
                    f1.record(stream)
                    torch.cuda.synchronize()
                    ost = dec.finish(ob)
                    kms = f0.elapsed_time(f1) / 5
                    for _ in range(2):
                        dec.run(ob)
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    for _ in range(5):
                        dec.run(ob)
                    torch.cuda.synchronize()
                    pms = (time.perf_counter() - t0) / 5 * 1e3
                    others.append({"workload": wl_names[name], "images": nb, "all_ok": all(x == _lib.OK for x in ost),
                                   "kernels_only_mpix_per_s": ompix / (kms * 1e-3), "ms_per_step": kms,
                                   "e2e_capi_device_out_mpix_per_s": ompix / (pms * 1e-3)})
                    if name == "progressive":  # the reference's CPU path on the same streams: 32 of them, all host threads
                        dt, how = cpu_reference_decode(oj[:32], cores, 1)
                        others[-1]["cpu_reference_mpix_per_s"] = 32 * HEIGHT * WIDTH / 1e6 / dt
                        others[-1]["cpu_reference"] = f"32 of the streams x 1 pass on {cores} threads; {how}"
                        if drv:  # the same streams through the framework call, and the reference's own CUDA decoder beside it
                            pdrv = PluginDriver(oj, local)
                            try:
                                r = pdrv.run("device", opts, 3)
                                others[-1]["plugin_device_out_mpix_per_s"] = r.get("mpix_per_s", r.get("error"))
                                others[-1]["plugin_ok"] = r.get("ok")
                                if os.path.exists(os.path.join(NVJPEG_EXT_DIR, "libnvjpeg_ext.so")):
                                    r = pdrv.run("device", opts, 2, ext_dir=NVJPEG_EXT_DIR, backend="hybrid")
                                    others[-1]["nvjpeg_cuda_decoder_mpix_per_s"] = r.get("mpix_per_s", r.get("error"))
                            finally:
                                pdrv.close()
                    del ob
                except Exception as e:  # an extra must not take the headline down
                    others.append({"workload": wl_names[name], "error": repr(e)[:200]})
            line["other_configs"] = others
        sys.stdout.flush()
        os.dup2(stdout_fd, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    if drv:
        drv.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
