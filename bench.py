#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric (decoded MPix/s, images/s) on its config #2:
batch 256 x 1920x1080 baseline 4:2:0 q90 synthetic JPEG -> interleaved RGB, per B200.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # libjpeg-turbo on the host cores

A "step" is one pass of the hot path over one batch.  `value` times the kernels only
(bitstreams already resident in HBM, outputs left in HBM); `e2e` times the public call with
HOST buffers: pinned bitstreams in, pinned pixels out, H2D and D2H inside the timed region.
Weak scaling: every rank decodes its own 256-image shard of a 256*N batch that is partitioned
by compressed byte count (nvimagecodec_b200/sharding.py); no data-path collective.
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
    n = min(PER_GPU, 128)
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
        "config": {"workload": WORKLOAD, "sample": f"{n} images per step"},
        "cpu_baseline": {"value": val, "unit": "MPix/s", "cores": cores, "kind": "reference", "sample": f"{n} of the 256 images per step; {how}"},
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
    assert torch.cuda.is_available(), "bench.py needs a GPU (the decoder has no CPU path)"
    torch.cuda.set_device(local)
    numa = "unset"
    if world > 1:
        # keep this rank's host threads (and, by first touch, its pinned staging) on the CPUs next to its GPU
        try:
            import pynvml

            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
            numa = "nvml cpu affinity, %d cpus" % len(os.sched_getaffinity(0))
        except Exception as e:  # affinity is an optimisation only
            numa = "unavailable (%s)" % type(e).__name__
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

    # ---- e2e: host pinned bitstreams in, host pinned pixels out, through the public call
    host_batch = dec.make_batch(srcs, fmts, rois, to_host=True)
    for _ in range(args.warmup):
        dec.run(host_batch)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st = dec.run(host_batch)
    torch.cuda.synchronize()
    t_e2e = time.perf_counter() - t0
    assert all(s == _lib.OK for s in st)
    # e2e with the pixels left in HBM (what nvImageCodec's GPU backends hand to a GPU consumer)
    for _ in range(2):
        dec.run(dev_batch)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        dec.run(dev_batch)
    torch.cuda.synchronize()
    t_e2e_dev = time.perf_counter() - t0
    clocks = sampler.stop()

    # ---- max over ranks, whole-job aggregate
    times = torch.tensor([ms_dev, t_e2e * 1e3, t_e2e_dev * 1e3], dtype=torch.float64, device="cuda")
    total_mpix = torch.tensor([mpix_step, float(n_img)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_reduce(total_mpix, op=dist.ReduceOp.SUM)
    ms_dev, ms_e2e, ms_e2e_dev = [float(x) for x in times.tolist()]
    job_mpix, job_imgs = [float(x) for x in total_mpix.tolist()]

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
            "color(K4b)": planes_bytes + out_bytes,
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
        ncu_name = {"scan(K0)": "scan_scatter", "huff_sync(K1)": "huff_sync", "verify_scan(K2)": "verify_scan", "huff_write(K3)": "huff_write", "dc_scan(K3b)": "dc_apply",
                    "idct(K4a)": "idct", "color(K4b)": "color"}
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
        line = {
            "metric": "decoded_mpix_per_s", "value": job_mpix * args.steps / (ms_dev * 1e-3), "unit": "MPix/s",
            "images_per_s": job_imgs * args.steps / (ms_dev * 1e-3),
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_dev / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOADS[args.workload], "images_per_gpu": n_img, "upsampling": "fancy" if args.fancy else "box",
                       "l2": "working set per step (bitstreams + %.1f GB coefficients + planes + pixels) is larger than the 126 MB L2, no flush needed"
                             % (2 * n_coefs / 1e9),
                       "sharding": "LPT by compressed bytes, one process per GPU, no collective", "host_affinity": numa},
            "e2e": {"value": job_mpix * args.steps / (ms_e2e * 1e-3), "unit": "MPix/s", "images_per_s": job_imgs * args.steps / (ms_e2e * 1e-3),
                    "h2d_bytes_per_step": host_batch.h2d_bytes, "d2h_bytes_per_step": host_batch.d2h_bytes,
                    "call": "Decoder.run(batch) == b200jpegDecodeBatch, pinned host in, pinned host out"},
            "e2e_device_out": {"value": job_mpix * args.steps / (ms_e2e_dev * 1e-3), "unit": "MPix/s",
                               "h2d_bytes_per_step": dev_batch.h2d_bytes, "d2h_bytes_per_step": 16 * n_img,
                               "note": "same call, pixels left in HBM (what a GPU consumer of nvImageCodec sees); statuses read back"},
            "gpu_launches": int(launches_per_step * args.steps),
            "roofline": roofline,
            "roofline_pipeline": {"bound": "hbm", "achieved": pipe, "peak": peak, "unit": "GB/s", "frac": pipe / peak,
                                  "algorithmic_bytes_per_step": algo_bytes, "formula": "C + 4*N_coef + P (SURVEY.md 8d)"},
            "stages": stage_list,
            "cpu_baseline": cpu,
            "clocks": clocks,
        }
        sys.stdout.flush()
        os.dup2(stdout_fd, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
