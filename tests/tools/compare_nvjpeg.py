"""Same framework, same call (nvimgcodecDecoderDecode on host bitstreams -> device P_RGB/I_RGB buffers), different JPEG
decoder extension: ours vs the reference's nvjpeg_ext over the toolkit's closed nvJPEG.  Run on the GPU box."""
import os, subprocess, sys, tempfile, shutil
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
jp = bench.make_corpus(n, os.cpu_count())
tmp = tempfile.mkdtemp()
lines = []
for i, d in enumerate(jp):
    p = os.path.join(tmp, f"{i}.jpg")
    open(p, "wb").write(d)
    lines.append(f"{p} I_RGB 0 0 0 0 device")
man = os.path.join(tmp, "manifest.txt")
open(man, "w").write("\n".join(lines) + "\n")
drv = os.path.join(ROOT, "oracle", "_ref", "host_driver")
env = dict(os.environ)
env["LD_LIBRARY_PATH"] = os.path.dirname(drv) + ":/usr/local/cuda/lib64:" + env.get("LD_LIBRARY_PATH", "")
ours = os.path.join(tmp, "ext_ours"); os.makedirs(ours)
shutil.copy(os.path.join(ROOT, "nvimagecodec_b200", "lib", "libb200jpeg_ext.so"), ours)
nvj = os.path.join(ROOT, "oracle", "_ref", "extensions_nvjpeg")
runs = [("b200_jpeg_decoder", ours, ":fancy_upsampling=1", "any"),
        ("nvjpeg_cuda_decoder (default: CPU Huffman <= 1 MPix)", nvj, ":fancy_upsampling=1", "hybrid"),
        ("nvjpeg_cuda_decoder hybrid_huffman_threshold=0 (GPU Huffman)", nvj, ":fancy_upsampling=1 nvjpeg_cuda_decoder:hybrid_huffman_threshold=0", "hybrid"),
        ("nvjpeg_hw_decoder", nvj, ":fancy_upsampling=1", "hw")]
for name, ext, opts, backend in runs:
    out = os.path.join(tmp, "out_" + str(abs(hash(name)) % 10000)); os.makedirs(out)
    r = subprocess.run([drv, ext, man, out, opts, "5", backend], capture_output=True, text=True, env=env, timeout=900)
    line = [l for l in r.stdout.splitlines() if l.startswith("BENCH")]
    print(name, "->", line[0] if line else ("FAILED rc=%d " % r.returncode + (r.stdout + r.stderr)[-300:].replace("\n", " | ")))
