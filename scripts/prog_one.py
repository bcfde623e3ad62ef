"""GPU box: one progressive 1080p image with B200JPEG_TRACE=1 -- prints how long every chain (scan) took."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200JPEG_TRACE"] = "1"
from nvimagecodec_b200 import Decoder, corpus  # noqa: E402

d = Decoder(0, options=":fancy_upsampling=1")
q = int(sys.argv[1]) if len(sys.argv) > 1 else 90
s = corpus.encode(corpus.synth(1080, 1920, 0), q, "420", progressive=True)
print("quality", q, "bytes", len(s), file=sys.stderr)
d.decode([s], "I_RGB")
d.decode([s], "I_RGB")
