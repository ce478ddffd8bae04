"""Device time of the tcgen05 flash attention on the encoder's shape (vitl: 32 frames x 1370 tokens x 16 heads).
VDA_LIB=/path/to/variant.so times an experimental build of the library."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from video_depth_anything_b200 import lib as L  # noqa: E402

if os.environ.get("VDA_LIB"):
    L.LIB_PATH = os.environ["VDA_LIB"]
lib = L.load()
dev = torch.device("cuda:0")
frames, heads = 32, 16
ntok = int(sys.argv[1]) if len(sys.argv) > 1 else 1370   # 2443 = the 518x924 network input of configs[3]
qkv = (torch.randn(frames * ntok, 3 * heads * 64, device=dev) * 1.3).bfloat16()
o = torch.empty(frames * ntok, heads * 64, device=dev, dtype=torch.bfloat16)
f = lambda: L.check(lib.vda_attention(qkv.data_ptr(), o.data_ptr(), L.BF16, frames, ntok, heads, 1, 1, L.stream_ptr()), "attn")  # noqa: E731
for _ in range(3):
    f()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = int(os.environ.get("ATTN_ITERS", "20"))
e0.record()
for _ in range(n):
    f()
e1.record()
if n >= 1000:   # long enough to sample the SM clock and board power under sustained load
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(0)
    cs, ps = [], []
    while not e1.query():
        cs.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
        ps.append(pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0)
    cs, ps = sorted(cs[len(cs) // 3:]), sorted(ps[len(ps) // 3:])
    print("sustained: sm %d MHz, %.0f W" % (cs[len(cs) // 2], ps[len(ps) // 2]))
torch.cuda.synchronize()
print("attention ntok=%d %.1f us  (%s)" % (ntok, e0.elapsed_time(e1) * 1e3 / n, os.environ.get("VDA_LIB", "in-tree library")))
