import sys, torch
sys.path.insert(0, "/root/repo")
from video_depth_anything_b200 import lib as L
lib = L.load(); dev = torch.device("cuda:0")
frames, ntok, heads = 32, 1370, 16
qkv = (torch.randn(frames * ntok, 3 * heads * 64, device=dev) * 1.3).bfloat16()
o = torch.empty(frames * ntok, heads * 64, device=dev, dtype=torch.bfloat16)
f = lambda: L.check(lib.vda_attention(qkv.data_ptr(), o.data_ptr(), L.BF16, frames, ntok, heads, 1, 1, L.stream_ptr()), "attn")
for _ in range(3): f()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): f()
e1.record(); torch.cuda.synchronize()
print("attention %.1f us" % (e0.elapsed_time(e1) * 50))
