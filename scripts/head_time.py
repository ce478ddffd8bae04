import sys, torch
sys.path.insert(0, "/root/repo")
from video_depth_anything_b200 import lib as L
lib = L.load(); dev = torch.device("cuda:0")
frames, ih, iw, C, oh, ow = 32, 296, 296, 128, 518, 518
x = torch.randn(frames*ih*iw, C, device=dev).bfloat16()
w = (torch.randn(32, 9*C, device=dev)*0.05)
w16 = w.half()
b = torch.randn(32, device=dev); hw = torch.rand(32, device=dev)
depth = torch.empty(frames, oh, ow, device=dev)
f = lambda: L.check(lib.vda_output_head_f16(x.data_ptr(), w16.data_ptr(), b.data_ptr(), hw.data_ptr(), 0.05, depth.data_ptr(), frames, ih, iw, C, oh, ow, L.stream_ptr()), "head")
for _ in range(3): f()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): f()
e1.record(); torch.cuda.synchronize()
print("output_head_f16 %.1f us" % (e0.elapsed_time(e1) * 100))
