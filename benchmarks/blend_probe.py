import os, sys, torch, numpy as np
sys.path.insert(0, '/root/repo')
from bioengine_worker_b200 import kernels as K, geometry as G
dev = torch.device('cuda:0')
H = W = 4096; B = 256; ov = 64
axes = [K.DeviceAxis(a, dev) for a in (G.identity_axis(1), G.fibsem_axis(H, B, ov), G.fibsem_axis(W, B, ov))]
T = axes[1].ax.T * axes[2].ax.T
kind = sys.argv[1] if len(sys.argv) > 1 else 'randn'
def make():
    x = torch.randn((T, 1, B, B), device=dev)
    if kind == 'far': x = x * 0.1 - 3
    if kind == 'smooth': x = torch.nn.functional.avg_pool2d(x, 9, 1, 4) * 6 - 0.8
    if kind == 'big': x = x * 50
    return x.contiguous()
logits = make()
w = torch.from_numpy(G.gauss_window_1d(B)).to(dev); one = torch.ones(1, dtype=torch.float64, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
logits2 = make()
table = K.blend_window((one, w, w))
from bioengine_worker_b200 import _lib
bits = torch.empty((H, W // 64), dtype=torch.int64, device=dev)
lib = _lib.load()
stream = torch.cuda.current_stream().cuda_stream
az, ay, ax = axes
def run(i=0):
    x = logits if i & 1 else logits2
    rc = lib.be_stitch_blend_f32(x.data_ptr(), None, None, bits.data_ptr(), 1, 1, 1, H, W, 1, ay.ax.T, ax.ax.T, 1, B, B, 1,
                                 ay.ax.stride, ax.ax.stride, one.data_ptr(), w.data_ptr(), w.data_ptr(), table.data_ptr(), 1, 0.5, stream)
    assert rc == 0
for i in range(400): run(i)   # warm-up (clocks)
torch.cuda.synchronize()
n = 200
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for i in range(n): run(i)
b.record(); torch.cuda.synchronize()
print(kind, os.environ.get('BE_MASK2D_BLOCKS', '-'), 'blend us per call (two 127 MB inputs alternating)', round(1e3 * a.elapsed_time(b) / n, 1))
