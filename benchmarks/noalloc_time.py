"""How long be_quantile_f32_noalloc (three-pass radix select, library-resident scratch) takes next to be_quantile_f32."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bioengine_worker_b200 import _lib
dev = torch.device("cuda:0")
q = torch.tensor([0.01, 0.998], dtype=torch.float64, device=dev)
for side in (2048, 4096):
    x = (torch.rand((side, side), device=dev) * 1000).contiguous()
    quant = torch.empty((1, 2), dtype=torch.float64, device=dev)
    norm = torch.empty((1, 2), dtype=torch.float64, device=dev)
    ws = torch.empty(_lib.query("be_quantile_workspace_bytes", 1, x.numel()), dtype=torch.uint8, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    for name in ("be_quantile_f32", "be_quantile_f32_noalloc"):
        args = [x.data_ptr(), 1, x.numel(), q.data_ptr(), 2, 1e-6, quant.data_ptr(), norm.data_ptr()]
        if name == "be_quantile_f32":
            args.append(ws.data_ptr())
        for _ in range(20):
            _lib.call(name, *args, s)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(50):
            _lib.call(name, *args, s)
        b.record(); torch.cuda.synchronize()
        print(f"{side}^2 {name}: {a.elapsed_time(b) / 50 * 1e3:.1f} us per call (back to back, input L2-warm for 2048^2)")
