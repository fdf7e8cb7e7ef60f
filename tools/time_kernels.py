#!/usr/bin/env python3
"""Time the tensor-core forward/backward kernels alone (CUDA events, after warm-up)."""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
import torch
from b200ds import _lib
lib = _lib.load()
B, F, H = int(os.environ.get("B", 65536)), 20, int(os.environ.get("H", 4096))
dev = torch.device("cuda", 0)
torch.manual_seed(0)
P = F * H + 2 * H + 1; Ppad = (P + 3) // 4 * 4
params = torch.randn(Ppad, device=dev) * 0.1
x = torch.randn(B, F, device=dev)
dz = torch.randn(B, device=dev) * 1e-5
xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=dev)
wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=dev)
z = torch.empty(B, device=dev); g = torch.empty(Ppad, device=dev)
sc = torch.zeros(lib.b200ds_mlp_backward_tc_scratch_bytes(B, F, H), dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream
_lib.check(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq.data_ptr(), st))
_lib.check(lib.b200ds_pack_params_bf16(params.data_ptr(), F, H, wq.data_ptr(), st))
fsc = torch.zeros(max(256, lib.b200ds_mlp_forward_tc_scratch_bytes(B, F, H)), dtype=torch.uint8, device=dev)
def fwd(): _lib.check(lib.b200ds_mlp_forward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, z.data_ptr(), fsc.data_ptr(), fsc.numel(), st))
def bwd(): _lib.check(lib.b200ds_mlp_backward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), dz.data_ptr(), B, F, H, g.data_ptr(), sc.data_ptr(), sc.numel(), st))
yl = (torch.rand(B, device=dev) > 0.5).float()
ls = torch.zeros(1, device=dev)
def lbwd(): _lib.check(lib.b200ds_mlp_loss_backward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), z.data_ptr(), yl.data_ptr(), B, F, H, ls.data_ptr(), g.data_ptr(), None, sc.data_ptr(), sc.numel(), st))
yb = torch.empty(B, device=dev)
def gen(): _lib.check(lib.b200ds_gen_batch_bf16(xq.data_ptr(), yb.data_ptr(), B, F, 1234, 0, None, st))
ctx = C.c_void_p()
_lib.check(lib.b200ds_ctx_create(0, 0, 1, None, C.byref(ctx)))
m = torch.zeros(Ppad, device=dev); v = torch.zeros(Ppad, device=dev)
ctr = torch.ones(1, dtype=torch.int64, device=dev)
def lbwd_partials(): _lib.check(lib.b200ds_mlp_loss_backward_partials_tc(xq.data_ptr(), wq.data_ptr(), z.data_ptr(), yl.data_ptr(), B, F, H, None, sc.data_ptr(), sc.numel(), st))
def update(): _lib.check(lib.b200ds_update_tc(ctx, sc.data_ptr(), sc.numel(), B, F, H, params.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), wq.data_ptr(), 1e-3, 0.9, 0.999, 1e-8, 0, ctr.data_ptr(), ls.data_ptr(), st))
def adam(): _lib.check(lib.b200ds_adam_step_counted(params.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), Ppad, 1e-3, 0.9, 0.999, 1e-8, ctr.data_ptr(), 1.0, 0, st))
def pack(): _lib.check(lib.b200ds_pack_params_bf16(params.data_ptr(), F, H, wq.data_ptr(), st))
# spin the clocks up
t0 = time.time()
while time.time() - t0 < 1.0:
    fwd(); (lbwd if os.environ.get("SPIN") == "lbwd" else bwd)()
torch.cuda.synchronize()
# each kernel N times inside ONE CUDA graph: device time per launch without the host's launch cost
# (back-to-back eager launches cannot resolve anything shorter than ~10 us of host work per call)
N = 20
for name, fn in (("gen_batch", gen), ("fwd", fwd), ("bwd+finish", bwd), ("loss+bwd+finish", lbwd), ("loss+bwd (partials)", lbwd_partials),
                 ("update (finish+adam+pack)", update), ("adam", adam), ("pack_w1", pack)):
    fn(); torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    cap = torch.cuda.Stream()
    cap.wait_stream(torch.cuda.current_stream())
    with torch.cuda.graph(gr, stream=cap):
        st_save = st
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(N): fn()
        st = st_save
    torch.cuda.current_stream().wait_stream(cap)
    for _ in range(3): gr.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): gr.replay()
    e1.record(); torch.cuda.synchronize()
    print(f"{os.environ.get('B200DS_LIB','default'):28s} B={B} H={H} {name:28s} {1e3 * e0.elapsed_time(e1) / (5 * N):8.2f} us", flush=True)
