#!/usr/bin/env python3
"""Timeline of the accumulator hand-offs inside the two tensor-core kernels (CTA 0), from a -DB200DS_TRACE build:

    tools/build_variant.sh trace "-DB200DS_TRACE"
    B200DS_LIB=$PWD/tools/bin/lib_trace.so python tools/tc_trace.py [fwd|bwd]

forward (one handshake per group of two chunks):  issuer: wait_begin, woke, issued | epilogue warps 0 (columns 0-63) and 4
(columns 64-127) of lane quadrant 0: group start, released, next group seen full, arithmetic done.
backward (per chunk): MMA1 issuer | MMA2 issuer | epilogue warp of quadrant 0 of the chunk's team: wait_begin, full, converted, A2 ready."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
import torch
from b200ds import _lib
lib = _lib.load()
which = sys.argv[1] if len(sys.argv) > 1 else "fwd"
B, F, H = 65536, 20, int(os.environ.get("H", 4096))
dev = torch.device("cuda", 0)
P = F * H + 2 * H + 1; Ppad = (P + 3) // 4 * 4
params = torch.randn(Ppad, device=dev) * 0.1
x = torch.randn(B, F, device=dev)
xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=dev)
wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=dev)
z = torch.empty(B, device=dev); g = torch.empty(Ppad, device=dev)
dz = torch.randn(B, device=dev) * 1e-5
st = torch.cuda.current_stream().cuda_stream
_lib.check(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq.data_ptr(), st))
_lib.check(lib.b200ds_pack_params_bf16(params.data_ptr(), F, H, wq.data_ptr(), st))
fsc = torch.zeros(max(256, lib.b200ds_mlp_forward_tc_scratch_bytes(B, F, H)), dtype=torch.uint8, device=dev)
sc = torch.zeros(lib.b200ds_mlp_backward_tc_scratch_bytes(B, F, H), dtype=torch.uint8, device=dev)
lib.b200ds_debug_trace.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
n = 24 * 128 * 4
buf = (C.c_longlong * n)()
for _ in range(3):
    if which == "fwd":
        _lib.check(lib.b200ds_mlp_forward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, z.data_ptr(), fsc.data_ptr(), fsc.numel(), st))
    else:
        _lib.check(lib.b200ds_mlp_backward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), dz.data_ptr(), B, F, H, g.data_ptr(), sc.data_ptr(), sc.numel(), st))
    _lib.check(lib.b200ds_debug_trace(None, 0, 1))
if which == "fwd":
    _lib.check(lib.b200ds_mlp_forward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, z.data_ptr(), fsc.data_ptr(), fsc.numel(), st))
else:
    _lib.check(lib.b200ds_mlp_backward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), dz.data_ptr(), B, F, H, g.data_ptr(), sc.data_ptr(), sc.numel(), st))
_lib.check(lib.b200ds_debug_trace(buf, n, 0))
t = [[[buf[((r * 128) + q) * 4 + e] for e in range(4)] for q in range(128)] for r in range(24)]
t0 = min(v for r in t for q in r for v in q if v > 0)
rel = lambda v: (v - t0) if v > 0 else -1      # noqa: E731
if which == "fwd":
    print("group | issuer: wait_begin woke issued | warp 0: start released next_full done | warp 4: start released next_full done")
    for q in range(16, 48):
        iss = t[q & 1][q]
        e0, e4 = t[2][q], t[6][q]
        print(f"{q:4d} | {rel(iss[0]):7d} {rel(iss[1]):7d} {rel(iss[2]):7d} | {rel(e0[0]):7d} {rel(e0[1]):7d} {rel(e0[2]):7d} {rel(e0[3]):7d} | {rel(e4[0]):7d} {rel(e4[1]):7d} {rel(e4[2]):7d} {rel(e4[3]):7d}")
else:
    print("chunk | MMA1: wait_begin woke issued | MMA2: wait_begin woke issued | epilogue quad 0: wait_begin full converted ready")
    for q in range(30, 70):
        a, b_, e = t[10][q], t[11][q], t[12][q]
        print(f"{q:4d} | {rel(a[0]):7d} {rel(a[1]):7d} {rel(a[2]):7d} | {rel(b_[0]):7d} {rel(b_[1]):7d} {rel(b_[2]):7d} | {rel(e[0]):7d} {rel(e[1]):7d} {rel(e[2]):7d} {rel(e[3]):7d}")
