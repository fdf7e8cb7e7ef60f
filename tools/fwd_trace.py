#!/usr/bin/env python3
"""Timeline of the forward kernel's accumulator hand-offs (CTA 0), from a -DB200DS_TRACE build of the library."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "distributed-gpu-test_b200"))
import torch
from b200ds import _lib
lib = _lib.load()
B, F, H = 65536, 20, int(os.environ.get("H", 4096))
dev = torch.device("cuda", 0)
P = F * H + 2 * H + 1; Ppad = (P + 3) // 4 * 4
params = torch.randn(Ppad, device=dev) * 0.1
x = torch.randn(B, F, device=dev)
xq = torch.empty(lib.b200ds_tc_xq_bytes(B), dtype=torch.uint8, device=dev)
wq = torch.empty(lib.b200ds_tc_wq_bytes(H), dtype=torch.uint8, device=dev)
z = torch.empty(B, device=dev)
st = torch.cuda.current_stream().cuda_stream
_lib.check(lib.b200ds_pack_x_bf16(x.data_ptr(), B, F, xq.data_ptr(), st))
_lib.check(lib.b200ds_pack_params_bf16(params.data_ptr(), F, H, wq.data_ptr(), st))
fsc = torch.zeros(max(256, lib.b200ds_mlp_forward_tc_scratch_bytes(B, F, H)), dtype=torch.uint8, device=dev)
for _ in range(3):
    _lib.check(lib.b200ds_mlp_forward_tc(xq.data_ptr(), wq.data_ptr(), params.data_ptr(), B, F, H, z.data_ptr(), fsc.data_ptr(), fsc.numel(), st))
torch.cuda.synchronize()
n = 16 * 128 * 4
buf = (C.c_longlong * n)()
lib.b200ds_debug_trace.argtypes = [C.c_void_p, C.c_size_t]
_lib.check(lib.b200ds_debug_trace(buf, n))
t = [[[buf[((r * 128) + q) * 4 + e] for e in range(4)] for q in range(128)] for r in range(6)]
t0 = min(v for r in t for q in r for v in q if v > 0)
print("chunk | issuer: wait_begin woke issued | epilogue team q%3 (last of its 4 warps to stamp): wait_begin full released ffma_done")
for q in range(40, 80):
    iss = t[q & 1][q]
    e = t[2 + q % 3][q]
    print(f"{q:4d} | {iss[0]-t0:7d} {iss[1]-t0:7d} {iss[2]-t0:7d} | team {q % 3} {e[0]-t0:7d} {e[1]-t0:7d} {e[2]-t0:7d} {e[3]-t0:7d}")
