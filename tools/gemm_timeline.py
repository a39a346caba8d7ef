#!/usr/bin/env python3
"""Development aid: per-k-block clock64 timeline of the GEMM pipeline (CTA (0,0)): when the
producer found the stage free, when the data had landed, when the MMAs had been issued."""
import ctypes, os, subprocess, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from distributed_training_pytorch_b200 import build
out = os.path.join(ROOT, "gpurun_out", "libddp_b200_gemmtl.so")
os.makedirs(os.path.dirname(out), exist_ok=True)
subprocess.run(["nvcc"] + build.NVCC_FLAGS + ["-DDDPB_GEMM_TIMELINE"] + os.environ.get("DIAG", "").split() + ["-o", out] + build.sources() + ["-ldl"], check=True)
L = ctypes.CDLL(out)
B, H = 256, int(os.environ.get("H", 4096))
x = torch.randn(B, H, device="cuda"); W = torch.randn(H, H, device="cuda"); b = torch.randn(H, device="cuda")
o = torch.empty(B, H, device="cuda"); ot = torch.empty(H, 256, device="cuda")
vp = ctypes.c_void_p
L.ddpb_linear_fwd.argtypes = [vp] * 5 + [ctypes.c_int] * 5 + [vp]
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    assert L.ddpb_linear_fwd(x.data_ptr(), W.data_ptr(), b.data_ptr(), o.data_ptr(), ot.data_ptr(), B, H, H, 256, 1, st) == 0
torch.cuda.synchronize()
buf = np.zeros((3, 512), np.int64)
L.ddpb_gemm_stamps.argtypes = [vp]
assert L.ddpb_gemm_stamps(buf.ctypes.data) == 0
nk = H // 32
t0 = buf[0, 0]
rel = buf[:, :nk] - t0
print("k-block: producer-issue, data-landed, mma-issued (cycles from first issue)")
for kb in list(range(0, 20)) + list(range(nk - 6, nk)):
    print(kb, rel[0, kb], rel[1, kb], rel[2, kb], " landed-issue latency:", rel[1, kb] - rel[0, kb])
d = np.diff(rel[1, 16:nk])
print("steady-state cycles per k-block (data-landed to data-landed): mean %.0f min %d max %d" % (d.mean(), d.min(), d.max()))
print("mean issue->landed latency: %.0f cycles" % (rel[1, 16:nk] - rel[0, 16:nk]).mean())

# Per-CTA wall-clock phases (globaltimer ns) for a short-K and the full-K launch.
L.ddpb_gemm_phases.argtypes = [vp]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for K in (128, 1024, H):
    xs = x[:, :K].contiguous(); Ws = W[:, :K].contiguous()
    for _ in range(2):
        flush.fill_(1)
        assert L.ddpb_linear_fwd(xs.data_ptr(), Ws.data_ptr(), b.data_ptr(), o.data_ptr(), ot.data_ptr(), B, K, H, 256, 1, st) == 0
    torch.cuda.synchronize()
    ph = np.zeros((1024, 8), np.int64)
    assert L.ddpb_gemm_phases(ph.ctypes.data) == 0
    ph = ph[ph[:, 0] > 0]
    t0 = ph[:, 0].min()
    r = ph[:, :5] - t0
    print("K=%d: %d CTAs; ns from first CTA entry; per phase (min / median / max):" % (K, len(ph)))
    for i, name in enumerate(["entry", "set-up done", "accumulator ready", "epilogue done", "exit"]):
        c = r[:, i][r[:, i] > -10**9]
        print("   %-18s %7d %7d %7d" % (name, c.min(), np.median(c), c.max()))
