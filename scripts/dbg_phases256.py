"""Timing experiment (GPU box): per-phase clock64 stamps of one tile of tc_edge_fwd256 from a -DEGNN_DBG_TIMING build.
    EGNN_NVCC_EXTRA=-DEGNN_DBG_TIMING python -m diffusion_model_b200.build --force && python scripts/dbg_phases256.py"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from diffusion_model_b200.EquivariantGraphNeuralNetwork import EGCL  # noqa: E402
from diffusion_model_b200 import _lib  # noqa: E402

lib = _lib.load()
torch.manual_seed(0)
d, H, n, deg = 256, 3, 4096, 128
layer = EGCL(2 * H + 1, d, d, 2 * H + 1, d, 1, H + d, d, H).to("cuda")
g = torch.Generator().manual_seed(1)
recv = torch.arange(n).repeat_interleave(deg)
send = torch.randint(0, n, (n * deg,), generator=g)
ei = torch.stack([recv, send]).cuda()
h, x = torch.randn(n, H, generator=g).cuda(), torch.randn(n, 3, generator=g).cuda()
names = {0: "tile start", 1: "GEMM1 observed", 2: "warp0 epilogue-1 done", 3: "barrier passed", 8: "issuer starts",
         9: "GEMM2 issued+committed", 4: "GEMM2 observed", 5: "tile end"}
for prec in ("bf16x3", "bf16"):
    layer.precision = prec
    with torch.no_grad():
        for _ in range(3):
            layer(ei, h, x)
    torch.cuda.synchronize()
    buf = (C.c_longlong * 32)()
    assert lib.egnn_dbg_read_stamps(buf) == 0
    for cta in (0, 1):
        st = {k: buf[cta * 16 + k] for k in names}
        t0 = st[0]
        print(prec, "CTA", "first (mlp_m)" if cta == 0 else "last (mlp_x)",
              " | ".join(f"{names[k]} +{st[k] - t0}" for k in (1, 2, 3, 8, 9, 4, 5)))

# backward (tc_edge_bwd256): one forward + backward, stamps of tile #3
bnames = {1: "GEMM1 observed", 10: "a1 barrier passed", 11: "image stores issued", 12: "GEMM2 issued", 2: "gather loads done", 3: "GEMM2 observed", 4: "gate pass+barrier", 5: "gz2 pass+barrier",
          6: "dgrad observed", 7: "gz1 pass+barrier", 8: "gu observed", 9: "tile end"}
for prec in ("bf16x3", "bf16"):
    layer.precision = prec
    hg, xg = h.clone().requires_grad_(True), x.clone().requires_grad_(True)
    for _ in range(2):
        layer.zero_grad(set_to_none=True)
        ho, xo = layer(ei, hg, xg)
        (ho.sum() + xo.sum()).backward()
    torch.cuda.synchronize()
    buf = (C.c_longlong * 32)()
    assert lib.egnn_dbg_read_stamps_bwd(buf) == 0
    for cta in (0, 1):
        t0 = buf[cta * 16]
        print(prec, "bwd CTA", "first (mlp_m)" if cta == 0 else "last (mlp_x)",
              " | ".join(f"{bnames[k]} +{buf[cta * 16 + k] - t0}" for k in (1, 10, 11, 12, 2, 3, 4, 5, 6, 7, 8, 9)))
