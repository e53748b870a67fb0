"""tcgen05 path: one-tile GEMM self-test (pins descriptor conventions on hardware) and, further down, parity of the
tensor-core edge kernels against the fp32 reference results."""
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _selftest(K, N, npass, swap):
    from diffusion_model_b200._lib import call, ptr, stream_ptr
    gen = torch.Generator().manual_seed(K * 1000 + N)
    A = torch.randn(128, K, generator=gen).to(DEV)
    B = torch.randn(N, K, generator=gen).to(DEV)
    D = torch.full((128, N), float("nan"), device=DEV)
    status = torch.zeros(1, dtype=torch.int32, device=DEV)
    call("egnn_tc_selftest", ptr(A), ptr(B), ptr(D), K, N, npass, swap, ptr(status), stream_ptr())
    torch.cuda.synchronize()
    ref = A.double() @ B.double().t()
    err = float((D.double() - ref).abs().max() / ref.abs().max())
    return err, int(status.item())


def test_tc_selftest_descriptor_convention(lib):
    out = {}
    os.makedirs("gpurun_out", exist_ok=True)
    for swap in (0, 1):
        for K, N in ((16, 16), (16, 256), (64, 64), (128, 128)):
            for npass in (1, 3):
                err, st = _selftest(K, N, npass, swap)
                out[f"K{K}_N{N}_p{npass}_swap{swap}"] = (err, st)
                json.dump(out, open("gpurun_out/tc_selftest.json", "w"), indent=1)
    for K, N in ((16, 16), (16, 256), (64, 64), (128, 128)):
        e1, s1 = out[f"K{K}_N{N}_p1_swap0"]
        e3, s3 = out[f"K{K}_N{N}_p3_swap0"]
        assert s1 == 0 and s3 == 0, out
        assert e1 < 2e-2, out   # single bf16 product
        assert e3 < 2e-5, out   # hi/lo split, 3 products
