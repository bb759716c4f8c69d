"""Pencil-decomposed rfft2 on one GPU: the fused path (row pass writing column-split blocks, exchange, in-place
column pass) run directly at world size 1 against the single-plan transform and the fp64 oracle.  The multi-rank
exchange itself is covered by tests/test_pencil_gloo.py (CPU, gloo) and, on real NCCL / peer memory, by the check
that bench.py --gpus N prints into its JSON line (`pencil.rel_l2`)."""
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,H,W", [(2, 720, 1440), (3, 512, 1024), (1, 8192, 8192), (2, 180, 90)])
def test_fused_pencil_forward_matches_single_plan(B, H, W):
    from trt_dft_plugins import capi
    from trt_dft_plugins.pencil import PencilRfft2
    torch.manual_seed(11)
    x = torch.randn(B, H, W, device="cuda")
    p = PencilRfft2(B, H, W)                       # no process group: world size 1, this rank owns everything
    z = p._forward_fused(x)
    torch.cuda.synchronize()
    ref = capi.rfftn(x, 2)
    assert O.rel_l2(z.cpu().numpy(), ref.cpu().numpy()) <= 1e-6
    want = O.torch_rfftn(x[B - 1].cpu(), 2, dtype=torch.float64)
    assert O.rel_l2(z[B - 1].cpu().numpy(), want.numpy()) <= 1e-5
