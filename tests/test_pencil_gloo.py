"""Pencil-decomposed rfft2: the all-to-all transpose logic at world size 2 over gloo (CPU).

The two local passes are stood in for by the oracle's torch.fft expressions -- this test pins the
exchange (uneven column split, pack / unpack order); the CUDA passes are covered by -m gpu tests."""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, str(ROOT / "tensorrt-dft-plugins_b200"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from trt_dft_plugins.pencil import PencilRfft2
    torch.manual_seed(1)
    B, H, W = 3, 10, 12                                # Mo = 7 columns: split 2 + 5 (even band starts); rows 5 + 5
    full = torch.randn(B, H, W)
    row_fft = lambda x: torch.view_as_real(torch.fft.rfft(x, dim=-1))
    col_fft = lambda z: torch.view_as_real(torch.fft.fft(torch.view_as_complex(z.contiguous()), dim=1))
    p = PencilRfft2(B, H, W, row_fft=row_fft, col_fft=col_fft)
    y = p.forward(full[:, p.h0:p.h0 + p.hn].contiguous())
    want = torch.view_as_real(torch.fft.rfft2(full, dim=(-2, -1)))[:, :, p.c0:p.c0 + p.cn]
    err = ((y - want).norm() / want.norm()).item()
    # and back: column band of the spectrum -> row band of the image (irfft2 semantics, scaled by 1/(H*W))
    inv_col = lambda z: torch.view_as_real(torch.fft.ifft(torch.view_as_complex(z.contiguous()), dim=1))
    inv_row = lambda t: torch.fft.irfft(torch.view_as_complex(t.contiguous()), n=W, dim=-1)
    xb = p.inverse(y.contiguous(), inv_row_fft=inv_row, inv_col_fft=inv_col)
    err_inv = ((xb - full[:, p.h0:p.h0 + p.hn]).norm() / full[:, p.h0:p.h0 + p.hn].norm()).item()
    torch.save({"err": err, "err_inv": err_inv, "shape": tuple(y.shape), "cn": p.cn}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_pencil_transpose_two_ranks(tmp_path):
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = torch.load(tmp_path / "r0.pt"), torch.load(tmp_path / "r1.pt")
    assert r0["cn"] + r1["cn"] == 7 and r0["shape"] == (3, 10, 2, 2) and r1["shape"] == (3, 10, 5, 2)
    assert r0["err"] < 1e-6 and r1["err"] < 1e-6
    assert r0["err_inv"] < 1e-6 and r1["err_inv"] < 1e-6
