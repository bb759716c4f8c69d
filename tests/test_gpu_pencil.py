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


@pytest.mark.parametrize("B,H,W", [(2, 720, 1440), (1, 8192, 8192), (3, 96, 200), (2, 34, 64)])
def test_peer_store_kernels_forward_and_inverse_on_one_gpu(B, H, W):
    """The kernels of the peer exchange (row pass storing column blocks into the owners' slabs, column pass storing
    rows into the owners' row bands) with the one rank's "peer" memory being local memory: forward against the
    single-plan rfft2, inverse back to the image."""
    from trt_dft_plugins import capi
    from trt_dft_plugins.pencil import PencilRfft2
    torch.manual_seed(12)
    x = torch.randn(B, H, W, device="cuda")
    p = PencilRfft2(B, H, W, exchange="peer")
    z = p.forward(x)
    assert p.exchange_used == "peer:local"
    torch.cuda.synchronize()
    ref = capi.rfftn(x, 2)
    assert O.rel_l2(z.cpu().numpy(), ref.cpu().numpy()) <= 1e-6
    zc = z.clone()
    xb = p.inverse(z)                          # row pass gathering from the slab (persistent kernel) or the store form
    torch.cuda.synchronize()
    assert O.rel_l2(xb.cpu().numpy(), x.cpu().numpy()) <= 1e-6
    p.inverse_exchange = "stores"              # column pass storing into the row band
    xs = p.inverse(zc)
    torch.cuda.synchronize()
    assert O.rel_l2(xs.cpu().numpy(), x.cpu().numpy()) <= 1e-6


def test_peer_split_addresses_with_several_destinations_on_one_gpu():
    """Several 'ranks' emulated on one GPU: each destination is its own local buffer.  Pins the address arithmetic of
    trt_dft_plan_set_peer_column_split / _row_split (uneven splits, an empty part, row offset inside the image)."""
    import ctypes
    from trt_dft_plugins import capi
    L = capi.lib()
    L.trt_dft_plan_set_peer_column_split.restype = ctypes.c_int
    L.trt_dft_plan_set_peer_column_split.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int64), ctypes.c_int64,
                                                     ctypes.c_int64, ctypes.c_int64, ctypes.POINTER(ctypes.c_void_p),
                                                     ctypes.POINTER(ctypes.c_int64)]
    L.trt_dft_plan_set_peer_row_split.restype = ctypes.c_int
    L.trt_dft_plan_set_peer_row_split.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int64), ctypes.c_int64,
                                                  ctypes.c_int64, ctypes.POINTER(ctypes.c_void_p)]
    torch.manual_seed(13)
    for (B, H, W, hn, h0) in ((2, 720, 1440, 300, 111), (1, 8192, 8192, 1024, 4096), (2, 60, 100, 17, 3)):
        Mo = W // 2 + 1
        starts = [0, Mo // 3, Mo // 3, Mo - 7, Mo]                 # 4 destinations, one empty
        x = torch.randn(B, hn, W, device="cuda")                   # this "rank"'s row band
        slabs = [torch.full((B, H, max(starts[j + 1] - starts[j], 1), 2), float("nan"), device="cuda") for j in range(4)]
        rows = capi.Plan(capi.FORWARD, (W,), B * hn)
        rc = L.trt_dft_plan_set_peer_column_split(rows.handle, 4, (ctypes.c_int64 * 5)(*starts), hn, H, h0,
                                                  (ctypes.c_void_p * 4)(*[s.data_ptr() for s in slabs]), None)
        assert rc == 0
        dummy = torch.empty(B * hn * Mo * 2, device="cuda")
        rows.exec_ptr(x.data_ptr(), dummy.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        plain = capi.rfftn(x, 1)                                   # [B, hn, Mo, 2]
        for j in range(4):
            cn = starts[j + 1] - starts[j]
            if cn == 0:
                continue
            got = slabs[j][:, h0:h0 + hn]
            assert O.rel_l2(got.cpu().numpy(), plain[:, :, starts[j]:starts[j + 1]].cpu().numpy()) <= 1e-6
            assert torch.isnan(slabs[j][:, :h0]).all() and torch.isnan(slabs[j][:, h0 + hn:]).all()   # nothing else touched
        # the same with even column starts and even slab pitches: runs leave as bulk copies out of staged rows (persistent
        # row kernel; shapes too small for it keep the 8-byte stores) -- the padding column may be written, nothing else
        bstarts = [0, (Mo // 3) // 2 * 2, (Mo // 3) // 2 * 2, (Mo - 7) // 2 * 2, Mo]
        pitches = [max((bstarts[j + 1] - bstarts[j] + 1) // 2 * 2, 2) for j in range(4)]
        bslabs = [torch.full((B, H, pitches[j], 2), float("nan"), device="cuda") for j in range(4)]
        rows2 = capi.Plan(capi.FORWARD, (W,), B * hn)
        rc = L.trt_dft_plan_set_peer_column_split(rows2.handle, 4, (ctypes.c_int64 * 5)(*bstarts), hn, H, h0,
                                                  (ctypes.c_void_p * 4)(*[s.data_ptr() for s in bslabs]), (ctypes.c_int64 * 4)(*pitches))
        assert rc == 0
        rows2.exec_ptr(x.data_ptr(), dummy.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
        rows2.exec_ptr(x.data_ptr(), dummy.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)      # staging rows are reused
        torch.cuda.synchronize()
        for j in range(4):
            cn = bstarts[j + 1] - bstarts[j]
            if cn == 0:
                continue
            got = bslabs[j][:, h0:h0 + hn, :cn]
            assert O.rel_l2(got.cpu().numpy(), plain[:, :, bstarts[j]:bstarts[j + 1]].cpu().numpy()) <= 1e-6
            assert torch.isnan(bslabs[j][:, :h0]).all() and torch.isnan(bslabs[j][:, h0 + hn:]).all()
        # mirror: a column band [B, H, cn] -> rows scattered into 4 row bands at column offset c0
        cn, c0 = 37, 5
        z = torch.randn(B, H, cn, 2, device="cuda")
        rstarts = [0, H // 4, H // 4, H - 3, H]
        bands = [torch.full((B, max(rstarts[j + 1] - rstarts[j], 1), Mo, 2), float("nan"), device="cuda") for j in range(4)]
        cols = capi.C2CPlan(capi.INVERSE, H, B, cn, scale=0.5)
        rc = L.trt_dft_plan_set_peer_row_split(cols.handle, 4, (ctypes.c_int64 * 5)(*rstarts), Mo, c0,
                                               (ctypes.c_void_p * 4)(*[b.data_ptr() for b in bands]))
        assert rc == 0
        cols.exec(z, torch.empty_like(z))
        torch.cuda.synchronize()
        want = 0.5 * torch.view_as_real(torch.fft.ifft(torch.view_as_complex(z.double()), dim=1, norm="forward"))
        for j in range(4):
            rn = rstarts[j + 1] - rstarts[j]
            if rn == 0:
                continue
            got = bands[j][:, :, c0:c0 + cn]
            assert O.rel_l2(got.cpu().numpy(), want[:, rstarts[j]:rstarts[j + 1]].cpu().numpy()) <= 1e-5
            assert torch.isnan(bands[j][:, :, :c0]).all() and torch.isnan(bands[j][:, :, c0 + cn:]).all()


def _two_gpu_worker(rank, world, port, out_dir):
    import os
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    sys.path.insert(0, str(root / "tensorrt-dft-plugins_b200"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from trt_dft_plugins import capi
    from trt_dft_plugins.pencil import PencilRfft2
    res = {}
    for (B, H, W) in ((2, 720, 1440), (1, 8192, 2048), (3, 96, 200)):
        g = torch.Generator(device=dev).manual_seed(21)
        full = torch.randn(B, H, W, device=dev, generator=g)          # same on both ranks
        ref = capi.rfftn(full, 2)
        for mode in ("peer", "nccl"):
            p = PencilRfft2(B, H, W, device=dev, exchange="auto" if mode == "peer" else "nccl")
            x_loc = full[:, p.h0:p.h0 + p.hn].contiguous()
            y = p.forward(x_loc)
            torch.cuda.synchronize()
            want = ref[:, :, p.c0:p.c0 + p.cn]
            e_f = ((y - want).double().norm() / want.double().norm()).item()
            yc = y.clone()                          # (the gathering inverse transforms the slab -- i.e. y -- in place)
            xb = p.inverse(yc.clone())
            torch.cuda.synchronize()
            e_i = ((xb - x_loc).double().norm() / x_loc.double().norm()).item()
            used = p.exchange_used
            if mode == "peer":                      # the inverse's other peer form: column kernel storing into row bands
                p.inverse_exchange = "stores"
                xs = p.inverse(yc)
                torch.cuda.synchronize()
                e_i = max(e_i, ((xs - x_loc).double().norm() / x_loc.double().norm()).item())
            res[(B, H, W, mode)] = (e_f, e_i, used)
    torch.save(res, os.path.join(out_dir, f"r{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_pencil_on_two_gpus_over_peer_memory_and_nccl(tmp_path):
    """The real exchange, both ways of running it and both directions, at world size 2 (skipped on a one-GPU box; the same
    check runs inside `bench.py --gpus N` and is printed as `pencil.rel_l2`)."""
    import os
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    port = 32500 + (os.getpid() % 2000)
    mp.spawn(_two_gpu_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        res = torch.load(tmp_path / f"r{r}.pt")
        for key, (e_f, e_i, used) in res.items():
            assert e_f <= 1e-6 and e_i <= 1e-6, (key, e_f, e_i, used)
            assert used.startswith("peer") or key[3] == "nccl", (key, used)
