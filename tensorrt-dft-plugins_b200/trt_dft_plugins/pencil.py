"""Pencil (slab) decomposed rfft2 for grids too large for one GPU pass -- BASELINE config 5,
`[4, 8192, 8192]` over 2/4/8 GPUs.

Layout.  Every rank holds a contiguous band of ROWS of every image, x_local [B, H_r, W].
  1. local row pass      R2C along W                          -> [B, H_r, W/2+1] complex
  2. all-to-all transpose (the only exchange on the path): rank r receives, from every rank s, the
     columns [c0_r, c1_r) of s's rows                          -> [B, H, C_r] complex on rank r
  3. local column pass   C2C of length H along the row axis     (data[outer=B][n=H][inner=C_r])
The result stays distributed by columns: rank r owns spectrum[:, :, c0_r:c1_r].  W/2+1 is odd for
every even W, so the column split is uneven (all_to_all_single with explicit split sizes).

The transforms are injected (`row_fft`, `col_fft`) so that the exchange logic can be tested on CPU
with the gloo backend; on GPUs they default to the library's plans (trt_dft_plugins.capi) and the
exchange runs over NCCL / NVLink.
"""
from typing import Callable, Optional

import torch
import torch.distributed as dist

from .multi_gpu import shard_batch


class PencilRfft2:
    def __init__(self, B: int, H: int, W: int, group=None, device=None,
                 row_fft: Optional[Callable] = None, col_fft: Optional[Callable] = None):
        self.B, self.H, self.W, self.Mo = int(B), int(H), int(W), int(W) // 2 + 1
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.device = device
        self.row_spans = [shard_batch(self.H, r, self.world) for r in range(self.world)]    # (h0, hn) per rank
        self.col_spans = [shard_batch(self.Mo, r, self.world) for r in range(self.world)]   # (c0, cn) per rank
        self.h0, self.hn = self.row_spans[self.rank]
        self.c0, self.cn = self.col_spans[self.rank]
        self._row_fft, self._col_fft = row_fft, col_fft
        self._plans = None

    # ---- default GPU transforms ---------------------------------------------------------------
    def _gpu_plans(self):
        if self._plans is None:
            from . import capi
            rows = capi.Plan(capi.FORWARD, (self.W,), self.B * self.hn)
            cols = capi.C2CPlan(capi.FORWARD, self.H, self.B, self.cn)
            self._plans = (rows, cols)
        return self._plans

    def row_pass(self, x_local):
        if self._row_fft is not None:
            return self._row_fft(x_local)
        rows, _ = self._gpu_plans()
        out = torch.empty(self.B, self.hn, self.Mo, 2, dtype=torch.float32, device=x_local.device)
        rows.exec(x_local.contiguous(), out)
        return out

    def col_pass(self, z):
        if self._col_fft is not None:
            return self._col_fft(z)
        _, cols = self._gpu_plans()
        cols.exec(z, z)                      # in place: each column tile is read completely before it is written
        return z

    # ---- the exchange ----------------------------------------------------------------------------
    def transpose(self, y_rows):
        """[B, H_r, Mo, 2] on every rank  ->  [B, H, C_r, 2] on every rank."""
        if self.world == 1:
            return y_rows
        B = self.B
        send = torch.cat([y_rows[:, :, c0:c0 + cn, :].reshape(-1) for (c0, cn) in self.col_spans])
        in_splits = [B * self.hn * cn * 2 for (_, cn) in self.col_spans]
        out_splits = [B * hn_s * self.cn * 2 for (_, hn_s) in self.row_spans]
        recv = torch.empty(sum(out_splits), dtype=y_rows.dtype, device=y_rows.device)
        dist.all_to_all_single(recv, send, output_split_sizes=out_splits, input_split_sizes=in_splits, group=self.group)
        z = torch.empty(B, self.H, self.cn, 2, dtype=y_rows.dtype, device=y_rows.device)
        off = 0
        for (h0_s, hn_s), n in zip(self.row_spans, out_splits):
            z[:, h0_s:h0_s + hn_s] = recv[off:off + n].view(B, hn_s, self.cn, 2)
            off += n
        return z

    def forward(self, x_local):
        """x_local: this rank's row band [B, H_r, W] (fp32).  Returns its column band [B, H, C_r, 2]."""
        assert tuple(x_local.shape) == (self.B, self.hn, self.W)
        if self._row_fft is not None or self._col_fft is not None or not x_local.is_cuda or self.world == 1:
            return self.col_pass(self.transpose(self.row_pass(x_local)))
        return self._forward_fused(x_local)

    # ---- GPU path: pack fused into the row pass, one all-to-all per image straight into place ----------
    def _fused_plans(self):
        if getattr(self, "_fplans", None) is None:
            from . import capi
            import ctypes
            rows = capi.Plan(capi.FORWARD, (self.W,), self.hn)                   # one image's row band per launch
            L = capi.lib()
            L.trt_dft_plan_set_column_split.restype = ctypes.c_int
            L.trt_dft_plan_set_column_split.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int64), ctypes.c_int64]
            starts = [c0 for (c0, _) in self.col_spans] + [self.Mo]
            arr = (ctypes.c_int64 * len(starts))(*starts)
            rc = L.trt_dft_plan_set_column_split(rows.handle, self.world, arr, self.hn)
            if rc != 0:
                raise capi.DftError(f"trt_dft_plan_set_column_split failed with status {rc}")
            cols = capi.C2CPlan(capi.FORWARD, self.H, 1, self.cn)               # one image's column band per launch
            self._fplans = (rows, cols)
        return self._fplans

    def _forward_fused(self, x_local):
        """Row pass of image b writes its spectrum already split by destination rank (the send buffer of the
        all-to-all); the all-to-all of image b lands directly in z[b] = [H, C_r] (blocks arrive in source-rank
        order = row order), so there is no pack or unpack copy, and it overlaps the row pass of image b+1 and
        the column pass of image b-1."""
        rows, cols = self._fused_plans()
        B, dev = self.B, x_local.device
        send = torch.empty(B, self.hn * self.Mo, 2, dtype=torch.float32, device=dev)
        z = torch.empty(B, self.H, self.cn, 2, dtype=torch.float32, device=dev)
        in_splits = [self.hn * cn * 2 for (_, cn) in self.col_spans]
        out_splits = [hn_s * self.cn * 2 for (_, hn_s) in self.row_spans]
        works = []
        for b in range(B):
            rows.exec(x_local[b], send[b])
            if self.world == 1:
                z[b].view(-1).copy_(send[b].view(-1))
                works.append(None)
            else:
                works.append(dist.all_to_all_single(z[b].view(-1), send[b].view(-1), output_split_sizes=out_splits,
                                                    input_split_sizes=in_splits, group=self.group, async_op=True))
        for b in range(B):
            if works[b] is not None:
                works[b].wait()
            cols.exec(z[b], z[b])
        return z
