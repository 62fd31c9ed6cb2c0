"""Tensor-parallel communicator whose all-reduce lives *inside* the decode mat-vec kernels.

North-star item: "the column->row GEMM pair fuses the collective into the kernel: GEMM -> reduce/all-reduce issues
NVSwitch P2P stores from inside the kernel".  Decode (T <= 4) path:

    o_proj / down GEMV (row-parallel, K shard)  --epilogue-->  st.global of the partial rows into EVERY peer's
                                                               reduce buffer + st.release.sys of an epoch flag
    next GEMV (QKV / gate+up / lm_head)         <--prologue--  ld.acquire.sys on the tp flags, then
                                                               h = h + sum_r partial[r] fused with RMSNorm

so a layer needs no separate collective launch and no NCCL call.  Buffers are symmetric allocations mapped
into every rank through CUDA IPC (`csrc/comm.cu`).  Larger token counts (prefill, big decode batches) use the
NCCL path of :class:`NcclComm` - that path is also the A/B baseline the spec asks to keep.

Reference counterpart: `ggml_cuda_op_mul_mat` broadcast + gather through a hub GPU (`ggml-cuda.cu:1451-1769`).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from ..ops import lib as _lib
from .comm import NcclComm

MAX_T = 4  # token rows per site in the fused path


class dlp_ar_in(C.Structure):
    _fields_ = [("buf", C.c_void_p), ("flags", C.c_void_p), ("epoch", C.c_void_p), ("h_out", C.c_void_p),
                ("error", C.c_void_p), ("tp", C.c_int32), ("reserved", C.c_int32)]


class dlp_ar_out(C.Structure):
    _fields_ = [("peer_buf", C.c_void_p * 8), ("peer_flags", C.c_void_p * 8), ("epoch", C.c_void_p),
                ("done_counter", C.c_void_p), ("rank", C.c_int32), ("tp", C.c_int32)]


class dlp_gemm_coll(C.Structure):
    _fields_ = [("x_flags", C.c_void_p), ("x_rows_per_src", C.c_int32), ("x_epoch", C.c_uint32), ("error", C.c_void_p),
                ("rs_peer_buf", C.c_void_p * 8), ("rs_peer_flags", C.c_void_p * 8), ("rs_done", C.c_void_p),
                ("rs_rows", C.c_int32), ("rs_rank", C.c_int32), ("rs_tp", C.c_int32), ("rs_epoch", C.c_uint32)]


class dlp_rs_norm_ag_desc(C.Structure):
    _fields_ = [("h_shard", C.c_void_p), ("red_buf", C.c_void_p), ("red_flags", C.c_void_p), ("red_epoch", C.c_uint32),
                ("pad0", C.c_uint32), ("norm_w", C.c_void_p), ("eps", C.c_float), ("pad1", C.c_uint32),
                ("ag_peer_buf", C.c_void_p * 8), ("ag_peer_flags", C.c_void_p * 8), ("ag_epoch", C.c_uint32), ("pad2", C.c_uint32),
                ("done", C.c_void_p), ("error", C.c_void_p), ("rank", C.c_int32), ("tp", C.c_int32), ("rows", C.c_int32),
                ("E", C.c_int32)]


class FusedComm(NcclComm):
    fused = True

    def __init__(self, grid, n_embd: int | None = None, n_sites: int | None = None):
        super().__init__(grid.tp_group, grid.tp, grid.tp_rank)
        self.grid = grid
        self.device = grid.device
        self.lib = _lib.load()
        self._declare()
        self.ready = False
        self.base = None
        self.peers: list[int] = []
        if n_embd is not None and n_sites is not None:
            self.setup(n_embd, n_sites)

    def _declare(self):
        L, vp, i32 = self.lib, C.c_void_p, C.c_int
        W = C.POINTER(_lib.dlp_weight)
        L.dlp_gemv_tp.restype = i32
        L.dlp_gemv_tp.argtypes = [W, W, vp, vp, C.c_float, vp, i32, i32, i32, C.POINTER(dlp_ar_in), C.POINTER(dlp_ar_out), vp]
        L.dlp_ar_consume.restype = i32
        L.dlp_ar_consume.argtypes = [vp, vp, C.POINTER(dlp_ar_in), i32, i32, vp]
        L.dlp_bump.restype = i32
        L.dlp_bump.argtypes = [vp, vp]
        L.dlp_symm_alloc.restype = i32
        L.dlp_symm_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
        L.dlp_ipc_export.restype = i32
        L.dlp_ipc_export.argtypes = [vp, vp]
        L.dlp_ipc_import.restype = i32
        L.dlp_ipc_import.argtypes = [vp, C.POINTER(vp)]
        L.dlp_gemm_sp.restype = i32
        L.dlp_gemm_sp.argtypes = [W, W, vp, i32, vp, i32, i32, i32, C.POINTER(dlp_gemm_coll), vp]
        L.dlp_rs_norm_ag.restype = i32
        L.dlp_rs_norm_ag.argtypes = [C.POINTER(dlp_rs_norm_ag_desc), vp]

    # ---- set-up: one symmetric allocation per rank, handles exchanged over the process group --------------
    def setup(self, n_embd: int, n_sites: int):
        tp = self.tp_size
        self.E, self.n_sites = n_embd, n_sites
        self.buf_bytes = tp * MAX_T * n_embd * 4
        self.site_stride = self.buf_bytes + 256          # reduce buffer + flag line
        total = n_sites * self.site_stride
        p = C.c_void_p()
        _lib.check(self.lib.dlp_symm_alloc(total, C.byref(p)), "symm_alloc")
        self.base = p.value
        handle = (C.c_char * 64)()
        _lib.check(self.lib.dlp_ipc_export(self.base, handle), "ipc_export")
        mine = bytes(handle)
        gathered: list = [None] * tp
        dist.all_gather_object(gathered, mine, group=self.group)
        self.peers = []
        for r, h in enumerate(gathered):
            if r == self.tp_rank:
                self.peers.append(self.base)
                continue
            q = C.c_void_p()
            hb = (C.c_char * 64).from_buffer_copy(h)
            _lib.check(self.lib.dlp_ipc_import(hb, C.byref(q)), "ipc_import")
            self.peers.append(q.value)
        self.epoch = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.done = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.error = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._in: dict[int, dlp_ar_in] = {}
        self._out: dict[int, dlp_ar_out] = {}
        dist.barrier(group=self.group)
        self.ready = True

    @classmethod
    def make_virtual(cls, tp: int, n_embd: int, n_sites: int, device="cuda") -> list["FusedComm"]:
        """`tp` communicators that live on ONE device and point at each other's buffers directly - no IPC, no process
        group.  Used to test the push / flag / consume protocol on a single GPU: the caller launches each phase for all
        virtual ranks in turn, so every flag is already published when a consumer looks at it (kernels that wait on each
        other must never be co-scheduled on one GPU)."""
        comms = []
        for r in range(tp):
            c = cls.__new__(cls)
            c.group, c.tp_size, c.tp_rank, c._gather_buf = None, tp, r, None
            c.grid, c.device, c.lib = None, torch.device(device), _lib.load()
            c._declare()
            c.E, c.n_sites = n_embd, n_sites
            c.buf_bytes = tp * MAX_T * n_embd * 4
            c.site_stride = c.buf_bytes + 256
            c._mem = torch.zeros(n_sites * c.site_stride, dtype=torch.uint8, device=device)
            c.base = c._mem.data_ptr()
            c.epoch = torch.zeros(1, dtype=torch.int32, device=device)
            c.done = torch.zeros(1, dtype=torch.int32, device=device)
            c.error = torch.zeros(1, dtype=torch.int32, device=device)
            c._in, c._out = {}, {}
            c.ready = True
            comms.append(c)
        for c in comms:
            c.peers = [o.base for o in comms]
        return comms

    @classmethod
    def make_virtual_sp(cls, tp: int, E: int, t_max: int, device="cuda") -> list["FusedComm"]:
        """Virtual ranks (one GPU) with the sequence-parallel gather / reduce buffers wired to each other."""
        comms = cls.make_virtual(tp, E, 1, device)
        for c in comms:
            total = c.sp_layout(E, t_max)
            c._sp_mem = torch.zeros(total, dtype=torch.uint8, device=device)
            c.sp_base = c._sp_mem.data_ptr()
            c._sp_state()
        for c in comms:
            c.sp_peers = [o.sp_base for o in comms]
        return comms

    def ar_in(self, site: int, h_out: torch.Tensor) -> dlp_ar_in:
        key = (site, h_out.data_ptr())
        d = self._in.get(key)
        if d is None:
            b = self.base + site * self.site_stride
            d = dlp_ar_in(b, b + self.buf_bytes, self.epoch.data_ptr(), h_out.data_ptr(), self.error.data_ptr(), self.tp_size, 0)
            self._in[key] = d
        return d

    def ar_out(self, site: int) -> dlp_ar_out:
        d = self._out.get(site)
        if d is None:
            d = dlp_ar_out()
            for r in range(self.tp_size):
                b = self.peers[r] + site * self.site_stride
                d.peer_buf[r] = b
                d.peer_flags[r] = b + self.buf_bytes
            d.epoch = self.epoch.data_ptr()
            d.done_counter = self.done.data_ptr()
            d.rank, d.tp = self.tp_rank, self.tp_size
            self._out[site] = d
        return d

    # ---- sequence-parallel prefill: GEMM -> reduce-scatter, all-gather -> GEMM ---------------------------------
    # Symmetric layout per rank:  G0 | G1 (gather buffers [Tmax, E] fp16)  R0 | R1 (reduce buffers [tp][Tmax/tp][E] fp32)
    #                             4 flag lines of 64 B (G0, G1, R0, R1)
    def sp_layout(self, E: int, t_max: int):
        tp = self.tp_size
        t_max = (t_max + tp - 1) // tp * tp
        g_bytes = t_max * E * 2
        r_bytes = t_max * E * 4
        self.sp_E, self.sp_tmax = E, t_max
        self.sp_off = {"G0": 0, "G1": g_bytes, "R0": 2 * g_bytes, "R1": 2 * g_bytes + r_bytes, "F": 2 * g_bytes + 2 * r_bytes}
        return 2 * g_bytes + 2 * r_bytes + 4 * 64

    def sp_setup(self, E: int, t_max: int):
        total = self.sp_layout(E, t_max)
        p = C.c_void_p()
        _lib.check(self.lib.dlp_symm_alloc(total, C.byref(p)), "symm_alloc")
        self.sp_base = p.value
        handle = (C.c_char * 64)()
        _lib.check(self.lib.dlp_ipc_export(self.sp_base, handle), "ipc_export")
        gathered: list = [None] * self.tp_size
        dist.all_gather_object(gathered, bytes(handle), group=self.group)
        self.sp_peers = []
        for r, h in enumerate(gathered):
            if r == self.tp_rank:
                self.sp_peers.append(self.sp_base)
                continue
            q = C.c_void_p()
            _lib.check(self.lib.dlp_ipc_import((C.c_char * 64).from_buffer_copy(h), C.byref(q)), "ipc_import")
            self.sp_peers.append(q.value)
        self._sp_state()
        dist.barrier(group=self.group)

    def _sp_state(self):
        self.g_epoch, self.r_epoch = [0, 0], [0, 0]
        self.sp_done = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.sp_ready = True

    def sp_tensor(self, g: int) -> torch.Tensor:
        """Zero-copy torch view ([Tmax, E] fp16) of this rank's gather buffer g."""
        class _Mem:
            pass
        m = _Mem()
        n = self.sp_tmax * self.sp_E * 2
        m.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (self.sp_base + self.sp_off[f"G{g}"], False), "version": 2}
        t = torch.as_tensor(m, device=self.device)
        self._keep = getattr(self, "_keep", []) + [m]
        return t.view(torch.float16).view(self.sp_tmax, self.sp_E)

    def sp_gather_view(self, g: int, T: int) -> int:
        """Device address of this rank's gather buffer g (the X operand of the next column-parallel GEMM)."""
        return self.sp_base + self.sp_off[f"G{g}"]

    def sp_norm_ag(self, h_shard: torch.Tensor, rows: int, norm_w: torch.Tensor | None, eps: float, g: int | None, red: int | None):
        """Fold reduce buffer `red` (if any) into h_shard, RMSNorm, push fp16 rows into every peer's gather buffer `g`."""
        from ..ops import api as ops
        d = dlp_rs_norm_ag_desc()
        d.h_shard = h_shard.data_ptr()
        if red is not None:
            d.red_buf = self.sp_base + self.sp_off[f"R{red}"]
            d.red_flags = self.sp_base + self.sp_off["F"] + (2 + red) * 64
            d.red_epoch = self.r_epoch[red]
        if norm_w is not None:
            self.g_epoch[g] += 1
            d.norm_w, d.eps, d.ag_epoch = norm_w.data_ptr(), eps, self.g_epoch[g]
            for r in range(self.tp_size):
                d.ag_peer_buf[r] = self.sp_peers[r] + self.sp_off[f"G{g}"]
                d.ag_peer_flags[r] = self.sp_peers[r] + self.sp_off["F"] + g * 64
        d.done, d.error = self.sp_done.data_ptr(), self.error.data_ptr()
        d.rank, d.tp, d.rows, d.E = self.tp_rank, self.tp_size, rows, self.sp_E
        _lib.check(self.lib.dlp_rs_norm_ag(C.byref(d), torch.cuda.current_stream().cuda_stream), "rs_norm_ag")
        ops._count()

    def sp_gemm(self, w, x, ldx: int, T: int, rows: int, epilogue: int, y: torch.Tensor | None = None, w_up=None,
                x_from_g: int | None = None, rs_to: int | None = None, x_owner: int | None = None):
        """tcgen05 GEMM with the collective fused in: X from gather buffer `x_from_g` (waits per owner) and / or the
        reduce-scatter push epilogue into reduce buffer `rs_to` of the owner ranks.  `x` is a tensor or a device address."""
        from ..ops import api as ops
        c = dlp_gemm_coll()
        c.error = self.error.data_ptr()
        if x_from_g is not None:
            c.x_flags = self.sp_base + self.sp_off["F"] + x_from_g * 64
            c.x_rows_per_src, c.x_epoch = rows, self.g_epoch[x_from_g]
            if x_owner is not None:   # X is a slice owned by a single rank: wait for that rank only
                c.x_flags += 4 * x_owner
        if rs_to is not None:
            self.r_epoch[rs_to] += 1
            for r in range(self.tp_size):
                c.rs_peer_buf[r] = self.sp_peers[r] + self.sp_off[f"R{rs_to}"]
                c.rs_peer_flags[r] = self.sp_peers[r] + self.sp_off["F"] + (2 + rs_to) * 64
            c.rs_done, c.rs_rows, c.rs_rank, c.rs_tp, c.rs_epoch = self.sp_done.data_ptr(), rows, self.tp_rank, self.tp_size, self.r_epoch[rs_to]
        xp = x if isinstance(x, int) else x.data_ptr()
        _lib.check(self.lib.dlp_gemm_sp(w.ref(), w_up.ref() if w_up is not None else None, xp, ldx, None if y is None else y.data_ptr(),
                                        0 if y is None else y.stride(0), epilogue, T, C.byref(c), torch.cuda.current_stream().cuda_stream),
                   "gemm_sp")
        ops._count()
        return y

    # ---- per-step -----------------------------------------------------------------------------------
    def begin_step(self):
        """Advance the epoch (device-side, graph-capturable); every rank calls it once per forward pass."""
        from ..ops import api as ops
        _lib.check(self.lib.dlp_bump(self.epoch.data_ptr(), torch.cuda.current_stream().cuda_stream), "bump")
        ops._count()

    def gemv(self, w, x, out_mode: int, T: int, out=None, w_up=None, norm_w=None, eps: float = 0.0, ar_in=None, ar_out=None):
        from ..ops import api as ops
        _lib.check(self.lib.dlp_gemv_tp(w.ref(), w_up.ref() if w_up is not None else None, x.data_ptr(),
                                        None if norm_w is None else norm_w.data_ptr(), eps,
                                        None if out is None else out.data_ptr(), 0 if out is None else out.stride(0), out_mode, T,
                                        None if ar_in is None else C.byref(ar_in), None if ar_out is None else C.byref(ar_out),
                                        torch.cuda.current_stream().cuda_stream), "gemv_tp")
        ops._count()
        return out

    def consume(self, h_in: torch.Tensor, h_out: torch.Tensor, site: int, T: int):
        from ..ops import api as ops
        d = self.ar_in(site, h_out)
        _lib.check(self.lib.dlp_ar_consume(h_in.data_ptr(), h_out.data_ptr(), C.byref(d), self.E, T,
                                           torch.cuda.current_stream().cuda_stream), "ar_consume")
        ops._count()

    def check_error(self):
        if int(self.error.item()):
            raise RuntimeError("fused all-reduce: a peer did not signal within the timeout (ranks diverged or died)")
