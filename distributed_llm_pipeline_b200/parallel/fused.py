"""Tensor-parallel communicator whose collectives live *inside* the kernels.

North-star item: "the column->row GEMM pair fuses the collective into the kernel: GEMM -> reduce/all-reduce issues
NVSwitch P2P or NVLS-multicast stores from inside the kernel".  Decode (T <= 4) path:

    o_proj / down GEMV (row-parallel, K shard)  --epilogue-->  partial rows stored into EVERY rank's buffer of the site:
                                                               one `multimem.st` on the NVLS multicast alias (the switch
                                                               replicates it) or one peer store per rank
    next GEMV (QKV / gate+up / lm_head)         <--prologue--  spins on the values themselves (a value is its own arrival
                                                               signal), then h = h + sum_r partial[r] fused with RMSNorm
    lm_head (vocab shard) -> dlp_argmax_tp                     sharded greedy sampling: 8-byte (value, id) words exchanged the
                                                               same way; no logits all-gather

so a decode step under TP has no collective launch, no flag, no system-scope fence and no NCCL call.  Buffers are symmetric
allocations: `torch.distributed._symmetric_memory` when it gives a multicast pointer (plumbing only - the stores are ours), else
cudaMalloc + CUDA IPC (`csrc/comm.cu`).  Larger token counts (prefill, big decode batches) use the sequence-parallel GEMM path
below; the NCCL path of :class:`NcclComm` is the A/B baseline the spec asks to keep.

Reference counterpart: `ggml_cuda_op_mul_mat` broadcast + gather through a hub GPU (`ggml-cuda.cu:1451-1769`).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.distributed as dist

from ..ops import lib as _lib
from .comm import NcclComm

MAX_T = 4  # token rows per site in the fused path
TOKX_MAX_R = 64  # rows of the sharded arg-max exchange (csrc/kernels.cuh)


AR_EMPTY = 0x80000000  # bit pattern of -0.0f: "not here yet" (csrc/kernels.cuh)


class dlp_ar_in(C.Structure):
    _fields_ = [("buf", C.c_void_p), ("epoch", C.c_void_p), ("h_out", C.c_void_p),
                ("error", C.c_void_p), ("tp", C.c_int32), ("reserved", C.c_int32)]


class dlp_ar_out(C.Structure):
    _fields_ = [("peer_buf", C.c_void_p * 8), ("mc_buf", C.c_void_p), ("epoch", C.c_void_p), ("rank", C.c_int32), ("tp", C.c_int32)]


class dlp_tokx(C.Structure):
    _fields_ = [("peer_slots", C.c_void_p * 8), ("mc_slots", C.c_void_p), ("slots", C.c_void_p), ("epoch", C.c_void_p),
                ("error", C.c_void_p), ("rank", C.c_int32), ("tp", C.c_int32)]


class dlp_gemm_coll(C.Structure):
    _fields_ = [("x_flags", C.c_void_p), ("x_rows_per_src", C.c_int32), ("x_epoch", C.c_uint32), ("error", C.c_void_p),
                ("rs_peer_buf", C.c_void_p * 8), ("rs_peer_flags", C.c_void_p * 8), ("rs_done", C.c_void_p),
                ("rs_rows", C.c_int32), ("rs_rank", C.c_int32), ("rs_tp", C.c_int32), ("rs_epoch", C.c_uint32), ("step", C.c_void_p),
                ("ws", C.c_void_p), ("ws_bytes", C.c_size_t)]


class dlp_rs_norm_ag_desc(C.Structure):
    _fields_ = [("h_shard", C.c_void_p), ("red_buf", C.c_void_p), ("red_flags", C.c_void_p), ("red_epoch", C.c_uint32),
                ("pad0", C.c_uint32), ("norm_w", C.c_void_p), ("eps", C.c_float), ("pad1", C.c_uint32),
                ("ag_peer_buf", C.c_void_p * 8), ("ag_peer_flags", C.c_void_p * 8), ("ag_epoch", C.c_uint32), ("pad2", C.c_uint32),
                ("done", C.c_void_p), ("error", C.c_void_p), ("rank", C.c_int32), ("tp", C.c_int32), ("rows", C.c_int32),
                ("E", C.c_int32), ("step", C.c_void_p)]


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
        L.dlp_argmax_tp.restype = i32
        L.dlp_argmax_tp.argtypes = [vp, i32, i32, i32, i32, C.POINTER(dlp_tokx), vp, vp]
        L.dlp_fill_u32.restype = i32
        L.dlp_fill_u32.argtypes = [vp, C.c_size_t, C.c_uint32, vp]

    # ---- set-up: one symmetric allocation per rank, handles exchanged over the process group --------------
    def _layout(self, n_embd: int, n_sites: int):
        tp = self.tp_size
        self.E, self.n_sites = n_embd, n_sites
        self.site_stride = 2 * tp * MAX_T * n_embd * 4          # [2 parities][tp][MAX_T][E] fp32
        self.tok_off = n_sites * self.site_stride                # token exchange slots: uint64 [2][tp][MAX_T]
        self.total_bytes = self.tok_off + ((2 * tp * TOKX_MAX_R * 8 + 255) // 256) * 256

    def _alloc_symmetric(self, nbytes: int):
        """(local base, [peer bases], multicast base or 0).  torch's symmetric memory is tried first because it also binds an NVLS
        multicast object when the fabric has one; cudaMalloc + CUDA IPC otherwise (DLP_SYMM=ipc forces it)."""
        import os
        tp = self.tp_size
        if os.environ.get("DLP_SYMM", "auto") != "ipc":
            try:
                import torch.distributed._symmetric_memory as symm_mem
                t = symm_mem.empty(nbytes, dtype=torch.uint8, device=self.device)
                hdl = symm_mem.rendezvous(t, group=self.group)
                peers = [int(p) for p in hdl.buffer_ptrs]
                mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
                if len(peers) == tp and peers[self.tp_rank] == t.data_ptr():
                    self._symm_keep = getattr(self, "_symm_keep", []) + [(t, hdl)]
                    self.symm_backend = "symm_mem+nvls" if mc else "symm_mem"
                    if os.environ.get("DLP_SYMM_MC", "1") == "0":
                        mc = 0
                    return t.data_ptr(), peers, mc
            except Exception as e:  # noqa: BLE001 - any failure of the optional path falls back to IPC
                self.symm_note = f"{type(e).__name__}: {e}"[:200]
        p = C.c_void_p()
        _lib.check(self.lib.dlp_symm_alloc(nbytes, C.byref(p)), "symm_alloc")
        handle = (C.c_char * 64)()
        _lib.check(self.lib.dlp_ipc_export(p.value, handle), "ipc_export")
        gathered: list = [None] * tp
        dist.all_gather_object(gathered, bytes(handle), group=self.group)
        peers = []
        for r, h in enumerate(gathered):
            if r == self.tp_rank:
                peers.append(p.value)
                continue
            q = C.c_void_p()
            _lib.check(self.lib.dlp_ipc_import((C.c_char * 64).from_buffer_copy(h), C.byref(q)), "ipc_import")
            peers.append(q.value)
        self.symm_backend = "cuda-ipc"
        return p.value, peers, 0

    def setup(self, n_embd: int, n_sites: int):
        self._layout(n_embd, n_sites)
        if getattr(self, "emulated", False):
            self._mem = torch.zeros(self.total_bytes, dtype=torch.uint8, device=self.device)
            self.base, self.mc_base, self.symm_backend = self._mem.data_ptr(), 0, "emulated"
            self.peers = [self.base] * self.tp_size      # ar_out()/tokx() shift each "peer" to its own slot
            self._arm()
            self.ready = True
            return
        self.base, self.peers, self.mc_base = self._alloc_symmetric(self.total_bytes)
        self._arm()
        dist.barrier(group=self.group)
        self.ready = True

    def _arm(self):
        """Initial state: every all-reduce word "empty", token slots zero, epoch 0."""
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(self.lib.dlp_fill_u32(self.base, self.tok_off // 4, AR_EMPTY, st), "fill")
        _lib.check(self.lib.dlp_fill_u32(self.base + self.tok_off, (self.total_bytes - self.tok_off) // 4, 0, st), "fill")
        self.epoch = torch.zeros(1, dtype=torch.int32, device=self.device)       # all-reduce sites: bumped once per fused decode step
        self.tok_epoch = torch.zeros(1, dtype=torch.int32, device=self.device)   # token exchange: bumped before every sharded arg-max
        self.error = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._in, self._out, self._tokx = {}, {}, None
        torch.cuda.synchronize()

    @classmethod
    def make_virtual(cls, tp: int, n_embd: int, n_sites: int, device="cuda") -> list["FusedComm"]:
        """`tp` communicators that live on ONE device and point at each other's buffers directly - no IPC, no process
        group.  Used to test the store / spin / re-arm protocol on a single GPU: the caller launches each phase for all
        virtual ranks in turn, so every value is already there when a consumer looks for it (kernels that wait on each
        other must never be co-scheduled on one GPU)."""
        comms = []
        for r in range(tp):
            c = cls.__new__(cls)
            c.group, c.tp_size, c.tp_rank, c._gather_buf = None, tp, r, None
            c.grid, c.device, c.lib = None, torch.device(device), _lib.load()
            c._declare()
            c._layout(n_embd, n_sites)
            c._mem = torch.zeros(c.total_bytes, dtype=torch.uint8, device=device)
            c.base, c.mc_base, c.symm_backend = c._mem.data_ptr(), 0, "virtual"
            c._arm()
            c.ready = True
            comms.append(c)
        for c in comms:
            c.peers = [o.base for o in comms]
        return comms

    @classmethod
    def make_emulated(cls, tp: int, n_embd: int | None = None, n_sites: int | None = None, device="cuda") -> "FusedComm":
        """Rank 0 of a `tp`-way group on ONE device with nobody else: its stores into "rank r's buffer" land in slot r of its own
        buffer, so its consumers find tp (identical) partials and never wait.  The kernels, shard shapes, launch counts and
        per-rank bytes are exactly those of a real tp-way step; only the NVLink hop is missing.  For per-rank timing and
        profiling of the tensor-parallel decode step on a single GPU (bench.py --emulate-tp)."""
        c = cls.__new__(cls)
        c.group, c.tp_size, c.tp_rank, c._gather_buf = None, tp, 0, None
        c.grid, c.device, c.lib = None, torch.device(device), _lib.load()
        c._declare()
        c.ready, c.emulated, c.no_sp = False, True, True
        c.base, c.peers = None, []
        if n_embd is not None and n_sites is not None:
            c.setup(n_embd, n_sites)
        return c

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
            d = dlp_ar_in(self.base + site * self.site_stride, self.epoch.data_ptr(), h_out.data_ptr(), self.error.data_ptr(), self.tp_size, 0)
            self._in[key] = d
        return d

    def ar_out(self, site: int) -> dlp_ar_out:
        d = self._out.get(site)
        if d is None:
            d = dlp_ar_out()
            emu = getattr(self, "emulated", False)
            for r in range(self.tp_size):   # (emulation: "rank r" is slot r of our own buffer)
                d.peer_buf[r] = self.peers[r] + site * self.site_stride + (r * MAX_T * self.E * 4 if emu else 0)
            d.mc_buf = (self.mc_base + site * self.site_stride) if self.mc_base else None
            d.epoch = self.epoch.data_ptr()
            d.rank, d.tp = self.tp_rank, self.tp_size
            self._out[site] = d
        return d

    def tokx(self) -> dlp_tokx:
        if self._tokx is None:
            d = dlp_tokx()
            emu = getattr(self, "emulated", False)
            for r in range(self.tp_size):
                d.peer_slots[r] = self.peers[r] + self.tok_off + (r * TOKX_MAX_R * 8 if emu else 0)
            d.mc_slots = (self.mc_base + self.tok_off) if self.mc_base else None
            d.slots = self.base + self.tok_off
            d.epoch, d.error = self.tok_epoch.data_ptr(), self.error.data_ptr()   # (own counter: the parities of the all-reduce
            d.rank, d.tp = self.tp_rank, self.tp_size                             #  sites must only advance with steps that use them)
            self._tokx = d
        return self._tokx

    def argmax_sharded(self, logits: torch.Tensor, v0: int, out: torch.Tensor) -> torch.Tensor:
        """Global arg-max over vocabulary-sharded logits [B, V_local] (B <= 64): the same token id on every rank, in one launch."""
        from ..ops import api as ops
        B, V = logits.shape
        st = torch.cuda.current_stream().cuda_stream
        _lib.check(self.lib.dlp_bump(self.tok_epoch.data_ptr(), st), "bump")
        ops._count()
        _lib.check(self.lib.dlp_argmax_tp(logits.data_ptr(), B, V, logits.stride(0), v0, C.byref(self.tokx()), out.data_ptr(),
                                          torch.cuda.current_stream().cuda_stream), "argmax_tp")
        ops._count()
        return out

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
        self.g_epoch, self.r_epoch = [0, 0], [0, 0]      # positions inside the current forward pass (see begin_sp_step)
        self.sp_done = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.sp_step = torch.zeros(1, dtype=torch.int32, device=self.device)   # forward-pass counter, bumped on the device
        if not hasattr(self, "error"):
            self.error = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.sp_ready = True

    def begin_sp_step(self):
        """Start of a sequence-parallel forward pass (every rank, once per pass): the device-side pass counter advances and the
        host-side epoch positions restart, so the epochs a launch carries are constants of its place in the pass - a captured
        graph of the pass replays correctly at any later step."""
        from ..ops import api as ops
        _lib.check(self.lib.dlp_bump(self.sp_step.data_ptr(), torch.cuda.current_stream().cuda_stream), "bump")
        ops._count()
        self.g_epoch, self.r_epoch = [0, 0], [0, 0]

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
        d.step = self.sp_step.data_ptr()
        _lib.check(self.lib.dlp_rs_norm_ag(C.byref(d), torch.cuda.current_stream().cuda_stream), "rs_norm_ag")
        ops._count()

    def sp_gemm(self, w, x, ldx: int, T: int, rows: int, epilogue: int, y: torch.Tensor | None = None, w_up=None,
                x_from_g: int | None = None, rs_to: int | None = None, x_owner: int | None = None):
        """tcgen05 GEMM with the collective fused in: X from gather buffer `x_from_g` (waits per owner) and / or the
        reduce-scatter push epilogue into reduce buffer `rs_to` of the owner ranks.  `x` is a tensor or a device address."""
        from ..ops import api as ops
        c = dlp_gemm_coll()
        c.error = self.error.data_ptr()
        c.step = self.sp_step.data_ptr()
        ws = ops.gemm_workspace(self.device)     # split-K workspace: a narrow shard has too few 128-row tiles to fill the SMs
        if ws is not None:
            c.ws, c.ws_bytes = ws.data_ptr(), ws.numel()
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

    # ---- single-device emulation of rank 0 (make_emulated): the NCCL-style entry points without a process group ----
    def allreduce_delta(self, delta, il, which):
        if getattr(self, "emulated", False):
            return delta.mul_(self.tp_size)          # tp identical partials
        return super().allreduce_delta(delta, il, which)

    def allgather_logits(self, logits, vocab_range, vocab_size):
        if getattr(self, "emulated", False):
            return logits.repeat(1, self.tp_size)[:, :vocab_size]
        return super().allgather_logits(logits, vocab_range, vocab_size)

    def check_error(self):
        if not getattr(self, "ready", False):      # never set up (e.g. a MoE model: its engine uses the NCCL entry points)
            return
        if int(self.error.item()):
            raise RuntimeError("fused all-reduce: a peer did not signal within the timeout (ranks diverged or died)")
