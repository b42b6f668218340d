"""Peer-memory communication layer.

The reference synchronises gradients with DDP's bucketed NCCL all-reduce and then runs
``torch.optim.Adam`` (cifar10_mpi_mobilenet_224.py:142-148, :179-180).  Here each rank owns one
device arena (``[flag pad | gradients | optional weights]``) that every peer on the node maps
through CUDA IPC; ``FusedAllReduceAdam.step`` launches ONE kernel (csrc/comm_adam.cu) that
barriers on the flag pads, sums all ranks' gradients straight out of peer memory over NVLink in a
fixed rank order, averages, applies Adam to the fp32 master weights, optionally refreshes a bf16
working copy, and zeroes the gradients -- no NCCL call and no separate optimiser kernel.

``PeerGroup`` does the one-time handle exchange through ``torch.distributed`` (object all-gather);
nothing per step touches the host beyond the kernel launch.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch

from ..ops import _lib
from ..ops._lib import COMM_MAX_BLOCKS, MAX_WORLD, CommAdamArgs, check
from ..optim import AdamHyper


class DeviceArena:
    """A cudaMalloc'd region owned by the native library, viewable as torch tensors and
    exportable to peer processes."""

    def __init__(self, nbytes: int) -> None:
        self.lib = _lib.require()
        self.nbytes = int((nbytes + 255) // 256 * 256)
        p = C.c_void_p()
        check(self.lib.b200dp_arena_alloc(self.nbytes, C.byref(p)))
        self.ptr = int(p.value)
        self._views: List[torch.Tensor] = []

    def export_handle(self) -> bytes:
        buf = C.create_string_buffer(64)
        check(self.lib.b200dp_ipc_export(self.ptr, buf))
        return bytes(buf.raw)

    def tensor(self, dtype: torch.dtype, numel: int, offset_bytes: int = 0) -> torch.Tensor:
        itemsize = torch.empty((), dtype=dtype).element_size()
        if offset_bytes % itemsize or offset_bytes + numel * itemsize > self.nbytes:
            raise ValueError("arena view out of range")
        typestr = {torch.float32: "<f4", torch.bfloat16: "<u2", torch.int32: "<i4", torch.uint8: "|u1",
                   torch.float64: "<f8", torch.int64: "<i8"}[dtype]
        holder = _CudaArray(self, self.ptr + offset_bytes, numel, typestr)
        t = torch.as_tensor(holder, device=torch.device("cuda", torch.cuda.current_device()))
        if dtype == torch.bfloat16:
            t = t.view(torch.bfloat16)
        return t

    def close(self) -> None:
        if self.ptr:
            self.lib.b200dp_arena_free(self.ptr)
            self.ptr = 0


class _CudaArray:
    def __init__(self, owner: DeviceArena, ptr: int, numel: int, typestr: str) -> None:
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}


class PeerGroup:
    """Maps every rank's arena into this process (one node, NVLink/NVSwitch peers)."""

    def __init__(self, ctx, arena: DeviceArena) -> None:
        self.ctx = ctx
        self.arena = arena
        self.world = ctx.world_size
        self.rank = ctx.rank
        self.lib = arena.lib
        self.ptrs: List[int] = [0] * self.world
        self._opened: List[int] = []
        if self.world == 1:
            self.ptrs[0] = arena.ptr
            return
        if self.world > MAX_WORLD:
            raise ValueError(f"world size {self.world} exceeds the peer-memory limit {MAX_WORLD}")
        import torch.distributed as dist
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, arena.export_handle())
        for r, h in enumerate(handles):
            if r == self.rank:
                self.ptrs[r] = arena.ptr
                continue
            p = C.c_void_p()
            check(self.lib.b200dp_ipc_open(h, C.byref(p)))
            self.ptrs[r] = int(p.value)
            self._opened.append(int(p.value))
        dist.barrier(device_ids=[ctx.device.index])

    def close(self) -> None:
        for p in self._opened:
            self.lib.b200dp_ipc_close(p)
        self._opened = []


def _pad4(n: int) -> int:
    return (n + 3) // 4 * 4


class FusedAllReduceAdam:
    """Owner of the per-rank comm arena and the device-resident step state.

    ``params`` / ``m`` / ``v`` are local fp32 vectors of ``numel`` elements (a multiple of 4).
    With ``algo='two_shot'`` the master weights live inside the arena so peers can store into them
    (``param_buffer()``); with one-shot they stay wherever the engine keeps them.
    """

    FLAG_BYTES = COMM_MAX_BLOCKS * MAX_WORLD * 4

    def __init__(self, ctx, numel: int, hyper: Optional[AdamHyper] = None, *,
                 params: Optional[torch.Tensor] = None, m: Optional[torch.Tensor] = None,
                 v: Optional[torch.Tensor] = None, work_bf16: Optional[torch.Tensor] = None,
                 grad_dtype: torch.dtype = torch.float32, algo: str = "one_shot",
                 zero_grads: bool = True, max_blocks: int = 0) -> None:
        if numel % 4:
            raise ValueError("numel must be a multiple of 4 (pad the flat vectors)")
        self.ctx = ctx
        self.lib = _lib.require()
        self.numel = numel
        self.h = hyper or AdamHyper()
        self.algo = {"one_shot": 0, "two_shot": 1}[algo]
        self.grad_dtype = grad_dtype
        gsize = 4 if grad_dtype == torch.float32 else 2
        dev = ctx.device
        off = self.FLAG_BYTES
        self._grad_off = off
        off += (numel * gsize + 255) // 256 * 256
        self._param_off = self._work_off = None
        if self.algo == 1:
            self._param_off = off
            off += (numel * 4 + 255) // 256 * 256
            if work_bf16 is not None:
                self._work_off = off
                off += (numel * 2 + 255) // 256 * 256
        self.arena = DeviceArena(off)
        self.group = PeerGroup(ctx, self.arena)
        self.grads = self.arena.tensor(grad_dtype, numel, self._grad_off)
        if self.algo == 1:
            self.params = self.arena.tensor(torch.float32, numel, self._param_off)
            if params is not None:
                self.params.copy_(params)
            self.work = self.arena.tensor(torch.bfloat16, numel, self._work_off) if self._work_off else None
        else:
            self.params = params if params is not None else torch.zeros(numel, dtype=torch.float32, device=dev)
            self.work = work_bf16
        self.m = m if m is not None else torch.zeros(numel, dtype=torch.float32, device=dev)
        self.v = v if v is not None else torch.zeros(numel, dtype=torch.float32, device=dev)
        self.out_avg: Optional[torch.Tensor] = None
        self.lr_dev = torch.full((1,), float(self.h.lr), dtype=torch.float32, device=dev)
        self._lr_host = float(self.h.lr)
        self.step_state = torch.zeros(COMM_MAX_BLOCKS, dtype=torch.int32, device=dev)
        self.epoch_state = torch.zeros(COMM_MAX_BLOCKS, dtype=torch.int32, device=dev)
        self.error_flag = torch.zeros(1, dtype=torch.int32, device=dev)
        self.args = CommAdamArgs()
        a = self.args
        a.world, a.rank, a.numel = ctx.world_size, ctx.rank, numel
        a.grad_dtype = 0 if grad_dtype == torch.float32 else 1
        a.algo = self.algo
        for r in range(ctx.world_size):
            base = self.group.ptrs[r]
            a.flags[r] = base
            a.grads[r] = base + self._grad_off
            if self.algo == 1:
                a.params[r] = base + self._param_off
                a.work_bf16[r] = (base + self._work_off) if self._work_off else None
        if self.algo == 0:
            a.params[ctx.rank] = self.params.data_ptr()
            a.work_bf16[ctx.rank] = self.work.data_ptr() if self.work is not None else None
        a.m, a.v, a.lr = self.m.data_ptr(), self.v.data_ptr(), self.lr_dev.data_ptr()
        a.step_state, a.epoch_state = self.step_state.data_ptr(), self.epoch_state.data_ptr()
        a.error_flag = self.error_flag.data_ptr()
        a.beta1, a.beta2, a.eps, a.weight_decay = self.h.beta1, self.h.beta2, self.h.eps, self.h.weight_decay
        a.zero_grads = 1 if zero_grads else 0
        a.apply_adam = 1
        a.out_avg = None
        a.spin_limit = 0
        a.max_blocks = max_blocks
        torch.cuda.synchronize()

    # ---- buffers -------------------------------------------------------------------------------
    def grad_buffer(self, _parity: int = 0) -> torch.Tensor:
        return self.grads

    def current_grad_buffer(self) -> torch.Tensor:
        return self.grads

    def param_buffer(self) -> torch.Tensor:
        return self.params

    # ---- per-step ------------------------------------------------------------------------------
    def set_lr(self, lr: float) -> None:
        if lr != self._lr_host:
            self.lr_dev.fill_(lr)
            self._lr_host = lr

    def step(self, lr: Optional[float] = None) -> None:
        """all-reduce(grads)/W -> Adam(params, m, v) [-> bf16 work] -> grads = 0, on the current stream."""
        if lr is not None:
            self.set_lr(lr)
        self.args.apply_adam = 1
        check(self.lib.b200dp_comm_adam_step(C.byref(self.args), _lib.current_stream()))

    def all_reduce_mean(self, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Plain averaged all-reduce of the gradient buffer into a local fp32 vector (no Adam)."""
        if out is None:
            if self.out_avg is None:
                self.out_avg = torch.empty(self.numel, dtype=torch.float32, device=self.ctx.device)
            out = self.out_avg
        self.args.apply_adam = 0
        self.args.out_avg = out.data_ptr()
        algo = self.args.algo
        self.args.algo = 0
        try:
            check(self.lib.b200dp_comm_adam_step(C.byref(self.args), _lib.current_stream()))
        finally:
            self.args.apply_adam = 1
            self.args.algo = algo
        return out

    def barrier(self) -> None:
        check(self.lib.b200dp_comm_barrier(C.byref(self.args), _lib.current_stream()))

    def broadcast(self, tensors_by_rank_ptr: List[int], src: torch.Tensor, root: int = 0) -> None:
        arr = (C.c_void_p * self.ctx.world_size)(*tensors_by_rank_ptr)
        check(self.lib.b200dp_comm_broadcast(C.byref(self.args), root, arr, src.data_ptr(), src.numel(),
                                             _lib.current_stream()))

    def grid(self) -> int:
        return check(self.lib.b200dp_comm_adam_grid(C.byref(self.args)))

    def check_error(self) -> None:
        from ..utils.watchdog import check_device_flag
        check_device_flag(int(self.error_flag.item()), self.ctx.rank)

    # ---- state -----------------------------------------------------------------------------------
    @property
    def steps(self) -> int:
        return int(self.step_state[0].item())

    def state_dict(self) -> Dict[str, object]:
        return {"m": self.m.detach().cpu(), "v": self.v.detach().cpu(), "step": self.steps,
                "hyper": vars(self.h).copy()}

    def load_state_dict(self, sd: Dict[str, object]) -> None:
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
        self.step_state.fill_(int(sd["step"]))

    def close(self) -> None:
        torch.cuda.synchronize()
        self.group.close()
        # the arena itself is freed when the last tensor view dies with the process; freeing it here
        # while torch tensors still alias it would leave dangling views
