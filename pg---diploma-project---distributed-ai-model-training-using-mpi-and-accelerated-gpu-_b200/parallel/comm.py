"""Peer-memory communication layer.

The reference synchronises gradients with DDP's bucketed NCCL all-reduce and then runs
``torch.optim.Adam`` (cifar10_mpi_mobilenet_224.py:142-148, :179-180).  Here each rank owns one
symmetric device arena

    [ flag pad | abort word | reduction slots | gradients | master weights | optional bf16 weights | extra fp32 ]

that every peer on the node maps, and the kernels of ``csrc/comm_adam.cu`` work on it directly over
NVLink / NVSwitch: barrier on the flag pads, sum all ranks' gradients, average, Adam on the fp32
master weights, zero the gradients -- no NCCL call and no separate optimiser kernel.

Memory comes from the CUDA virtual-memory API (``csrc/vmm.cu``): allocations are exported as POSIX
file descriptors, which the ranks pass to each other over AF_UNIX sockets, and -- where the driver
offers it -- ONE multicast object is bound over all of them, which is what lets the kernels use
``multimem.ld_reduce`` (the NVSwitch adds the W copies) and ``multimem.st`` (the NVSwitch replicates a
store).  Without VMM the arena falls back to ``cudaMalloc`` + CUDA IPC and the peer load/store
kernels.  ``torch.distributed`` is used once, to exchange socket names and capability bits.
"""
from __future__ import annotations

import ctypes as C
import os
import socket
import threading
from typing import Dict, List, Optional

import torch

from ..ops import _lib
from ..ops._lib import COMM_MAX_BLOCKS, MAX_WORLD, REDUCE_MAX, CommAdamArgs, check
from ..optim import AdamHyper

_TYPESTR = {torch.float32: "<f4", torch.bfloat16: "<u2", torch.int32: "<i4", torch.uint8: "|u1",
            torch.float64: "<f8", torch.int64: "<i8"}


class _CudaArray:
    def __init__(self, owner, ptr: int, numel: int, typestr: str) -> None:
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}


def _view(owner, ptr: int, dtype: torch.dtype, numel: int) -> torch.Tensor:
    t = torch.as_tensor(_CudaArray(owner, ptr, numel, _TYPESTR[dtype]),
                        device=torch.device("cuda", torch.cuda.current_device()))
    return t.view(torch.bfloat16) if dtype == torch.bfloat16 else t


class DeviceArena:
    """A cudaMalloc'd region owned by the native library, viewable as torch tensors and exportable
    to peer processes through CUDA IPC (the fallback when VMM is unavailable)."""

    def __init__(self, nbytes: int) -> None:
        self.lib = _lib.require()
        self.nbytes = int((nbytes + 255) // 256 * 256)
        p = C.c_void_p()
        check(self.lib.b200dp_arena_alloc(self.nbytes, C.byref(p)))
        self.ptr = int(p.value)

    def export_handle(self) -> bytes:
        buf = C.create_string_buffer(64)
        check(self.lib.b200dp_ipc_export(self.ptr, buf))
        return bytes(buf.raw)

    def tensor(self, dtype: torch.dtype, numel: int, offset_bytes: int = 0) -> torch.Tensor:
        itemsize = torch.empty((), dtype=dtype).element_size()
        if offset_bytes % itemsize or offset_bytes + numel * itemsize > self.nbytes:
            raise ValueError("arena view out of range")
        return _view(self, self.ptr + offset_bytes, dtype, numel)

    def close(self) -> None:
        if self.ptr:
            self.lib.b200dp_arena_free(self.ptr)
            self.ptr = 0


def _exchange_fds(ctx, my_fds: List[int], tag: str) -> Dict[int, List[int]]:
    """Every rank hands `my_fds` to every peer (SCM_RIGHTS over abstract AF_UNIX sockets)."""
    import torch.distributed as dist
    world, rank = ctx.world_size, ctx.rank
    name = f"\0b200dp-{os.getpid()}-{tag}-{rank}"
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(name)
    srv.listen(world)

    def serve() -> None:
        for _ in range(world - 1):
            conn, _ = srv.accept()
            with conn:
                socket.send_fds(conn, [b"fd"], list(my_fds))

    th = threading.Thread(target=serve, daemon=True)
    th.start()
    names: List[Optional[str]] = [None] * world
    dist.all_gather_object(names, name)          # everybody is listening once this returns
    got: Dict[int, List[int]] = {}
    for r in range(world):
        if r == rank:
            continue
        with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
            c.connect(names[r])
            _, fds, _, _ = socket.recv_fds(c, 16, max(len(my_fds), 1) + 1)
            got[r] = list(fds)
    th.join(timeout=60)
    srv.close()
    return got


class SymmArena:
    """One arena per rank, mapped into every rank; optionally aliased by an NVSwitch multicast mapping.

    ``ptrs[r]`` is rank r's arena as mapped in THIS process (``ptrs[rank]`` is the local one),
    ``mc_ptr`` the multicast alias (0 when unavailable), ``kind`` one of ``local`` (world 1),
    ``ipc``, ``vmm``, ``vmm+nvls``.  ``B200DP_COMM_MEM=ipc|vmm`` forces a smaller capability.
    """

    def __init__(self, ctx, nbytes: int) -> None:
        self.ctx = ctx
        self.lib = _lib.require()
        self.world, self.rank = ctx.world_size, ctx.rank
        self.ptrs: List[int] = [0] * self.world
        self.mc_ptr = 0
        self._mc_handle = 0
        self._mem_handle = 0
        self._opened: List[int] = []
        self._ipc: Optional[DeviceArena] = None
        if self.world > MAX_WORLD:
            raise ValueError(f"world size {self.world} exceeds the peer-memory limit {MAX_WORLD}")
        if self.world == 1:
            self._ipc = DeviceArena(nbytes)
            self.nbytes = self._ipc.nbytes
            self.ptrs[0] = self._ipc.ptr
            self.kind = "local"
            return
        import torch.distributed as dist
        flags, gran = C.c_int(0), C.c_size_t(0)
        check(self.lib.b200dp_vmm_query(C.byref(flags), C.byref(gran)))
        force = os.environ.get("B200DP_COMM_MEM", "auto")
        mine = 0 if force == "ipc" else (flags.value & 1 if force == "vmm" else flags.value)
        caps: List[Optional[tuple]] = [None] * self.world
        dist.all_gather_object(caps, (mine, int(gran.value)))
        common = min(c[0] & 1 for c in caps), min((c[0] >> 1) & 1 for c in caps)
        gran_all = max(c[1] for c in caps)
        if not common[0]:
            self._init_ipc(nbytes)
            return
        self.nbytes = (nbytes + gran_all - 1) // gran_all * gran_all
        p, fd, h = C.c_void_p(), C.c_int(-1), C.c_ulonglong(0)
        check(self.lib.b200dp_vmm_alloc(self.nbytes, C.byref(p), C.byref(fd), C.byref(h)))
        self.ptrs[self.rank] = int(p.value)
        self._mem_handle = int(h.value)
        got = _exchange_fds(ctx, [fd.value], "mem")
        for r, fds in got.items():
            q = C.c_void_p()
            check(self.lib.b200dp_vmm_import(fds[0], self.nbytes, C.byref(q)))
            self.lib.b200dp_close_fd(fds[0])
            self.ptrs[r] = int(q.value)
            self._opened.append(int(q.value))
        self.lib.b200dp_close_fd(fd.value)
        self.kind = "vmm"
        if common[1]:
            self._init_multicast()
        dist.barrier(device_ids=[ctx.device.index])

    def _init_ipc(self, nbytes: int) -> None:
        import torch.distributed as dist
        self._ipc = DeviceArena(nbytes)
        self.nbytes = self._ipc.nbytes
        handles: List[Optional[bytes]] = [None] * self.world
        dist.all_gather_object(handles, self._ipc.export_handle())
        for r, h in enumerate(handles):
            if r == self.rank:
                self.ptrs[r] = self._ipc.ptr
                continue
            p = C.c_void_p()
            check(self.lib.b200dp_ipc_open(h, C.byref(p)))
            self.ptrs[r] = int(p.value)
            self._opened.append(int(p.value))
        self.kind = "ipc"
        dist.barrier(device_ids=[self.ctx.device.index])

    def _init_multicast(self) -> None:
        """Rank 0 creates the multicast object; everybody adds its device, binds its allocation, maps the alias."""
        import torch.distributed as dist
        mc = C.c_ulonglong(0)
        ok = 1
        if self.rank == 0:
            fd = C.c_int(-1)
            try:
                check(self.lib.b200dp_mc_create(self.world, self.nbytes, C.byref(mc), C.byref(fd)))
            except RuntimeError:
                ok = 0
            _exchange_fds(self.ctx, [fd.value] if ok else [], "mc")      # an empty list tells the peers it failed
            if ok:
                self.lib.b200dp_close_fd(fd.value)
        else:
            got = _exchange_fds(self.ctx, [], "mc")
            try:
                if not got[0]:
                    raise RuntimeError("rank 0 could not create the multicast object")
                check(self.lib.b200dp_mc_import(got[0][0], self.nbytes, C.byref(mc)))
            except RuntimeError:
                ok = 0
            for fds in got.values():
                for f in fds:
                    self.lib.b200dp_close_fd(f)
        if ok:
            try:
                check(self.lib.b200dp_mc_add_device(mc.value))
            except RuntimeError:
                ok = 0
        t = torch.tensor([ok], device=self.ctx.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)        # also the "every device has been added" barrier
        if int(t.item()) == 0:
            return
        try:
            check(self.lib.b200dp_mc_bind(mc.value, self._mem_handle, self.nbytes))
        except RuntimeError:
            ok = 0
        t = torch.tensor([ok], device=self.ctx.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)        # every rank has bound its memory
        if int(t.item()) == 0:
            return
        p = C.c_void_p()
        check(self.lib.b200dp_mc_map(mc.value, self.nbytes, C.byref(p)))
        self.mc_ptr = int(p.value)
        self._mc_handle = int(mc.value)
        self.kind = "vmm+nvls"

    def tensor(self, dtype: torch.dtype, numel: int, offset_bytes: int = 0) -> torch.Tensor:
        itemsize = torch.empty((), dtype=dtype).element_size()
        if offset_bytes % itemsize or offset_bytes + numel * itemsize > self.nbytes:
            raise ValueError("arena view out of range")
        return _view(self, self.ptrs[self.rank] + offset_bytes, dtype, numel)

    def close(self) -> None:
        """Unmap everything and free the local arena.  Tensor views made from it must be dead by now."""
        if self._ipc is not None:
            for p in self._opened:
                self.lib.b200dp_ipc_close(p)
            self._opened = []
            self._ipc.close()
            self._ipc = None
            return
        if self.mc_ptr:
            self.lib.b200dp_vmm_free(self.mc_ptr)
            self.mc_ptr = 0
        if self._mc_handle:
            self.lib.b200dp_mc_release(self._mc_handle)
            self._mc_handle = 0
        for p in self._opened:
            self.lib.b200dp_vmm_free(p)
        self._opened = []
        if self.ptrs and self.ptrs[self.rank]:
            self.lib.b200dp_vmm_free(self.ptrs[self.rank])
            self.ptrs[self.rank] = 0


from ..utils.watchdog import CommError  # noqa: E402  (a peer missed a device barrier; the failing launch changed nothing)


def _a256(n: int) -> int:
    return (n + 255) // 256 * 256


class FusedAllReduceAdam:
    """Owner of the per-rank comm arena and the device-resident step state.

    ``grads`` / ``params`` (and ``work``, ``extra``) are views of the arena, so peers can read the
    gradients and store updated weights; ``m`` / ``v`` are local fp32 vectors.  ``numel`` must be a
    multiple of 4.  ``algo``: ``one_shot``, ``two_shot``, ``nvls`` or ``auto`` (NVLS when the arena
    has a multicast alias and W >= 4, else two-shot for W >= 4, else one-shot).
    """

    FLAG_BYTES = COMM_MAX_BLOCKS * MAX_WORLD * 4
    ABORT_BYTES = 256            # one word per rank: set by any peer whose barrier timed out
    SLOT_BYTES = MAX_WORLD * REDUCE_MAX * 8

    def __init__(self, ctx, numel: int, hyper: Optional[AdamHyper] = None, *,
                 params: Optional[torch.Tensor] = None, m: Optional[torch.Tensor] = None,
                 v: Optional[torch.Tensor] = None, work_bf16: bool = False,
                 grad_dtype: torch.dtype = torch.float32, algo: str = "one_shot",
                 zero_grads: bool = True, max_blocks: int = 0, extra_numel: int = 0) -> None:
        if numel % 4:
            raise ValueError("numel must be a multiple of 4 (pad the flat vectors)")
        self.ctx = ctx
        self.lib = _lib.require()
        self.numel = numel
        self.h = hyper or AdamHyper()
        self.grad_dtype = grad_dtype
        gsize = 4 if grad_dtype == torch.float32 else 2
        dev = ctx.device
        off = self.FLAG_BYTES
        self._abort_off = off
        off += self.ABORT_BYTES
        self._slot_off = off
        off += self.SLOT_BYTES
        self._grad_off = off
        off += _a256(numel * gsize)
        self._param_off = off
        off += _a256(numel * 4)
        self._work_off = None
        if work_bf16:
            self._work_off = off
            off += _a256(numel * 2)
        self._extra_off = off
        off += _a256(max(extra_numel, 4) * 4)
        self.arena = SymmArena(ctx, off)
        if algo == "auto":
            # measured at W = 8, 9 MB: one-shot 128 us, NCCL all-reduce alone 72 us; two-shot moves 4x less data
            algo = ("nvls" if self.arena.mc_ptr else "two_shot") if ctx.world_size >= 4 else "one_shot"
        if algo == "nvls" and (not self.arena.mc_ptr or grad_dtype != torch.float32 or work_bf16):
            if ctx.world_size > 1:
                raise RuntimeError(f"algo 'nvls' needs a multicast arena (have '{self.arena.kind}'), fp32 gradients "
                                   "and no bf16 working copy")
        self.algo_name = algo
        self.algo = {"one_shot": 0, "two_shot": 1, "nvls": 2}[algo]
        self.grads = self.arena.tensor(grad_dtype, numel, self._grad_off)
        self.params = self.arena.tensor(torch.float32, numel, self._param_off)
        if params is not None:
            self.params.copy_(params)
        self.work = self.arena.tensor(torch.bfloat16, numel, self._work_off) if self._work_off else None
        self.extra = self.arena.tensor(torch.float32, extra_numel, self._extra_off) if extra_numel else None
        self.m = m if m is not None else torch.zeros(numel, dtype=torch.float32, device=dev)
        self.v = v if v is not None else torch.zeros(numel, dtype=torch.float32, device=dev)
        self.out_avg: Optional[torch.Tensor] = None
        self.lr_dev = torch.full((1,), float(self.h.lr), dtype=torch.float32, device=dev)
        self._lr_host = float(self.h.lr)
        self.step_state = torch.zeros(4, dtype=torch.int32, device=dev)
        self.epoch_state = torch.zeros(COMM_MAX_BLOCKS, dtype=torch.int32, device=dev)
        self.error_flag = torch.zeros(1, dtype=torch.int32, device=dev)
        # host-mapped mirror of the error flag: pinned memory is addressable from the device at the same address
        self.error_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self._red_in = torch.zeros(REDUCE_MAX, dtype=torch.float64, device=dev)
        self._red_out = torch.zeros(REDUCE_MAX, dtype=torch.float64, device=dev)
        self.args = CommAdamArgs()
        a = self.args
        a.world, a.rank, a.numel = ctx.world_size, ctx.rank, numel
        a.grad_dtype = 0 if grad_dtype == torch.float32 else 1
        a.algo = self.algo
        for r in range(ctx.world_size):
            base = self.arena.ptrs[r]
            a.flags[r] = base
            a.abort_words[r] = base + self._abort_off
            a.slots[r] = base + self._slot_off
            a.grads[r] = base + self._grad_off
            a.params[r] = base + self._param_off
            a.work_bf16[r] = (base + self._work_off) if self._work_off else None
        if self.arena.mc_ptr:
            a.mc_grads = self.arena.mc_ptr + self._grad_off
            a.mc_params = self.arena.mc_ptr + self._param_off
        a.m, a.v, a.lr = self.m.data_ptr(), self.v.data_ptr(), self.lr_dev.data_ptr()
        a.step_state, a.epoch_state = self.step_state.data_ptr(), self.epoch_state.data_ptr()
        a.error_flag = self.error_flag.data_ptr()
        a.error_host = self.error_host.data_ptr()
        a.beta1, a.beta2, a.eps, a.weight_decay = self.h.beta1, self.h.beta2, self.h.eps, self.h.weight_decay
        a.zero_grads = 1 if zero_grads else 0
        a.apply_adam = 1
        a.bump_step = 1
        a.out_avg = None
        a.spin_limit = int(os.environ.get("B200DP_COMM_SPIN_LIMIT", "0"))     # polls before a barrier gives up (0 = library default)
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

    def _broadcast(self, offset_bytes: int, numel: int, root: int) -> None:
        if self.ctx.world_size == 1:
            return
        W = self.ctx.world_size
        arr = (C.c_void_p * W)(*[self.arena.ptrs[r] + offset_bytes for r in range(W)])
        mc = (self.arena.mc_ptr + offset_bytes) if self.arena.mc_ptr else None
        check(self.lib.b200dp_comm_broadcast(C.byref(self.args), root, arr, mc, self.arena.ptrs[self.ctx.rank] + offset_bytes,
                                             (numel + 3) // 4 * 4, _lib.current_stream()))

    def broadcast_params(self, root: int = 0) -> None:
        """K-C3: the root's master weights overwrite everybody's (peer stores, or one multimem.st per 16 bytes)."""
        self._broadcast(self._param_off, self.numel, root)

    def broadcast_extra(self, root: int = 0) -> None:
        """K-C4: the root's `extra` region (BatchNorm running statistics) overwrites everybody's."""
        if self.extra is not None:
            self._broadcast(self._extra_off, self.extra.numel(), root)

    def reduce_f64(self, values: torch.Tensor, op: str = "sum") -> torch.Tensor:
        """All-reduce of up to 32 doubles over the slot tables (K-C6 / K-C7, max-over-ranks timers)."""
        k = values.numel()
        if k > REDUCE_MAX:
            raise ValueError(f"at most {REDUCE_MAX} values")
        self._red_in[:k].copy_(values.to(torch.float64).flatten())
        check(self.lib.b200dp_comm_reduce_f64(C.byref(self.args), self._red_in.data_ptr(), self._red_out.data_ptr(), k,
                                              {"sum": 0, "max": 1}[op], _lib.current_stream()))
        return self._red_out[:k].clone()

    def grid(self) -> int:
        return check(self.lib.b200dp_comm_adam_grid(C.byref(self.args)))

    def poll_error(self) -> None:
        """Host-side check of the mirrored error flag; no device synchronisation."""
        code = int(self.error_host[0])
        if code:
            raise CommError(f"rank {self.ctx.rank}: peer {code - 1} did not reach a device barrier in time "
                            "(dead or stalled rank); the failing launch left weights, moments and gradients untouched")

    def check_error(self) -> None:
        code = int(self.error_flag.item())
        if code:
            self.error_host[0] = code
        self.poll_error()

    # ---- state -----------------------------------------------------------------------------------
    @property
    def steps(self) -> int:
        return int(self.step_state[0].item())

    def state_dict(self) -> Dict[str, object]:
        """Full Adam state.  With the two-shot forms a rank only ever updates the moments of the slices it owns (the
        rest stays zero), so the full vectors are the element-wise sum over ranks."""
        m, v = self.m.detach().clone(), self.v.detach().clone()
        if self.algo >= 1 and self.ctx.initialized:
            import torch.distributed as dist
            dist.all_reduce(m)
            dist.all_reduce(v)
        return {"m": m.cpu(), "v": v.cpu(), "step": self.steps, "hyper": vars(self.h).copy(),
                "algo": self.algo_name, "world": self.ctx.world_size}

    def load_state_dict(self, sd: Dict[str, object]) -> None:
        # every rank keeps the full vectors; it only reads the slices it owns
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
        self.step_state.fill_(int(sd["step"]))

    def close(self) -> None:
        torch.cuda.synchronize()
        self.grads = self.params = self.work = self.extra = None     # drop the arena views before unmapping
        self.arena.close()
