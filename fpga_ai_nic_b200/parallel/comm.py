"""Communicator: one symmetric heap per GPU, mapped into every peer over CUDA IPC.

Role in the reference: OPAE buffer allocation + MPF address translation + the IKL ring routing
script (sw/mlp_mpi_example_f32.cpp:304-350, hw/cci_mpf_app_conf.vh, sw/setup_route.sh).  Here
torch.distributed is only the rendezvous (it carries 64-byte IPC handles once); all gradient
traffic afterwards is peer loads/stores issued by our own kernels over NVLink.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

from .. import _native as N
from ..ops import kernels as K

_DT_CODES = {torch.float32: "<f4", torch.bfloat16: "<u2", torch.uint8: "|u1", torch.int32: "<i4", torch.uint32: "<u4"}


class _CudaBlob:
    """Minimal __cuda_array_interface__ carrier so torch can view native memory without a copy."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        self._owner = owner


def tensor_from_ptr(ptr: int, shape, dtype: torch.dtype, owner=None, device=None) -> torch.Tensor:
    nbytes = int(torch.tensor([], dtype=dtype).element_size())
    for s in shape:
        nbytes *= int(s)
    dev = device if device is not None else torch.device("cuda", torch.cuda.current_device())
    raw = torch.as_tensor(_CudaBlob(ptr, nbytes, owner), device=dev)
    return raw.view(dtype).view(*shape)


from ..utils.numa import bind_to_gpu  # noqa: E402


class Communicator:
    def __init__(self, heap_bytes: int, rank: int | None = None, world_size: int | None = None, group=None):
        import torch.distributed as dist
        self.lib = N.cuda()
        if world_size is None:
            world_size = dist.get_world_size(group) if dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.rank, self.world_size = rank, world_size
        self.device = torch.device("cuda", torch.cuda.current_device())
        # one process per GPU: keep this rank's CPU threads (and the pinned staging buffers it allocates from now
        # on) on the NUMA node next to its GPU - eight host -> device input copies from one socket halve each other
        self.numa = bind_to_gpu(self.device.index) if world_size > 1 else {"bound": False, "reason": "single rank"}
        h = C.c_void_p()
        N.check(self.lib.ainic_comm_create(C.byref(h), rank, world_size, heap_bytes), "comm_create")
        self.handle = h
        if world_size > 1:
            buf = (C.c_uint8 * 64)()
            N.check(self.lib.ainic_comm_export(self.handle, buf), "comm_export")
            mine = bytes(buf)
            gathered: list = [None] * world_size
            dist.all_gather_object(gathered, mine, group=group)
            blob = b"".join(gathered)
            N.check(self.lib.ainic_comm_connect(self.handle, blob), "comm_connect")
            dist.barrier(group=group)
        self._closed = False

    # -- symmetric allocations -------------------------------------------------------------
    def alloc_ptr(self, nbytes: int, align: int = 4096) -> int:
        p = self.lib.ainic_comm_alloc(self.handle, nbytes, align)
        if not p:
            raise N.NativeError("symmetric heap exhausted: " + self.lib.ainic_last_error().decode())
        return p

    def empty(self, shape, dtype=torch.float32) -> torch.Tensor:
        """A tensor inside the symmetric heap.  Every rank must allocate in the same order."""
        if isinstance(shape, int):
            shape = (shape,)
        nbytes = torch.tensor([], dtype=dtype).element_size()
        for s in shape:
            nbytes *= int(s)
        return tensor_from_ptr(self.alloc_ptr(nbytes), shape, dtype, owner=self, device=self.device)

    def free(self, t) -> None:
        """Return a block obtained from ``empty`` / ``alloc_ptr`` to the heap (a tensor or a raw pointer).  Same call in
        the same order on every rank, and no kernel anywhere may still be using the block - with more than one rank
        that means: behind a barrier.  The tensor must not be used afterwards."""
        p = t.data_ptr() if isinstance(t, torch.Tensor) else int(t)
        N.check(self.lib.ainic_comm_free(self.handle, p), "comm_free")

    def trim(self, group=None) -> int:
        """Collective: give staging areas that a larger bucket outgrew back to the heap (they are only retired while
        a peer might still read them).  Synchronises the device and, with more than one rank, runs a barrier first."""
        torch.cuda.synchronize(self.device)
        if self.world_size > 1:
            import torch.distributed as dist
            dist.barrier(group=group)
        return int(self.lib.ainic_comm_release_retired(self.handle))

    def heap_stats(self) -> dict:
        out = (C.c_size_t * 4)()
        N.check(self.lib.ainic_comm_heap_stats(self.handle, out), "comm_heap_stats")
        return {"in_use": out[0], "high_water": out[1], "free_blocks": out[2], "capacity": out[3]}

    def peer_ptr(self, t: torch.Tensor, peer: int) -> int:
        return self.lib.ainic_comm_peer_ptr(self.handle, t.data_ptr(), peer)

    # -- fused all-reduce (+ SGD) -----------------------------------------------------------
    def allreduce_fused(self, grad: torch.Tensor, weight: torch.Tensor | None = None, lr: float = 0.0,
                        weight_bf16: torch.Tensor | None = None, grad_out: torch.Tensor | None = None,
                        codec: int = K.CODEC_BFP, bfp_mode: int = K.CLEAN, algo: int = K.ALGO_AUTO, max_ctas: int = 0,
                        stream=None, grad_compressed: torch.Tensor | None = None, grad_scattered: bool = False) -> None:
        """One fused launch: compress -> peer exchange -> fp32 reduce in rank order -> decompress -> SGD.
        ``grad_compressed``: this rank's gradient already as frames inside the symmetric heap (GEMM
        epilogue output); with ``grad_scattered`` it is this rank's inbox filled by every rank's
        GEMM -> reduce-scatter epilogue (ainic_gemm_scatter_t layout)."""
        a = N.AllreduceArgs(grad.data_ptr(), weight.data_ptr() if weight is not None else None,
                            weight_bf16.data_ptr() if weight_bf16 is not None else None,
                            grad_out.data_ptr() if grad_out is not None else None, grad.numel(), lr, codec, bfp_mode,
                            algo, max_ctas, grad_compressed.data_ptr() if grad_compressed is not None else None,
                            1 if grad_scattered else 0)
        N.check(self.lib.ainic_allreduce_fused(self.handle, C.byref(a), N.stream_ptr(stream)), "allreduce_fused")
        K._count()

    # -- NVLS (in-switch reduce + multicast) -------------------------------------------------------
    def enable_nvls(self, heap_bytes: int, group=None) -> bool:
        """Create the multicast-bound heap.  Collective.  Returns False (with ``nvls_error`` set) when the
        box does not expose NVLS; the P2P algorithms keep working either way."""
        import torch.distributed as dist
        self.nvls = None
        self.nvls_error = ""
        if self.world_size < 2:
            self.nvls_error = "single rank"
            return False
        sup = torch.tensor([int(self.lib.ainic_nvls_supported())], device=self.device)
        dist.all_reduce(sup, op=dist.ReduceOp.MIN, group=group)
        if int(sup.item()) == 0:
            self.nvls_error = "CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED is 0 (or the driver lacks cuMulticast*)"
            return False
        name = f"ainic_nvls_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}".encode()
        h = C.c_void_p()
        rc = self.lib.ainic_nvls_stage1(C.byref(h), self.rank, self.world_size, heap_bytes, name)
        ok = torch.tensor([1 if rc == 0 else 0], device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if rc != 0:
            self.nvls_error = self.lib.ainic_last_error().decode()
        if int(ok.item()) == 0:
            if rc == 0:
                self.lib.ainic_nvls_destroy(h)
                self.nvls_error = "a peer rank could not set up NVLS"
            return False
        for stage in (self.lib.ainic_nvls_stage2, self.lib.ainic_nvls_stage3):
            rc = stage(h)
            ok = torch.tensor([1 if rc == 0 else 0], device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)   # doubles as the barrier between stages
            if int(ok.item()) == 0:
                self.nvls_error = self.lib.ainic_last_error().decode() or "a peer rank failed"
                return False
        self.nvls = h
        return True

    def nvls_empty(self, shape, dtype=torch.float32) -> torch.Tensor:
        """A tensor inside the NVLS heap (same allocation order on every rank)."""
        if isinstance(shape, int):
            shape = (shape,)
        nbytes = torch.tensor([], dtype=dtype).element_size()
        for s in shape:
            nbytes *= int(s)
        p = self.lib.ainic_nvls_alloc(self.nvls, nbytes)
        if not p:
            raise N.NativeError("NVLS heap exhausted")
        return tensor_from_ptr(p, shape, dtype, owner=self, device=self.device)

    def allreduce_nvls(self, grad: torch.Tensor, weight: torch.Tensor | None = None, lr: float = 0.0,
                       weight_bf16: torch.Tensor | None = None, grad_out: torch.Tensor | None = None, max_ctas: int = 0,
                       stream=None) -> None:
        """In-switch all-reduce + SGD of a bucket that lives in the NVLS heap (``nvls_empty``).  ``grad`` is fp32, or
        bf16 (half the wire and HBM bytes: the switch accumulates in fp32 and returns / multicasts bf16 sums)."""
        codec = K.CODEC_BF16 if grad.dtype == torch.bfloat16 else K.CODEC_F32
        a = N.AllreduceArgs(grad.data_ptr(), weight.data_ptr() if weight is not None else None,
                            weight_bf16.data_ptr() if weight_bf16 is not None else None,
                            grad_out.data_ptr() if grad_out is not None else None, grad.numel(), lr, codec, 0,
                            K.ALGO_NVLS, max_ctas)
        N.check(self.lib.ainic_allreduce_nvls(self.handle, self.nvls, C.byref(a), N.stream_ptr(stream)), "allreduce_nvls")
        K._count()

    def counters(self) -> dict:
        out = (C.c_uint64 * 8)()
        N.check(self.lib.ainic_comm_counters(self.handle, out), "comm_counters")
        return {"busy_ns": out[0], "wait_ns": out[1], "launches": out[2], "timeouts": out[3], "last_ns": out[5],
                "to_barrier0_ns": out[6], "to_barrier1_ns": out[7]}

    def flow_trace(self, nchunks: int) -> dict:
        """Per-chunk time line of the last flow call (needs AINIC_FLOW_TRACE=1): microseconds since the first doorbell."""
        out = (C.c_uint64 * 160)()
        N.check(self.lib.ainic_comm_flow_trace(self.handle, out), "comm_flow_trace")
        names = ["doorbell", "reduce_go", "reduce_done", "gather_go", "gather_done"]
        t0 = min((out[k] for k in range(nchunks) if out[k]), default=0)
        return {n: [round((out[w * 32 + k] - t0) / 1e3, 1) if out[w * 32 + k] else None for k in range(nchunks)]
                for w, n in enumerate(names)}

    def close(self) -> None:
        if not self._closed:
            self._closed = True
            self.lib.ainic_comm_destroy(self.handle)

    def __del__(self):  # best effort; explicit close() preferred
        try:
            if not self._closed and os.environ.get("AINIC_NO_DEL_CLOSE") is None:
                self.close()
        except Exception:
            pass
