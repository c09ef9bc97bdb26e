"""Python face of the NIC-style command interface (csrc/api/nic.cpp).

Signatures follow the reference driver (sw/mlp_mpi_example_f32.cpp:65-180):
``setup`` = all_reduce_setup, ``all_reduce(buf, weight_out, flags)`` returns a request handle,
``wait(handle)`` spins on one of eight round-robin done slots, plus the counter getters and an
MPI-style ``iallreduce`` (sw/mlp_mpi_example_f32.cpp:615-620).

``Nic(backend="fake")`` needs no GPU: N fake NICs of one process are linked into a fabric and
compute with the CPU oracle (how the reference's withheld RTL testbench simulates three NICs in one
simulator, readme.pdf p.2-5).  ``Nic(backend="gpu", comm=...)`` launches the fused sm_100a kernels.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native as N
from ..ops import kernels as K

DONE_SLOTS = 8


class NicError(RuntimeError):
    def __init__(self, rc: int, what: str):
        super().__init__(f"{what}: {N.ERRORS.get(rc, rc)}")
        self.rc = rc


class Nic:
    def __init__(self, backend: str = "gpu", comm=None, rank: int = 0, nranks: int = 1):
        self.backend = backend
        self.lib = N.cuda() if backend == "gpu" else N.oracle()
        self.comm = comm
        h = C.c_void_p()
        rc = self.lib.ainic_nic_open(C.byref(h), comm.handle if comm is not None else None,
                                     0 if backend == "gpu" else 1, rank, nranks)
        if rc:
            raise NicError(rc, "nic_open")
        self.handle = h
        self.rank, self.nranks = rank, nranks

    @staticmethod
    def link(nics: list["Nic"]) -> None:
        """Connect fake NICs (one per rank-thread) into one fabric."""
        arr = (C.c_void_p * len(nics))(*[n.handle for n in nics])
        rc = nics[0].lib.ainic_nic_fake_link(arr, len(nics))
        if rc:
            raise NicError(rc, "fake_link")
        for i, n in enumerate(nics):
            n.rank, n.nranks = i, len(nics)

    def setup(self, count: int, codec: str = "bfp", bfp_mode: int = K.CLEAN, algo: str = "auto", lr: float = 0.1) -> None:
        rc = self.lib.ainic_nic_setup(self.handle, count, K.CODEC_NAMES[codec], bfp_mode, K.ALGO_NAMES[algo], lr)
        if rc:
            raise NicError(rc, "nic_setup")

    def done_word(self, slot: int) -> int:
        return int(self.lib.ainic_nic_done_buffer(self.handle)[slot * 16])

    def _stream(self, stream):
        return N.stream_ptr(stream) if self.backend == "gpu" else None

    def all_reduce(self, buf: torch.Tensor, weight_out: torch.Tensor | None, flags: int = 0, stream=None) -> N.Rqst:
        rq = N.Rqst()
        rc = self.lib.ainic_nic_all_reduce(self.handle, buf.data_ptr(), weight_out.data_ptr() if weight_out is not None else None,
                                           flags, buf.numel(), C.byref(rq), self._stream(stream))
        if rc:
            raise NicError(rc, "nic_all_reduce")
        if self.backend == "gpu":
            K._count()
        return rq

    def iallreduce(self, buf: torch.Tensor, stream=None) -> N.Rqst:
        rq = N.Rqst()
        rc = self.lib.ainic_iallreduce(self.handle, buf.data_ptr(), buf.numel(), C.byref(rq), self._stream(stream))
        if rc:
            raise NicError(rc, "iallreduce")
        if self.backend == "gpu":
            K._count()
        return rq

    def wait(self, rq: N.Rqst, timeout_s: float = 30.0) -> None:
        rc = self.lib.ainic_nic_wait(self.handle, C.byref(rq), timeout_s)
        if rc:
            raise NicError(rc, "nic_wait")

    def abort(self) -> None:
        self.lib.ainic_nic_abort(self.handle)

    def latency_ns(self) -> int:
        return int(self.lib.ainic_nic_latency_ns(self.handle))

    def stall_ns(self) -> int:
        return int(self.lib.ainic_nic_stall_ns(self.handle))

    def close(self) -> None:
        if self.handle:
            self.lib.ainic_nic_close(self.handle)
            self.handle = None
