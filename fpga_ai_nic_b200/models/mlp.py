"""The reference's MLP (sw/mlp_mpi_example_f32.cpp) on B200.

``MLPTrainer`` drives the native executor ``ainic_mlp_*`` (csrc/api/mlp.cu): tcgen05 GEMMs for the
fully-connected layers, a fused softmax/cross-entropy kernel, and one fused all-reduce + SGD kernel
per layer bucket, overlapped with the rest of the backward pass.  ``MLPConfig.reference()`` is the
reference run (10 x 2048 x 2048, 1792 samples per rank, fuse 0; sw/run.sh:16).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import torch

from .. import _native as N
from ..ops import kernels as K
from ..ops import oracle as O
from ..parallel.comm import Communicator, tensor_from_ptr


@dataclass
class MLPConfig:
    sizes: list = field(default_factory=lambda: [2048] * 11)   # C[0..L]
    mb: int = 1792                                             # samples per rank
    fuse: int = 0                                              # 0 none, 1 bias, 2 relu, 3 bias+relu
    lr: float = 0.1                                            # hw/weight_update.sv:439-444 (-0.1)
    loss_weight: float = 0.1                                   # sw/mlp_mpi_example_f32.cpp:250
    codec: str = "auto"                                        # auto: f32 on 1 rank, bfp otherwise
    bfp_mode: int = K.CLEAN
    algo: str = "auto"
    seed: int = 1                                              # libxsmm_rng_set_seed(1), :273
    init_scale: float = 1.0                                    # multiplies the reference init (wide layers need < 1)
    dtype: str = "bf16"                                        # "bf16": bf16 activations / weight shadows (kind::f16 MMA);
                                                               # "f32": the reference's all-fp32 data flow on kind::tf32 MMA
    bucket_mb: float | None = None                             # fp32 MiB of gradients per all-reduce call; 0 = per
                                                               # layer (reference, :754); None = 80 from 2 ranks, else 0

    @staticmethod
    def reference(nlayers: int = 10, width: int = 2048, mb: int = 1792) -> "MLPConfig":
        return MLPConfig(sizes=[width] * (nlayers + 1), mb=mb)

    @property
    def nlayers(self) -> int:
        return len(self.sizes) - 1

    def flops_per_step(self, global_mb: int) -> float:
        """The reference's accounting: 6*MB*C*K per layer, 4* for layer 0 (:794-798)."""
        f = 0.0
        for l in range(self.nlayers):
            f += (4.0 if l == 0 else 6.0) * global_mb * self.sizes[l] * self.sizes[l + 1]
        return f

    def param_count(self) -> int:
        return sum(self.sizes[l] * self.sizes[l + 1] for l in range(self.nlayers))

    def resolved_bucket_mb(self, world: int) -> float:
        if self.bucket_mb is not None:
            return float(self.bucket_mb)
        if os.environ.get("AINIC_BUCKET_MB"):
            return float(os.environ["AINIC_BUCKET_MB"])
        return 48.0 if world == 2 else 80.0 if world > 2 else 0.0      # same rule as csrc/api/mlp.cu

    def bucket_plan(self, world: int = 1) -> list:
        """(lowest layer, highest layer) of every fused all-reduce call of one step, in issue order - the
        same walk csrc/api/bucket_plan.h does: collect layers from the last one down until the bucket
        holds ``bucket_mb`` of fp32 gradients, or follow the explicit schedule AINIC_BUCKET_LAYERS
        ("4,3,2,1" = layers per bucket, big buckets first); layer 0 always closes the last bucket."""
        limit = int(self.resolved_bucket_mb(world) * (1 << 20))
        sched = [int(v) for v in os.environ.get("AINIC_BUCKET_LAYERS", "").replace(",", " ").split() if int(v) > 0]
        plan, top, pend, n_in, si = [], self.nlayers - 1, 0, 0, 0
        for l in range(self.nlayers - 1, -1, -1):
            pend += self.sizes[l] * self.sizes[l + 1] * 4
            n_in += 1
            if sched:
                close = n_in >= (sched[si] if si < len(sched) else 1)
            else:
                close = pend >= limit
            if l > 0 and not close:
                continue
            plan.append((l, top))
            top, pend, n_in, si = l - 1, 0, 0, si + 1
        return plan

    def heap_bytes(self, world: int = 8) -> int:
        p = self.param_count() + 1024 * self.nlayers
        biggest = max(self.sizes[l] * self.sizes[l + 1] for l in range(self.nlayers))
        # a bucket closes once it holds >= bucket_mb, so it is smaller than bucket_mb + the biggest layer
        bucket = int(self.resolved_bucket_mb(world) * (1 << 20) / 4)
        largest = max(sum(self.sizes[l] * self.sizes[l + 1] + 1024 for l in range(lo, hi + 1)) for lo, hi in self.bucket_plan(world))
        frames = (min(p, max(bucket + biggest, largest)) + 1023) // 512 + (biggest + 511) // 512
        # flow mode (one persistent all-reduce per step over every layer) stages the whole model, but needs
        # only half the generic per-frame staging (ainic_comm_reserve_flow)
        frames = max(frames, ((p + 511) // 512 + 1) // 2 + self.nlayers + 2)
        # fp32 weights + fp32 gradients, two compressed-gradient buffers per layer (GEMM epilogue output),
        # the communicator's staging for the largest bucket, slack for alignment
        return p * 8 + p * 3 + frames * (2 * 1024 + 2 * 2048) + (8 << 20) + self.nlayers * 32768


class MLPTrainer:
    def __init__(self, cfg: MLPConfig, comm: Communicator | None = None):
        self.cfg = cfg
        self.lib = N.cuda()
        if comm is None:
            import torch.distributed as dist
            world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
            comm = Communicator(cfg.heap_bytes(world))
        self.comm = comm
        self.world = self.comm.world_size
        codec = cfg.codec
        if codec == "auto":
            codec = "f32" if self.world == 1 else "bfp"   # nothing crosses a wire on one rank
        self.codec = K.CODEC_NAMES[codec]
        self.algo = K.ALGO_NAMES[cfg.algo]
        sizes = (C.c_int * len(cfg.sizes))(*cfg.sizes)
        h = C.c_void_p()
        # the bucket size is read at create time (AINIC_BUCKET_MB) so that the communicator's staging is
        # sized once; a later set_bucket_mb() may grow it, which abandons the old area inside the heap
        saved = os.environ.get("AINIC_BUCKET_MB")
        if cfg.bucket_mb is not None:
            os.environ["AINIC_BUCKET_MB"] = repr(float(cfg.bucket_mb))
        try:
            N.check(self.lib.ainic_mlp_create2(C.byref(h), self.comm.handle, cfg.nlayers, sizes, cfg.mb, cfg.fuse, cfg.lr,
                                               cfg.loss_weight, self.codec, cfg.bfp_mode, self.algo,
                                               {"bf16": 0, "f32": 1}[cfg.dtype]), "mlp_create")
        finally:
            if cfg.bucket_mb is not None:
                if saved is None:
                    os.environ.pop("AINIC_BUCKET_MB", None)
                else:
                    os.environ["AINIC_BUCKET_MB"] = saved
        self.handle = h
        self.device = self.comm.device
        self.act_dtype = torch.float32 if cfg.dtype == "f32" else torch.bfloat16
        self.loss_dev = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._views: dict = {}
        self._wver = 0        # bumped whenever the weights may have changed (forward_mx re-quantises its copies)
        self.init_weights()

    # -- buffers ------------------------------------------------------------------------------
    _WHAT = {"weight": (0, torch.float32), "grad": (1, torch.float32), "shadow": (2, torch.bfloat16),
             "bias": (3, torch.float32), "dbias": (4, torch.float32), "act": (5, torch.bfloat16),
             "dact": (6, torch.bfloat16), "mask": (7, torch.uint8)}

    def buffer(self, what: str, layer: int) -> torch.Tensor:
        key = (what, layer)
        if key not in self._views:
            code, dt = self._WHAT[what]
            if what in ("act", "dact"):
                dt = self.act_dtype
            ptr = self.lib.ainic_mlp_buffer(self.handle, code, layer)
            if not ptr:
                raise KeyError(key)
            s = self.cfg.sizes
            if what in ("weight", "grad", "shadow"):
                shape = (s[layer + 1], s[layer])
            elif what in ("bias", "dbias"):
                shape = (s[layer + 1],)
            elif what == "mask":
                shape = (self.cfg.mb, s[layer + 1])
            else:
                shape = (self.cfg.mb, s[layer])
            self._views[key] = tensor_from_ptr(ptr, shape, dt, owner=self, device=self.device)
        return self._views[key]

    def init_weights(self) -> None:
        """0.05 - U[0,1)/10 per buffer (sw/mlp_mpi_example_f32.cpp:182-189, :415-450).  Every rank
        uses the same seed, which is what rank 0's MPI_Bcast achieves in the reference (:463-470)."""
        for l in range(self.cfg.nlayers):
            w = self.buffer("weight", l)
            w.copy_(O.init_buf(w.numel(), seed=self.cfg.seed * 1000 + l).view_as(w) * self.cfg.init_scale)
            self.buffer("bias", l).copy_(O.init_buf(self.cfg.sizes[l + 1], seed=self.cfg.seed * 1000 + 500 + l))
        self.sync_shadows()

    def broadcast_parameters(self, src: int = 0) -> None:
        """Rank ``src``'s weights and biases replace everyone's - the reference's MPI_Bcast of
        fil_libxsmm / bias_libxsmm after rank 0 initialised them (sw/mlp_mpi_example_f32.cpp:463-470).
        Start-up plumbing, so it goes through torch.distributed; the shadows are rebuilt afterwards."""
        import torch.distributed as dist
        if self.world > 1:
            for l in range(self.cfg.nlayers):
                dist.broadcast(self.buffer("weight", l), src=src)
                dist.broadcast(self.buffer("bias", l), src=src)
        self.sync_shadows()

    def scatter_batch(self, x: torch.Tensor | None, y: torch.Tensor | None, src: int = 0):
        """Split a global minibatch held by rank ``src`` ([world*mb, C0] fp32/bf16 and [world*mb] int32) into
        per-rank shards, like the reference's MPI_Scatter of the activations (:453-460).  Returns this
        rank's (x bf16 [mb, C0], y int32 [mb]) on the trainer's device."""
        import torch.distributed as dist
        mb, c0 = self.cfg.mb, self.cfg.sizes[0]
        xs = torch.empty(mb, c0, dtype=torch.bfloat16, device=self.device)
        ys = torch.empty(mb, dtype=torch.int32, device=self.device)
        if self.world == 1:
            xs.copy_(x.to(self.device).reshape(mb, c0))
            ys.copy_(y.to(self.device))
            return xs, ys
        is_src = self.comm.rank == src
        xl = list(x.to(self.device).bfloat16().reshape(self.world, mb, c0).unbind(0)) if is_src else None
        yl = list(y.to(self.device).to(torch.int32).reshape(self.world, mb).unbind(0)) if is_src else None
        dist.scatter(xs, xl, src=src)
        dist.scatter(ys, yl, src=src)
        return xs, ys

    def sync_shadows(self) -> None:
        self._wver += 1
        N.check(self.lib.ainic_mlp_sync_shadows(self.handle, N.stream_ptr()), "mlp_sync_shadows")
        torch.cuda.current_stream().synchronize()

    # -- execution ------------------------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        assert x.dtype == self.act_dtype and x.is_cuda and x.is_contiguous()
        N.check(self.lib.ainic_mlp_forward(self.handle, x.data_ptr(), N.stream_ptr()), "mlp_forward")
        return self.buffer("act", self.cfg.nlayers)

    def forward_mx(self, x: torch.Tensor) -> torch.Tensor:
        """Inference forward pass on the block-scaled fp8 tensor-core kind (tcgen05.mma.kind::mxf8f6f4.block_scale):
        weights are quantised once per weight version (e4m3 + one ue8m0 scale per 32 inputs, from the bf16 shadows),
        activations by one quantiser launch per layer, accumulation is fp32, layer outputs are bf16.  The GEMM is
        1.3-1.6 x the bf16 one (profiles/README.md R2.4); training keeps bf16 operands.
        Returns the logits (bf16, a buffer reused by the next call).  Layer widths must be multiples of 128."""
        cfg = self.cfg
        assert cfg.dtype == "bf16" and x.dtype == torch.bfloat16 and x.is_cuda and x.is_contiguous()
        assert all(s % 128 == 0 for s in cfg.sizes), "MX forward: layer widths must be multiples of 128"
        st = self.__dict__.setdefault("_mx", {"version": None, "w": [], "act": {}, "out": {}})
        version = self._wver
        if st["version"] != version:          # (re)quantise the weights: training steps / loads change them
            self.wait()
            prev = st["w"] or [None] * cfg.nlayers
            st["w"] = [K.mx_quantize_rows(self.buffer("shadow", l), out=prev[l]) for l in range(cfg.nlayers)]
            st["version"] = version
        h = x
        for l in range(cfg.nlayers):
            rows = h.shape[0]
            key = (l, rows)
            aq = st["act"][key] = K.mx_quantize_rows(h, out=st["act"].get(key))
            if key not in st["out"]:
                st["out"][key] = torch.empty(rows, cfg.sizes[l + 1], dtype=torch.bfloat16, device=h.device)
            wq, wsf = st["w"][l]
            h = K.gemm_mxfp8(K.GEMM_NT, aq[0], aq[1], wq, wsf, out=st["out"][key])
            if cfg.fuse in (1, 3):
                h.add_(self.buffer("bias", l).to(h.dtype))
            if cfg.fuse >= 2:
                h.relu_()
        return h

    def step(self, x: torch.Tensor, labels: torch.Tensor, want_loss: bool = True,
             loss_out: torch.Tensor | None = None) -> torch.Tensor | None:
        """One training iteration (enqueue only).  Returns the device scalar holding this rank's
        summed cross-entropy (divide by mb for the mean); ``loss_out`` (fp32[1], device) replaces the
        trainer's own scalar, e.g. to keep several steps' losses in flight."""
        assert x.dtype == self.act_dtype and x.is_cuda and x.is_contiguous()
        assert labels.dtype == torch.int32 and labels.is_cuda
        dst = loss_out if loss_out is not None else self.loss_dev
        self._wver += 1
        N.check(self.lib.ainic_mlp_step(self.handle, x.data_ptr(), labels.data_ptr(),
                                        dst.data_ptr() if want_loss else None, N.stream_ptr()), "mlp_step")
        return dst if want_loss else None

    def set_bucket_mb(self, mb: float) -> None:
        """Gradient bucket size (fp32 MiB collected from the last layer down before one fused
        all-reduce is issued; 0 = one call per layer).  Collective: same value on every rank."""
        N.check(self.lib.ainic_mlp_set_bucket_bytes(self.handle, int(mb * (1 << 20))), "mlp_set_bucket_bytes")

    @property
    def bucket_mb(self) -> float:
        return self.lib.ainic_mlp_bucket_bytes(self.handle) / float(1 << 20)

    def wait(self) -> None:
        """Make the current stream wait for all outstanding weight updates (reference: wait(), :157-180)."""
        N.check(self.lib.ainic_mlp_wait_updates(self.handle, N.stream_ptr()), "mlp_wait_updates")

    # -- profiling / health ------------------------------------------------------------------------
    def profile(self, enable: bool | None = None) -> dict:
        """Per-phase device times (the reference's DETAILED_PROFILE).  ``profile(True)`` starts, ``profile(False)``
        stops, ``profile()`` reads: {'steps', 'fwd_fc', 'loss', 'bwdupd_fc', 'update_tail'} in ms (accumulated)."""
        out = (C.c_double * 4)()
        n = self.lib.ainic_mlp_profile(self.handle, -1 if enable is None else int(enable), out)
        return {"steps": int(n), "fwd_fc": out[0], "loss": out[1], "bwdupd_fc": out[2], "update_tail": out[3]}

    def check_health(self) -> dict:
        """Failure detection: raises if any fused all-reduce gave up waiting for a peer (the reference's
        wait() would spin forever).  Returns the communicator's counters otherwise."""
        ctr = self.comm.counters()
        if ctr["timeouts"]:
            raise N.NativeError(f"{ctr['timeouts']} flag waits timed out inside fused all-reduce kernels: a peer rank is "
                                "stuck or dead; weights on this rank are no longer trustworthy")
        return ctr

    @property
    def launches(self) -> int:
        return int(self.lib.ainic_mlp_launches(self.handle))

    # -- checkpoint / resume (the reference has none: SURVEY section 5) ---------------------------
    def state_dict(self) -> dict:
        self.wait()
        torch.cuda.synchronize()
        return {"sizes": list(self.cfg.sizes),
                "weights": [self.buffer("weight", l).cpu().clone() for l in range(self.cfg.nlayers)],
                "biases": [self.buffer("bias", l).cpu().clone() for l in range(self.cfg.nlayers)]}

    def load_state_dict(self, sd: dict) -> None:
        if list(sd["sizes"]) != list(self.cfg.sizes):
            raise ValueError(f"checkpoint is for layer sizes {list(sd['sizes'])}, trainer has {list(self.cfg.sizes)}")
        for l in range(self.cfg.nlayers):
            for name, t in (("weight", sd["weights"][l]), ("bias", sd["biases"][l])):
                dst = self.buffer(name, l)
                if tuple(t.shape) != tuple(dst.shape) or t.dtype != dst.dtype:
                    raise ValueError(f"checkpoint {name}[{l}] is {tuple(t.shape)} {t.dtype}, expected {tuple(dst.shape)} {dst.dtype}")
        self.wait()
        torch.cuda.synchronize()
        for l in range(self.cfg.nlayers):
            self.buffer("weight", l).copy_(sd["weights"][l])
            self.buffer("bias", l).copy_(sd["biases"][l])
        self.sync_shadows()

    def close(self, collective: bool = True) -> None:
        """Destroy the native trainer.  Its weight / gradient arenas go back to the communicator's heap, so another
        trainer can be built on the same communicator.  With more than one rank that needs every rank to be idle:
        ``collective=True`` (default) runs a barrier first and must then be called on every rank; with
        ``collective=False`` the arenas stay allocated until the communicator is destroyed."""
        if self.handle:
            free_heap = 1
            if self.world > 1:
                import torch.distributed as dist
                if collective and dist.is_available() and dist.is_initialized():
                    self.wait()
                    torch.cuda.synchronize(self.device)
                    dist.barrier()
                else:
                    free_heap = 0
            self._views.clear()
            self.lib.ainic_mlp_destroy_ex(self.handle, free_heap)
            self.handle = None
