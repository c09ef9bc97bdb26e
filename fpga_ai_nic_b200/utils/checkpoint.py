"""Checkpoint / resume.  The reference has none - its only persistent state is the on-card DDR
weight copy loaded once through the `flags` bit (hw/weight_update.sv:624-637) - so
this is an addition: flat fp32 master weights are what defines the model, exactly the buffers the
NIC interface exchanges, and they are saved as such (rank 0 writes; replicas are bit-identical)."""
from __future__ import annotations

import os
import tempfile

import torch


def save(trainer, path: str, step: int = 0, extra: dict | None = None) -> None:
    sd = trainer.state_dict()
    sd["step"] = step
    sd["extra"] = extra or {}
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    fd, tmp = tempfile.mkstemp(dir=d, suffix=".tmp")
    os.close(fd)
    torch.save(sd, tmp)
    os.replace(tmp, path)           # atomic: a crash never leaves a half-written checkpoint


def load(trainer, path: str) -> dict:
    # tensors, ints, lists and plain dicts only: no arbitrary unpickling of an untrusted file
    sd = torch.load(path, map_location="cpu", weights_only=True)
    trainer.load_state_dict(sd)
    return {"step": sd.get("step", 0), "extra": sd.get("extra", {})}
