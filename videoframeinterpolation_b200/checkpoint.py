"""Checkpoint compatibility with the reference trainer.

The reference writes ``{'model': state_dict, 'best_psnr', 'step', 'epoch', 'optimizer': ...}`` (Trainer.py:43-54) where
``state_dict`` comes from the DDP-wrapped model when training was distributed, i.e. every key carries a ``module.``
prefix (Trainer.py:17-21), and it reads the optimiser back under the key ``'optim'`` (Trainer.py:26-29), which its own
``save_model`` never writes.  ``load_reference_checkpoint`` accepts all of these spellings; parameter names are the
reference's own (videoframeinterpolation_b200.dat mirrors them), so no key remapping is needed beyond the prefix.
"""
from __future__ import annotations

from typing import Any, Dict, Mapping, Optional, Tuple

import torch


def strip_ddp_prefix(sd: Mapping[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Remove DistributedDataParallel's ``module.`` prefix (only when every key carries it)."""
    keys = list(sd.keys())
    if keys and all(k.startswith("module.") for k in keys):
        return {k[len("module."):]: v for k, v in sd.items()}
    return dict(sd)


def split_checkpoint(chkpt: Mapping[str, Any]) -> Tuple[Dict[str, torch.Tensor], Optional[dict], dict]:
    """(model state_dict without DDP prefix, optimiser state or None, remaining metadata)."""
    if "model" in chkpt and isinstance(chkpt["model"], Mapping):
        sd = chkpt["model"]
        optim = chkpt.get("optim", chkpt.get("optimizer"))       # the reference saves 'optimizer' but loads 'optim'
        meta = {k: v for k, v in chkpt.items() if k not in ("model", "optim", "optimizer")}
    else:                                                          # a bare state_dict
        sd, optim, meta = chkpt, None, {}
    return strip_ddp_prefix(sd), optim, meta


def load_reference_checkpoint(model: torch.nn.Module, path_or_dict, strict: bool = True, map_location="cpu") -> dict:
    """Load a reference ``.pth`` (or an already loaded dict) into a vfi_b200 model; returns the metadata
    (``best_psnr``, ``step``, ``epoch``) plus ``'optimizer'`` when the file carried one."""
    chkpt = torch.load(path_or_dict, map_location=map_location, weights_only=False) if isinstance(path_or_dict, (str, bytes)) or hasattr(path_or_dict, "read") else path_or_dict
    sd, optim, meta = split_checkpoint(chkpt)
    model.load_state_dict(sd, strict=strict)
    if optim is not None:
        meta["optimizer"] = optim
    return meta


def save_reference_checkpoint(model: torch.nn.Module, path, epoch: int = 0, step: int = 0, best_psnr: float = 0.0,
                              optimizer: Optional[torch.optim.Optimizer] = None, ddp_prefix: bool = False) -> None:
    """Write the file layout of Trainer.save_model (Trainer.py:43-54); ``ddp_prefix`` reproduces a DDP-trained file."""
    sd = model.state_dict()
    if ddp_prefix:
        sd = {"module." + k: v for k, v in sd.items()}
    chkpt = {"model": sd, "best_psnr": best_psnr, "step": step, "epoch": epoch}
    if optimizer is not None:
        chkpt["optimizer"] = optimizer.state_dict()
    torch.save(chkpt, path)
