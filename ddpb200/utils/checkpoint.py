"""state_dict checkpoints.

The reference never saves anything (no torch.save / state_dict call);
the layout to preserve is the implied one: 34 fp32 tensors under the keys of SURVEY.md §2.6
(``layers.<i>.weight`` ... ``fc1.bias``), prefixed ``module.`` under the DDP-style wrapper,
plus one ``momentum_buffer`` per parameter for SGD.  Tensors are written contiguous in the
standard (K,C,3,3) layout, so a checkpoint loads into the reference's stock ``VGG11()`` and
vice versa.
"""
from __future__ import annotations

import torch


def plain_state_dict(module: torch.nn.Module) -> dict:
    return {k: v.detach().to("cpu").clone(memory_format=torch.contiguous_format)
            for k, v in module.state_dict().items()}


def save_checkpoint(path, module: torch.nn.Module, optimizer=None, **extra) -> None:
    blob = {"model": plain_state_dict(module), **extra}
    if optimizer is not None:
        sd = optimizer.state_dict()
        sd["state"] = {k: {n: (t.detach().cpu() if torch.is_tensor(t) else t) for n, t in v.items()}
                       for k, v in sd["state"].items()}
        blob["optimizer"] = sd
    torch.save(blob, path)


def load_checkpoint(path, module: torch.nn.Module, optimizer=None) -> dict:
    blob = torch.load(path, map_location="cpu", weights_only=False)
    sd = blob["model"] if "model" in blob else blob
    own = module.state_dict()
    # accept checkpoints written with or without the DDP prefix
    if set(sd) != set(own):
        if all(k.startswith("module.") for k in sd):
            sd = {k[len("module."):]: v for k, v in sd.items()}
        elif all(("module." + k) in own for k in sd):
            sd = {"module." + k: v for k, v in sd.items()}
    with torch.no_grad():
        missing = set(own) - set(sd)
        unexpected = set(sd) - set(own)
        if missing or unexpected:
            raise KeyError(f"checkpoint mismatch: missing {sorted(missing)}, unexpected {sorted(unexpected)}")
        for k, dst in own.items():
            dst.copy_(sd[k])        # in place: keeps arena views intact
    eng = getattr(getattr(module, "module", module), "_engine", None)
    if eng is not None:
        eng.repack(full=True)
    if optimizer is not None and "optimizer" in blob:
        optimizer.load_state_dict(blob["optimizer"])
    return blob
