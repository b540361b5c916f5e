"""Checkpoint / resume.  The reference has none on any live path; the format here is
``torch.save`` of ``{'round', 'state_dict' (reference keys, incl. BN buffers), 'mask'
(param-name -> tensor), 'args', RNG states, the training loaders' shuffle-generator states, 'stat_info', optional per-client
state_dicts}``."""
from __future__ import annotations

import os
import random
from typing import Dict, Optional

import numpy as np
import torch


def save_checkpoint(path: str, round_idx: int, w_global, mask, args, stat_info, personal: Optional[Dict] = None,
                    loaders: Optional[Dict] = None, extra: Optional[Dict] = None) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    ck = {
        "round": int(round_idx),
        "state_dict": {k: v.cpu() for k, v in w_global.state_dict().items()},
        "mask": ({n: mask.view(n).cpu().clone() for n in mask.layout.param_names} if mask is not None else None),
        "args": dict(vars(args)) if hasattr(args, "__dict__") else args,
        "rng": {"python": random.getstate(), "numpy": np.random.get_state(), "torch": torch.get_rng_state()},
        "stat_info": {k: v for k, v in stat_info.items() if k != "phase_time_s"},
        "loaders": ({int(c): st.cpu().clone() for c, st in loaders.items()} if loaders else None),
        "extra": extra or None,
        "personal": ({int(c): {k: v.cpu() for k, v in (s if isinstance(s, dict) else s.state_dict()).items()} for c, s in personal.items()}
                     if personal else None),
    }
    tmp = path + ".tmp"
    torch.save(ck, tmp)
    os.replace(tmp, path)


def load_checkpoint(path: str, map_location="cpu", restore_rng: bool = False) -> dict:
    ck = torch.load(path, map_location=map_location, weights_only=False)
    if restore_rng and "rng" in ck:
        random.setstate(ck["rng"]["python"])
        np.random.set_state(ck["rng"]["numpy"])
        torch.set_rng_state(ck["rng"]["torch"].cpu())
    return ck
