"""Checkpoint / resume of a node's learning state.

The reference has none (``enable_checkpointing=False``, lightning_learner.py:188).
Because a learner's whole model is one flat vector, a checkpoint is just that vector plus the
optimizer moments, the round counters and the out-of-band integer buffers.

    save_checkpoint(node, "ckpt.pt")            # after / between experiments, or from a round listener
    load_checkpoint(node, "ckpt.pt")            # before set_start_learning: the node's model starts from it
"""

from __future__ import annotations

from typing import Any, Dict

import torch


def learner_state(learner: Any) -> Dict[str, Any]:
    st: Dict[str, Any] = {
        "layout_hash": learner.layout.hash,
        "flat": learner.flat.detach().to("cpu", torch.float32).clone(),
        "extra": {k: v.cpu().clone() for k, v in learner._extra().items()},
    }
    for name in ("adam_m", "adam_v", "adam_step"):
        t = getattr(learner, name, None)
        if t is not None:
            st[name] = t.detach().cpu().clone()
    return st


def save_checkpoint(node: Any, path: str) -> None:
    learner = node.state.learner
    if learner is None:
        raise RuntimeError("node has no learner yet (nothing to checkpoint)")
    torch.save(
        {
            "addr": node.addr,
            "round": node.state.round,
            "total_rounds": node.state.total_rounds,
            "learner": learner_state(learner),
        },
        path,
    )


def load_checkpoint(node: Any, path: str) -> Dict[str, Any]:
    """Restore the model weights into ``node.model`` and, when the node has a live fused learner,
    its Adam moments and step count as well.  Eager learners build a fresh optimizer per round (as
    the reference does, Q16), so for them a checkpoint is weights only.  The file holds tensors,
    strings and integers only and is read with ``weights_only=True`` (no code runs on load).  The
    round counters are returned to the caller, which decides how many rounds are left to run."""
    from p2pfl_b200.learning.params import FlatParams, ParamLayout

    ckpt = torch.load(path, map_location="cpu", weights_only=True)
    ls = ckpt["learner"]
    layout = ParamLayout.from_state_dict(node.model.state_dict())
    if layout.hash != ls["layout_hash"]:
        raise ValueError("checkpoint was written for a different model layout")
    params = FlatParams(ls["flat"], layout, ls["extra"])
    node.model.load_state_dict({k: v.clone() for k, v in params.state_dict().items()})
    learner = node.state.learner
    if learner is not None:
        learner.set_parameters(params)
        for name in ("adam_m", "adam_v", "adam_step"):
            if name in ls and getattr(learner, name, None) is not None:
                getattr(learner, name).copy_(ls[name].to(getattr(learner, name).device))
    return ckpt
