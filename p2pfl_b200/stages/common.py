"""Helpers shared by the stages."""

from __future__ import annotations

from typing import Any, Dict, List

from p2pfl_b200.management.logger import logger


def pack_model(protocol: Any, learner: Any, params: Any) -> Any:
    """Serialise ``params`` the way ``protocol`` wants it: a device payload for
    data planes that move device memory themselves, pickled bytes otherwise."""
    if getattr(protocol, "device_payloads", False) and hasattr(learner, "export_payload"):
        # params=None: the live model, by reference (peers pull it straight from the arena)
        return learner.export_payload(params if params is not None else learner.live_params())
    return learner.encode_parameters(params=params if params is not None else learner.export_model())


def broadcast_metrics(protocol: Any, state: Any, results: Dict[str, float], lazy: bool = False) -> None:
    """Share evaluation results with every node
    (reference: stages/base_node/train_stage.py:103-112).  ``lazy``: another broadcast
    follows shortly (the pre-vote) and the two may share one message per neighbour."""
    if not results:
        return
    flat: List[str] = []
    for k, v in results.items():
        flat += [str(k), str(v)]
    msg = protocol.build_msg("metrics", flat, round=state.round)
    send_async(protocol, msg, lazy)  # nobody waits on metrics


def announce_aggregated(protocol: Any, state: Any, held: List[str]) -> None:
    """Tell the round's trainers whose models this node holds (``models_aggregated``,
    reference: train_stage.py:83-87, add_model_command.py:84-91).  Flooded like in the reference
    when every trainer is a direct neighbour anyway; otherwise sent to the trainers only."""
    from p2pfl_b200.settings import Settings

    msg = protocol.build_msg("models_aggregated", held, round=state.round)
    peers = [n for n in state.train_set if n != state.addr]
    multicast = getattr(protocol, "multicast_async", None)
    if Settings.DIRECT_TRAINSET_STATUS and multicast is not None and peers:
        direct = protocol.get_neighbors(only_direct=True)
        if any(n not in direct for n in peers):
            multicast(msg, peers)
            return
    send_async(protocol, msg)


def send_async(protocol: Any, msg: Any, lazy: bool = False) -> None:
    """Fire-and-forget broadcast where the protocol has one (plain broadcast otherwise)."""
    fn = getattr(protocol, "broadcast_async", None)
    if fn is None:
        protocol.broadcast(msg)
    elif lazy:
        try:
            fn(msg, lazy=True)
        except TypeError:  # a protocol without lazy queueing
            fn(msg)
    else:
        fn(msg)


def evaluate_and_share(ctx: Any, lazy: bool = False) -> None:
    st = ctx.state
    if st.learner is None:
        return
    import time as _t

    logger.debug(st.addr, "evaluating")
    t0 = _t.time()
    results = st.learner.evaluate()
    ctx.timings["evaluate_learner"] = ctx.timings.get("evaluate_learner", 0.0) + _t.time() - t0
    logger.debug(st.addr, f"evaluated: {results}")
    broadcast_metrics(ctx.communication_protocol, st, results, lazy)
