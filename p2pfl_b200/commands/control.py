"""Handlers for the small control messages (SURVEY §2.6 catalogue).

One class per wire command, same command strings as the reference so a mixed
deployment interoperates: ``beat``, ``start_learning``, ``stop_learning``,
``model_initialized``, ``vote_train_set``, ``models_aggregated``,
``models_ready``, ``metrics``.
"""

from __future__ import annotations

from typing import Any, Callable

from p2pfl_b200.commands.command import Command
from p2pfl_b200.management.logger import logger
from p2pfl_b200.node_state import NodeState
from p2pfl_b200.settings import Settings


# value of a beat's `round` field that says "the originator speaks this framework's extensions";
# reference nodes ignore the field (their handler takes the time argument only)
EXTENDED_MARK = -2

# the commands a reference node understands; anything else is only sent to extended peers
REFERENCE_COMMANDS = frozenset(
    ["beat", "start_learning", "stop_learning", "model_initialized", "vote_train_set", "models_aggregated",
     "models_ready", "metrics", "init_model", "add_model"]
)  # fmt: skip


# Engine-level envelope for several control messages to one extended neighbour (never sent to a
# reference node; no counterpart in the reference): args = [pack_bundled(m), ...].
BUNDLE_CMD = "bundle"


def pack_bundled(msg: Any) -> str:
    import json

    # (the source is packed too: a relayed message keeps its originator, not the relayer)
    return json.dumps([msg.cmd, int(msg.round), int(msg.ttl), int(msg.hash), [str(a) for a in msg.args], msg.source])


def unpack_bundled(source: str, packed: str) -> Any:
    import json

    from p2pfl_b200.communication.messages import Message

    fields = json.loads(packed)
    cmd, rnd, ttl, h, args = fields[:5]
    origin = str(fields[5]) if len(fields) > 5 else source
    return Message(source=origin, cmd=str(cmd), args=[str(a) for a in args], round=int(rnd), ttl=int(ttl), hash=int(h))


class HeartbeatCommand(Command):
    """``beat [time]`` (reference: commands/heartbeat_command.py:39-52)."""

    def __init__(self, heartbeater: Any, on_extended: Any = None) -> None:
        self._hb = heartbeater
        self._on_extended = on_extended

    @staticmethod
    def get_name() -> str:
        return "beat"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        self._hb.beat(source, float(args[0]))
        if round == EXTENDED_MARK and self._on_extended is not None:
            self._on_extended(source)


class NeighborsHintCommand(Command):
    """``neighbors_hint [n1, n2, ...]``: the sender's direct neighbours (never relayed, only sent
    between nodes of this framework).  Lets a relay skip nodes the originator reaches itself."""

    def __init__(self, on_neighbors: Any) -> None:
        self._on_neighbors = on_neighbors

    @staticmethod
    def get_name() -> str:
        return "neighbors_hint"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        self._on_neighbors(source, list(args), int(round))


class StartLearningCommand(Command):
    """``start_learning [rounds, epochs]`` (reference: commands/start_learning_command.py:39-60)."""

    def __init__(self, start_learning_fn: Callable[..., None]) -> None:
        self._start = start_learning_fn

    @staticmethod
    def get_name() -> str:
        return "start_learning"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        # `round` carries the experiment's data-plane sequence base (-1 / absent from reference nodes)
        self._start(int(args[0]), int(args[1]), round if round is not None and round > 0 else 0)


class StopLearningCommand(Command):
    """``stop_learning`` (reference: commands/stop_learning_command.py:40-61);
    tolerates an idle node (Q22)."""

    def __init__(self, stop_learning_fn: Callable[[], None]) -> None:
        self._stop = stop_learning_fn

    @staticmethod
    def get_name() -> str:
        return "stop_learning"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        self._stop()


class ModelInitializedCommand(Command):
    """``model_initialized``: the sender holds the initial weights
    (reference: commands/model_initialized_command.py:48)."""

    def __init__(self, state: NodeState) -> None:
        self._state = state

    @staticmethod
    def get_name() -> str:
        return "model_initialized"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        st = self._state
        with st.changed:
            st.nei_status[source] = max(st.nei_status.get(source, -1), -1)
            st.changed.notify_all()


class VoteTrainSetCommand(Command):
    """``vote_train_set [cand, weight, ...]`` for ``round``
    (reference: commands/vote_train_set_command.py:54-67).  Votes are filed under
    their round, so an early vote for the next round is kept."""

    def __init__(self, state: NodeState) -> None:
        self._state = state

    @staticmethod
    def get_name() -> str:
        return "vote_train_set"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        st = self._state
        if len(args) % 2:
            raise ValueError("vote_train_set needs (candidate, weight) pairs")
        votes = {args[i]: int(args[i + 1]) for i in range(0, len(args), 2)}
        # ballots are filed under their round: early ones (a faster node pre-voting for a later
        # round) are kept until this node gets there; stale ones are dropped.  Under the state's
        # lock: the learning thread switches rounds (and prunes this table) concurrently.
        with st.changed:
            if st.round is None or not (st.round <= round <= st.round + 2 * max(1, Settings.PREVOTE_AHEAD) + 1):
                logger.debug(st.addr, f"vote from {source} for round {round} ignored (mine: {st.round})")
                return
            st.train_set_votes.setdefault(round, {})[source] = votes
            st.changed.notify_all()


class ModelsAggregatedCommand(Command):
    """``models_aggregated [contributors...]``: what the sender already holds
    this round (reference: commands/models_agregated_command.py:49-51)."""

    def __init__(self, state: NodeState) -> None:
        self._state = state

    @staticmethod
    def get_name() -> str:
        return "models_aggregated"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        st = self._state
        with st.changed:  # the round may switch between the comparison and the write otherwise
            if round == st.round:
                st.models_aggregated[source] = list(args)
                st.changed.notify_all()
            elif st.round is not None and round == st.round + 1:
                st.early_aggregated[source] = (round, list(args))  # installed when this node reaches that round


class ModelsReadyCommand(Command):
    """``models_ready``: the sender finished aggregation of ``round``
    (reference: commands/models_ready_command.py:52-54).  Records the message's
    round, monotonically (Q18)."""

    def __init__(self, state: NodeState) -> None:
        self._state = state

    @staticmethod
    def get_name() -> str:
        return "models_ready"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        st = self._state
        with st.changed:
            if st.round is None:
                return
            if round in (st.round - 1, st.round, st.round + 1):
                st.nei_status[source] = max(st.nei_status.get(source, -1), round)
                st.changed.notify_all()


class MetricsCommand(Command):
    """``metrics [name, value, ...]`` (reference: commands/metrics_command.py:50-55)."""

    def __init__(self, state: NodeState) -> None:
        self._state = state

    @staticmethod
    def get_name() -> str:
        return "metrics"

    def execute(self, source: str, round: int, *args: Any, **kwargs: Any) -> None:
        for i in range(0, len(args) - 1, 2):
            logger.log_metric(source, args[i], float(args[i + 1]), round=round if round >= 0 else None)
