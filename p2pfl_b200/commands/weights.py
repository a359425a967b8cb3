"""Handlers for model-carrying messages: ``init_model`` and ``add_model``.

A payload that cannot be decoded or does not match the local layout stops the
receiving node, as in the reference (init_model_command.py:104-114,
add_model_command.py:95-105) - the fault tests rely on it.
"""

from __future__ import annotations

from typing import Any, Callable, List, Optional

from p2pfl_b200.commands.command import Command
from p2pfl_b200.exceptions import DecodingParamsError, ModelNotMatchingError, PayloadUnavailableError
from p2pfl_b200.management.logger import logger
from p2pfl_b200.node_state import NodeState


class InitModelCommand(Command):
    """Initial weights from the experiment's initiator (or a neighbour that
    already has them)."""

    def __init__(self, state: NodeState, stop_fn: Callable[[], None], aggregator: Any, protocol: Any) -> None:
        self._state = state
        self._stop = stop_fn
        self._aggregator = aggregator
        self._protocol = protocol

    @staticmethod
    def get_name() -> str:
        return "init_model"

    def execute(
        self,
        source: str,
        round: int,
        *args: Any,
        weights: Any = None,
        contributors: Optional[List[str]] = None,
        weight: int = 1,
        **kwargs: Any,
    ) -> None:
        st = self._state
        if weights is None:
            raise ValueError("init_model without weights")
        if st.model_initialized.is_set():
            return
        with st.stash_lock:
            learner = st.learner if st.round is not None else None
            if learner is None:
                # the learner is still being built: keep the payload, StartLearningStage applies it
                st.pending_init = lambda: self.execute(
                    source, round, weights=weights, contributors=contributors, weight=weight
                )
                logger.debug(st.addr, "init_model stashed: learner not ready yet")
                return
        if st.model_initialized.is_set():
            logger.debug(st.addr, "init_model ignored: already initialised")
            return
        if round != st.round:
            logger.debug(st.addr, f"init_model for round {round} ignored (mine: {st.round})")
            return
        try:
            learner.set_parameters(_decode_from(learner, weights, source))
        except PayloadUnavailableError as e:
            # not a bad model: device memory this node cannot read (yet).  Dropped; the initiator's
            # gossip loop offers the weights again until this node announces ``model_initialized``.
            logger.warning(st.addr, f"initial model from {source} not readable yet ({e}); waiting for the next offer")
            return
        except (DecodingParamsError, ModelNotMatchingError) as e:
            logger.error(st.addr, f"initial model unusable ({e!r}); stopping node")
            self._stop()
            return
        except Exception as e:
            logger.error(st.addr, f"unexpected error applying the initial model ({e!r}); stopping node")
            self._stop()
            return
        st.model_initialized.set()
        st.notify()
        logger.info(st.addr, f"initial model received from {source}")
        self._protocol.broadcast(self._protocol.build_msg("model_initialized"))


def _decode_from(learner: Any, weights: Any, source: str) -> Any:
    """Decode a model payload; a device reference may only name the arena of the node that sent it
    (model messages are never relayed, so any other owner is a forged or confused reference)."""
    from p2pfl_b200.parallel.payload import DeviceRef

    if DeviceRef.is_ref(weights):
        owner = DeviceRef.from_bytes(weights).addr
        if owner != source:
            raise DecodingParamsError(f"device reference of {owner} arrived in a message from {source}")
    return learner.decode_parameters(weights)


class AddModelCommand(Command):
    """A (partial or full) aggregate for the current round."""

    def __init__(self, state: NodeState, stop_fn: Callable[[], None], aggregator: Any, protocol: Any) -> None:
        self._state = state
        self._stop = stop_fn
        self._aggregator = aggregator
        self._protocol = protocol

    @staticmethod
    def get_name() -> str:
        return "add_model"

    def execute(
        self,
        source: str,
        round: int,
        *args: Any,
        weights: Any = None,
        contributors: Optional[List[str]] = None,
        weight: int = 1,
        **kwargs: Any,
    ) -> None:
        st = self._state
        if weights is None or contributors is None:
            raise ValueError("add_model needs weights and contributors")
        learner = st.learner
        if st.round is None or learner is None:
            logger.debug(st.addr, "add_model ignored: not learning")
            return
        if round != st.round:
            logger.debug(st.addr, f"add_model from {source} for round {round} ignored (mine: {st.round})")
            return
        with st.stash_lock:
            if not self._aggregator.get_train_set():
                # vote still running here: keep the payload, Train/Wait stage replays it.  A device
                # reference is resolved NOW: the sender only keeps the memory it names stable until
                # this call returns, and may be training on it again by the time of the replay.
                from p2pfl_b200.parallel.payload import DeviceRef

                if DeviceRef.is_ref(weights):
                    try:
                        weights = _decode_from(learner, weights, source)
                    except Exception as e:
                        logger.error(st.addr, f"early model from {source} could not be fetched ({e!r}); dropped")
                        return
                st.pending_models.append(
                    (round, lambda: self.execute(source, round, weights=weights, contributors=contributors, weight=weight))
                )
                logger.debug(st.addr, f"add_model from {source} stashed: train set not armed yet")
                return
        if set(contributors) >= set(st.train_set) and st.train_set:
            # whoever sends the finished aggregate of this round holds it: no need to wait for its
            # ``models_ready`` (which travels on another thread and often arrives later) before
            # deciding that it needs nothing from this node - without this, the last node of a line
            # regularly sent the model straight back to the neighbour it had just got it from
            with st.changed:
                st.nei_status[source] = max(st.nei_status.get(source, -1), round)
                st.changed.notify_all()
        if not self._aggregator.would_accept(list(contributors)):
            logger.debug(st.addr, f"add_model from {source} not needed (contributors {contributors}); not decoded")
            return
        try:
            model = _decode_from(learner, weights, source)
            learner.check_parameters(model)
            held = self._aggregator.add_model(model, list(contributors), int(weight))
        except PayloadUnavailableError as e:
            logger.warning(st.addr, f"model from {source} not readable here ({e}); dropped, the sender will offer it again")
            return
        except (DecodingParamsError, ModelNotMatchingError) as e:
            logger.error(st.addr, f"received model unusable ({e!r}); stopping node")
            self._stop()
            return
        except Exception as e:
            logger.error(st.addr, f"unexpected error adding a model ({e!r}); stopping node")
            self._stop()
            return
        if held and st.addr in st.train_set:
            # who holds whose model only steers the exchange among the round's TRAINERS (TrainStage:
            # candidates are train-set members); a node that merely receives the finished aggregate
            # has nothing to tell them (the reference floods this notice from every receiver,
            # add_model_command.py:84-91: on a 64-node line with 16 trainers that is 48 useless floods
            # per round)
            from p2pfl_b200.stages.common import announce_aggregated

            announce_aggregated(self._protocol, st, held)
