"""``Node``: the public façade (reference: p2pfl/node.py:58-378).

    node = Node(model, data, address="127.0.0.1", learner=..., aggregator=FedAvg, protocol=...)
    node.start(); node.connect(addr); node.set_start_learning(rounds=2, epochs=1)

Wires a communication protocol, an aggregator, the node state and the nine
command handlers together and runs the stage workflow on a learning thread.
A fresh ``LearningWorkflow`` is built per experiment (Q2).
"""

from __future__ import annotations

import threading
import time
from typing import Any, Dict, Optional, Type

from p2pfl_b200.commands import (
    AddModelCommand,
    InitModelCommand,
    MetricsCommand,
    ModelInitializedCommand,
    ModelsAggregatedCommand,
    ModelsReadyCommand,
    StartLearningCommand,
    StopLearningCommand,
    VoteTrainSetCommand,
)
from p2pfl_b200.exceptions import LearnerNotSetException, NodeRunningException, ZeroRoundsException
from p2pfl_b200.learning.aggregators.aggregator import Aggregator
from p2pfl_b200.learning.aggregators.fedavg import FedAvg
from p2pfl_b200.management.logger import logger
from p2pfl_b200.node_state import NodeState
from p2pfl_b200.settings import Settings
from p2pfl_b200.stages.stage import LearningContext
from p2pfl_b200.stages.workflows import LearningWorkflow


def _default_protocol() -> Type:
    from p2pfl_b200.communication.grpc.grpc_communication_protocol import GrpcCommunicationProtocol

    return GrpcCommunicationProtocol


def _default_learner() -> Type:
    from p2pfl_b200.learning.torch_learner import TorchLearner

    return TorchLearner


def _thread_uses_cuda() -> None:
    try:
        from p2pfl_b200 import ops

        ops.thread_uses_cuda()
    except Exception:
        pass


class Node:
    def __init__(
        self,
        model: Any,
        data: Any,
        address: str = "127.0.0.1",
        learner: Optional[Type] = None,
        aggregator: Type[Aggregator] = FedAvg,
        protocol: Optional[Type] = None,
        simulation: bool = False,
        **kwargs: Any,
    ) -> None:
        """``kwargs`` are forwarded to the learner constructor when it accepts
        them (e.g. ``device="cuda:3"``)."""
        protocol = protocol or _default_protocol()
        self._communication_protocol = protocol(address)
        self.addr = self._communication_protocol.get_address()

        self.data = data
        self.model = model
        self.learner_class = learner or _default_learner()
        self.learner_kwargs: Dict[str, Any] = dict(kwargs)
        self.aggregator = aggregator(node_name=self.addr)

        self.state = NodeState(self.addr, simulation=simulation)
        self.aggregator.add_listener(self.state.notify)
        if hasattr(self._communication_protocol, "set_neighbors_listener"):
            self._communication_protocol.set_neighbors_listener(self.state.notify)

        self.learning_workflow: Optional[LearningWorkflow] = None
        self._learning_thread: Optional[threading.Thread] = None
        self.__running = False
        self.last_timings: Dict[str, float] = {}
        self.live_timings: Dict[str, float] = {}

        self._communication_protocol.add_command(
            [
                StartLearningCommand(self.__start_learning_thread),
                StopLearningCommand(self.__stop_learning),
                ModelInitializedCommand(self.state),
                VoteTrainSetCommand(self.state),
                ModelsAggregatedCommand(self.state),
                ModelsReadyCommand(self.state),
                MetricsCommand(self.state),
                InitModelCommand(self.state, self.stop, self.aggregator, self._communication_protocol),
                AddModelCommand(self.state, self.stop, self.aggregator, self._communication_protocol),
            ]
        )
        if hasattr(self._communication_protocol, "bind_node"):
            self._communication_protocol.bind_node(self)

    # ------------------------------------------------------------ neighbours
    def connect(self, addr: str) -> bool:
        """Connect to ``addr`` (bidirectional through the handshake).  Adding
        nodes while learning is running is not fully supported, as in the
        reference (node.py:141-142)."""
        self.assert_running(True)
        if self.state.round is not None:
            logger.info(self.addr, "connecting while learning is running is not fully supported")
        logger.info(self.addr, f"connecting to {addr}")
        return self._communication_protocol.connect(addr)

    def get_neighbors(self, only_direct: bool = False) -> Dict[str, Any]:
        return self._communication_protocol.get_neighbors(only_direct)

    def disconnect(self, addr: str) -> None:
        self.assert_running(True)
        logger.info(self.addr, f"removing {addr}")
        self._communication_protocol.disconnect(addr, disconnect_msg=True)

    # ------------------------------------------------------------- lifecycle
    def assert_running(self, running: bool) -> None:
        if self.__running != running:
            raise NodeRunningException(f"node is {'not ' if running else ''}running")

    def is_running(self) -> bool:
        return self.__running

    def start(self, wait: bool = False) -> None:
        self.assert_running(False)
        if Settings.GIL_SWITCH_INTERVAL and Settings.GIL_SWITCH_INTERVAL > 0:
            import sys

            sys.setswitchinterval(Settings.GIL_SWITCH_INTERVAL)
        self.__running = True
        logger.register_node(self.addr, self.state, self.state.simulation)
        self._communication_protocol.start()
        logger.info(self.addr, "node started")
        if wait:
            self._communication_protocol.wait_for_termination()
            logger.info(self.addr, "server terminated")

    def stop(self) -> None:
        """Stop learning and networking; never raises (reference: node.py:227-246)."""
        logger.info(self.addr, "stopping node")
        try:
            self.__running = False
            self.__stop_learning()
            self._communication_protocol.stop()
            self.state.clear()
            logger.unregister_node(self.addr)
        except Exception as e:
            logger.debug(self.addr, f"error while stopping: {e!r}")

    # ------------------------------------------------------------- learning
    def set_data(self, data: Any) -> None:
        self.data = data
        if self.state.learner is not None:
            self.state.learner.set_data(data)

    def set_model(self, model: Any) -> None:
        self.model = model
        if self.state.learner is not None:
            self.state.learner.set_model(model)

    def set_start_learning(self, rounds: int = 1, epochs: int = 1) -> None:
        """Start an experiment on the whole network; this node's weights are the
        initial weights (reference: node.py:297-329)."""
        self.assert_running(True)
        if rounds < 1:
            raise ZeroRoundsException("rounds must be greater than 0")
        if self.state.round is not None:
            logger.info(self.addr, "learning already started")
            return
        if self.learner_class is None:
            raise LearnerNotSetException("learner not set")
        logger.info(self.addr, f"starting learning: {rounds} rounds x {epochs} epochs")
        proto = self._communication_protocol
        # data-plane sequence numbers must grow from one experiment to the next (arenas and
        # their arrival flags are reused): derive the base from the wall clock
        seq_base = (int(time.time()) % (1 << 18)) * 4096
        # the reference's handler takes exactly (rounds, epochs); the base rides in the message's
        # `round` field, which the reference passes along and ignores for this command
        proto.broadcast(proto.build_msg("start_learning", [str(rounds), str(epochs)], round=seq_base))
        self.state.model_initialized.set()
        proto.broadcast(proto.build_msg("model_initialized"))
        self.__start_learning_thread(rounds, epochs, seq_base)

    def set_stop_learning(self) -> None:
        if self.state.round is not None:
            proto = self._communication_protocol
            proto.broadcast(proto.build_msg("stop_learning"))
            self.__stop_learning()
        else:
            logger.info(self.addr, "learning is not running")

    def wait_learning_finished(self, timeout: Optional[float] = None) -> bool:
        """Event-based replacement for polling ``state.round is None``."""
        ok = self.state.finished.wait(timeout)
        t = self._learning_thread
        if ok and t is not None and t is not threading.current_thread():
            t.join(timeout)
        return ok

    # ---------------------------------------------------------- internals
    def __start_learning_thread(self, rounds: int, epochs: int, seq_base: int = 0) -> None:
        with self.state.start_thread_lock:
            if self.state.round is not None or not self.state.finished.is_set():
                return
            self.state.finished.clear()
            self.state.seq_base = seq_base
        t = threading.Thread(
            target=self.__start_learning, args=(rounds, epochs), name=f"learning_thread-{self.addr}", daemon=True
        )
        self._learning_thread = t
        t.start()

    def __start_learning(self, rounds: int, epochs: int) -> None:
        _thread_uses_cuda()
        workflow = LearningWorkflow()
        self.learning_workflow = workflow
        ctx = LearningContext(
            rounds=rounds,
            epochs=epochs,
            state=self.state,
            model=self.model,
            data=self.data,
            communication_protocol=self._communication_protocol,
            aggregator=self.aggregator,
            learner_class=self.learner_class,
            early_stopping_fn=lambda: self.state.round is None,
            learner_kwargs=self.learner_kwargs,
        )
        self.live_timings = ctx.timings  # the running experiment's stage timers (read-only for observers)
        try:
            workflow.run(ctx)
        except Exception:
            import traceback

            logger.error(self.addr, f"learning thread failed:\n{traceback.format_exc()}")
            self.stop()
        finally:
            self.last_timings = dict(ctx.timings)
            if self.state.round is not None:
                # workflow ended without closing the experiment (aborted start)
                self.state.clear()
            self.state.finished.set()
            self.state.notify()

    def __stop_learning(self) -> None:
        """Local interrupt (reference: node.py:366-378 + stop_learning_command.py:40-61)."""
        st = self.state
        if st.round is None and st.learner is None:
            return
        logger.info(self.addr, "stopping learning")
        learner = st.learner
        if learner is not None:
            try:
                learner.interrupt_fit()
            except Exception:
                pass
        self.aggregator.clear()
        st.clear()
        logger.experiment_finished(self.addr)
