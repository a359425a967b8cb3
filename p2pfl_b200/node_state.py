"""Per-node mutable state shared by stages, commands and the logger.

Field names follow the reference's ``NodeState`` (p2pfl/node_state.py:26-115)
because tests and helpers read them directly (``state.round is None`` is the
"experiment finished" signal, p2pfl/utils.py:105).  The reference uses bare
``threading.Lock`` objects as latches; here they are ``Event``s plus one
``Condition`` (``changed``) that every mutation notifies, so stage code can
block on a predicate instead of sleeping a fixed period.
"""

from __future__ import annotations

import threading
from typing import Any, Callable, Dict, List, Optional
from p2pfl_b200.utils.lockcheck import make_condition, make_lock, make_rlock


class NodeState:
    """State record of one node."""

    def __init__(self, addr: str, simulation: bool = False) -> None:
        self.addr = addr
        self.simulation = simulation
        self.status = "Idle"
        self.actual_exp_name: Optional[str] = None
        self.round: Optional[int] = None
        self.total_rounds: Optional[int] = None
        self.learner: Any = None
        # base of the data-plane sequence numbers of this experiment (agreed through start_learning)
        self.seq_base: int = 0
        # round whose fused device exchange completed with every contributor (None otherwise)
        self.fused_complete_round: Optional[int] = None

        # who aggregated what this round: node -> list of contributors
        self.models_aggregated: Dict[str, List[str]] = {}
        # `models_aggregated` notices for the NEXT round that arrived before this node got there
        # (a fast trainer's notice travels point to point and can overtake a slow node's round change)
        self.early_aggregated: Dict[str, Any] = {}
        # last round for which a neighbour is known to hold the model (-1 = initialised)
        self.nei_status: Dict[str, int] = {}
        self.train_set: List[str] = []
        # votes are keyed by round so an early vote for round r+1 survives the
        # round switch (the reference keeps one flat dict and relies on timing)
        self.train_set_votes: Dict[int, Dict[str, Dict[str, int]]] = {}

        # payloads that arrived before this node could use them (learner not
        # built yet / train set not armed yet); replayed by the stages
        self.stash_lock = make_lock("state.stash")
        self.pending_init: Optional[Callable[[], None]] = None
        self.pending_models: List[Any] = []

        # observers of round boundaries: fn(rounds_completed), called on the learning thread
        self.round_listeners: List[Callable[[int], None]] = []

        # one condition for "something changed" - all waits are predicate waits
        self.changed = make_condition("state.changed")
        self.start_thread_lock = make_lock("state.start")
        # set once this node holds the experiment's initial weights
        self.model_initialized = threading.Event()
        # set when the experiment ended on this node (normally or by stop)
        self.finished = threading.Event()
        self.finished.set()

    # ------------------------------------------------------------------ events
    def notify(self) -> None:
        """Wake every thread blocked in :meth:`wait_for`."""
        with self.changed:
            self.changed.notify_all()

    def wait_for(self, predicate: Callable[[], bool], timeout: Optional[float]) -> bool:
        """Block until ``predicate()`` holds or ``timeout`` seconds elapse."""
        with self.changed:
            return self.changed.wait_for(predicate, timeout)

    # -------------------------------------------------------------- experiment
    def set_experiment(self, exp_name: str, total_rounds: int) -> None:
        """Begin an experiment (reference: node_state.py:83-95).  Unlike the
        reference every per-experiment field is reset here (Q19)."""
        with self.changed:
            self.status = "Learning"
            self.actual_exp_name = exp_name
            self.total_rounds = total_rounds
            self.round = 0
            self.models_aggregated = {}
            self.early_aggregated = {}
            self.train_set = []
            self.train_set_votes = {}
            self.finished.clear()
            self.changed.notify_all()

    def increase_round(self) -> None:
        """Advance one round (reference: node_state.py:97-108)."""
        if self.round is None:
            raise ValueError("round not initialised")
        # The per-round tables are also written by the message handlers (ballots and notices of
        # faster peers arrive at any time): both sides hold ``changed`` while they touch them, so the
        # round switch neither iterates a dictionary that grows under it nor loses a notice filed
        # between "copy the early ones" and "clear them".
        with self.changed:
            self.round += 1
            self.models_aggregated = {src: args for src, (rnd, args) in self.early_aggregated.items() if rnd == self.round}
            self.early_aggregated = {}
            # drop votes of rounds that are over
            for r in [r for r in self.train_set_votes if r < self.round]:
                del self.train_set_votes[r]
            self.changed.notify_all()
        for fn in list(self.round_listeners):
            fn(self.round)

    def clear(self) -> None:
        """Back to idle (reference: node_state.py:110-115, plus Q19 full reset)."""
        with self.changed:
            self.status = "Idle"
            self.actual_exp_name = None
            self.round = None
            self.total_rounds = None
            self.models_aggregated = {}
            self.early_aggregated = {}
            self.nei_status = {}
            self.train_set = []
            self.train_set_votes = {}
            self.pending_init = None
            self.pending_models = []
            self.model_initialized.clear()
            self.finished.set()
            self.changed.notify_all()

    def __repr__(self) -> str:
        return (
            f"NodeState(addr={self.addr}, status={self.status}, exp={self.actual_exp_name}, "
            f"round={self.round}/{self.total_rounds}, train_set={self.train_set})"
        )
