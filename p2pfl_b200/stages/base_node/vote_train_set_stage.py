"""Stage 2: elect this round's train set
(reference: p2pfl/stages/base_node/vote_train_set_stage.py:43-178).

Every node samples up to ``TRAIN_SET_SIZE`` candidates among all nodes it
knows (direct and indirect) plus itself, gives them decreasing random weights,
and floods its ballot.  When the ballots of all known nodes are in (or
``VOTE_TIMEOUT`` passes) the tallies are summed; the top ``TRAIN_SET_SIZE``
win, ties broken by name, so every node that saw the same ballots computes the
same set.  The wait is a predicate wait woken by ``VoteTrainSetCommand``, not
the reference's 2-second polling.
"""

from __future__ import annotations

import math
import random
import time
from typing import Dict, List, Optional, Type

from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings
from p2pfl_b200.stages.stage import LearningContext, Stage
from p2pfl_b200.stages.stage_factory import StageFactory


def tally_votes(ballots: Dict[str, Dict[str, int]], k: int) -> List[str]:
    """Sum ballots and return the ``k`` winners (votes desc, then name desc -
    the order the reference's two stable sorts produce, vote_train_set_stage.py:141-162)."""
    totals: Dict[str, int] = {}
    for ballot in ballots.values():
        for cand, w in ballot.items():
            totals[cand] = totals.get(cand, 0) + int(w)
    ranked = sorted(totals.items(), key=lambda kv: kv[0], reverse=True)
    ranked = sorted(ranked, key=lambda kv: kv[1], reverse=True)
    return [c for c, _ in ranked[:k]]


class VoteTrainSetStage(Stage):
    @staticmethod
    def name() -> str:
        return "VoteTrainSetStage"

    @staticmethod
    def execute(ctx: LearningContext) -> Optional[Type[Stage]]:
        st = ctx.state
        t0 = time.time()
        try:
            VoteTrainSetStage._vote(ctx)
            train_set = VoteTrainSetStage._aggregate_votes(ctx)
            if ctx.early_stopping_fn():
                return None
            st.train_set = VoteTrainSetStage._validate(ctx, train_set)
            st.notify()
            logger.info(st.addr, f"train set of {len(st.train_set)} nodes: {st.train_set}")
        finally:
            ctx.timings["vote"] = ctx.timings.get("vote", 0.0) + time.time() - t0
        if st.addr in st.train_set:
            return StageFactory.get_stage("TrainStage")
        return StageFactory.get_stage("WaitAggregatedModelsStage")

    @staticmethod
    def _vote(ctx: LearningContext) -> None:
        VoteTrainSetStage.cast_ballot(ctx, ctx.state.round)

    @staticmethod
    def cast_ballot(ctx: LearningContext, rnd: int, background: bool = False) -> None:
        """Sample candidates, weight them, record and broadcast the ballot for round ``rnd``
        (a no-op when this node already voted for that round, e.g. by pre-voting)."""
        st, proto = ctx.state, ctx.communication_protocol
        if st.addr in st.train_set_votes.get(rnd, {}):
            return
        candidates = list(proto.get_neighbors(only_direct=False).keys())
        if st.addr not in candidates:
            candidates.append(st.addr)
        k = min(Settings.TRAIN_SET_SIZE, len(candidates))
        picked = random.sample(candidates, k)
        ballot = {c: math.floor(random.randint(0, 1000) / (i + 1)) for i, c in enumerate(picked)}
        with st.changed:
            st.train_set_votes.setdefault(rnd, {})[st.addr] = ballot
        flat: List[str] = []
        for c, w in ballot.items():
            flat += [c, str(w)]
        logger.debug(st.addr, f"voting for round {rnd}: {ballot}")
        msg = proto.build_msg("vote_train_set", flat, round=rnd)
        if background:  # a pre-vote is not urgent: do not hold the learning thread for N-1 round trips
            getattr(proto, "broadcast_async", proto.broadcast)(msg)
        else:
            proto.broadcast(msg)

    @staticmethod
    def _aggregate_votes(ctx: LearningContext) -> List[str]:
        st, proto = ctx.state, ctx.communication_protocol
        rnd = st.round

        def expected() -> set:
            return set(proto.get_neighbors(only_direct=False).keys()) | {st.addr}

        def have_all() -> bool:
            if ctx.early_stopping_fn():
                return True
            got = set(st.train_set_votes.get(rnd, {}).keys())
            return expected() <= got

        complete = st.wait_for(have_all, Settings.VOTE_TIMEOUT)
        if ctx.early_stopping_fn():
            return []
        ballots = dict(st.train_set_votes.get(rnd, {}))
        known = expected()
        if not complete:
            logger.info(st.addr, f"vote timeout; missing ballots from {sorted(known - set(ballots))}")
        ballots = {voter: b for voter, b in ballots.items() if voter in known}
        return tally_votes(ballots, Settings.TRAIN_SET_SIZE)

    @staticmethod
    def _validate(ctx: LearningContext, train_set: List[str]) -> List[str]:
        st, proto = ctx.state, ctx.communication_protocol
        known = set(proto.get_neighbors(only_direct=False).keys()) | {st.addr}
        return [n for n in train_set if n in known]
