"""Stage 4: install the aggregate and diffuse it to neighbours that lack it
(reference: p2pfl/stages/base_node/gossip_model_stage.py:41-132)."""

from __future__ import annotations

import time
from typing import Any, List, Optional, Type

from p2pfl_b200.exceptions import NoModelsToAggregateError
from p2pfl_b200.management import tracer
from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings
from p2pfl_b200.stages.common import pack_model
from p2pfl_b200.stages.common import send_async
from p2pfl_b200.stages.stage import LearningContext, Stage
from p2pfl_b200.stages.stage_factory import StageFactory


class GossipModelStage(Stage):
    @staticmethod
    def name() -> str:
        return "GossipModelStage"

    @staticmethod
    def execute(ctx: LearningContext) -> Optional[Type[Stage]]:
        st, proto, agg = ctx.state, ctx.communication_protocol, ctx.aggregator
        if st.learner is None or st.round is None:
            return None
        t0 = time.time()
        logger.info(st.addr, "waiting for the aggregation")
        try:
            params = agg.wait_and_get_aggregation()
        except NoModelsToAggregateError:
            if ctx.early_stopping_fn():
                return None
            raise
        if ctx.early_stopping_fn():
            return None
        st.learner.set_parameters(params)
        contributors = agg.get_aggregated_models()
        # after a complete fused exchange nobody waits for this round's models_ready: it may ride
        # with the next round's metrics / pre-vote bundle
        # ... provided every direct neighbour is such a trainer: a neighbour outside the train set waits
        # for this notice in its own diffusion loop (it would offer the model to this node otherwise)
        train_set = set(st.train_set)
        lazy = (
            getattr(st, "fused_complete_round", None) == st.round
            and st.total_rounds is not None
            and st.round + 1 < st.total_rounds
            and all(n in train_set for n in proto.get_neighbors(only_direct=True))
        )
        ready = proto.build_msg("models_ready", [], round=st.round)
        if Settings.DIRECT_ROUND_STATUS:
            ready.ttl = 1  # only direct neighbours ever look at it (candidates() / pending() below)
        send_async(proto, ready, lazy)
        t1 = time.time()
        GossipModelStage._diffuse(ctx, params, contributors, t1)
        t2 = time.time()
        ctx.timings["aggregate"] = ctx.timings.get("aggregate", 0.0) + t1 - t0
        ctx.timings["diffuse"] = ctx.timings.get("diffuse", 0.0) + t2 - t1
        if tracer.enabled():
            tracer.complete("wait for the aggregate", st.addr, t0, t1, "phase", {"round": st.round})
            tracer.complete("diffuse", st.addr, t1, t2, "phase", {"round": st.round})
        if ctx.early_stopping_fn():
            return None
        return StageFactory.get_stage("RoundFinishedStage")

    @staticmethod
    def _diffuse(ctx: LearningContext, params: Any, contributors: List[str], t_start: float) -> None:
        st, proto = ctx.state, ctx.communication_protocol
        fixed_round = st.round
        train_set = set(st.train_set)
        # after a complete fused exchange every trainer holds the aggregate by construction
        trainers_done = getattr(st, "fused_complete_round", None) == fixed_round

        def candidates() -> List[str]:
            # trainers normally finish on their own; give them a grace period
            # before spending a full-model transfer on them
            grace_over = time.time() - t_start >= Settings.TRAINER_DIFFUSION_GRACE
            out = []
            for n in proto.get_neighbors(only_direct=True):
                if st.nei_status.get(n, -1) >= fixed_round:  # Q17: .get
                    continue
                if trainers_done and n in train_set:
                    continue
                if n in train_set and not grace_over:
                    continue
                out.append(n)
            return out

        def pending() -> List[str]:
            return [
                n
                for n in proto.get_neighbors(only_direct=True)
                if st.nei_status.get(n, -1) < fixed_round and not (trainers_done and n in train_set)
            ]

        cache: List[Any] = []

        def model_fn(_: str) -> Any:
            if st.learner is None or st.round is None:
                return None
            if not cache:
                cache.append(pack_model(proto, st.learner, params))
            return proto.build_weights("add_model", st.round, cache[0], contributors, 1)

        def wait_fn(t: float) -> None:
            # "nobody left to serve" counts as well: the last ``models_ready`` often arrives WHILE the
            # models of this tick are being sent, i.e. before this snapshot is taken - waiting for a
            # further change then slept a whole gossip period (50 ms of a 55 ms round in
            # bench/control_plane_bench.py --train-set 4)
            snap = pending()
            st.wait_for(lambda: ctx.early_stopping_fn() or not candidates() or pending() != snap, t)

        # Non-trainers are served at once; trainers only once the grace period is
        # over and they still have not announced ``models_ready``.
        while not ctx.early_stopping_fn():
            proto.gossip_weights(ctx.early_stopping_fn, candidates, pending, model_fn, wait_fn=wait_fn)
            left = Settings.TRAINER_DIFFUSION_GRACE - (time.time() - t_start)
            if not pending() or left <= 0:
                break
            snap = pending()
            st.wait_for(lambda: ctx.early_stopping_fn() or not pending() or pending() != snap, left)
