"""Stage 1: set the experiment up and spread the initial weights
(reference: p2pfl/stages/base_node/start_learning_stage.py:43-136)."""

from __future__ import annotations

import inspect
import time
from typing import Any, List, Optional, Type

from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings
from p2pfl_b200.stages.common import pack_model
from p2pfl_b200.stages.stage import LearningContext, Stage
from p2pfl_b200.stages.stage_factory import StageFactory


def build_learner(ctx: LearningContext) -> Any:
    kwargs = {}
    try:
        accepted = inspect.signature(ctx.learner_class.__init__).parameters
        kwargs = {k: v for k, v in ctx.learner_kwargs.items() if k in accepted}
    except (TypeError, ValueError):
        pass
    return ctx.learner_class(ctx.model, ctx.data, ctx.state.addr, ctx.epochs, **kwargs)


class StartLearningStage(Stage):
    @staticmethod
    def name() -> str:
        return "StartLearningStage"

    @staticmethod
    def execute(ctx: LearningContext) -> Optional[Type[Stage]]:
        st = ctx.state
        t0 = time.time()
        with st.start_thread_lock:
            if st.round is not None:
                logger.debug(st.addr, "learning already started; duplicate start ignored")
                return None
            st.set_experiment("experiment", ctx.rounds)
            old = st.learner
            if old is not None:
                try:
                    old.close()  # hands its device arena back to the pool for the new learner
                except Exception:
                    pass
            st.learner = build_learner(ctx)
        publish = getattr(ctx.communication_protocol, "publish_arena", None)
        if publish is not None:
            publish(st.learner)
        logger.experiment_started(st.addr)

        # initial weights: the initiator set the event itself; everybody else
        # gets them through InitModelCommand (possibly stashed before the
        # learner existed)
        with st.stash_lock:
            pending, st.pending_init = st.pending_init, None
        if pending is not None and not st.model_initialized.is_set():
            pending()
        logger.info(st.addr, "waiting for the initial weights")
        ok = st.wait_for(lambda: st.model_initialized.is_set() or ctx.early_stopping_fn(), Settings.INIT_MODEL_TIMEOUT)
        if ctx.early_stopping_fn():
            return None
        if not ok:
            logger.error(st.addr, "initial weights never arrived; leaving the experiment")
            st.clear()
            return None

        StartLearningStage._gossip_init_model(ctx)

        rest = Settings.WAIT_HEARTBEATS_CONVERGENCE - (time.time() - t0)
        if rest > 0:
            time.sleep(rest)
        return StageFactory.get_stage("VoteTrainSetStage")

    @staticmethod
    def _gossip_init_model(ctx: LearningContext) -> None:
        st, proto = ctx.state, ctx.communication_protocol

        def candidates() -> List[str]:
            return [n for n in proto.get_neighbors(only_direct=True) if n not in st.nei_status]

        cache: List[Any] = []

        def model_fn(_: str) -> Any:
            if st.learner is None or st.round is None:
                return None
            if not cache:  # encode once, not per candidate per tick (Q13)
                cache.append(pack_model(proto, st.learner, None))
            return proto.build_weights("init_model", st.round, cache[0], ctx.aggregator.get_aggregated_models(), 1)

        proto.gossip_weights(
            ctx.early_stopping_fn,
            candidates,
            candidates,
            model_fn,
            wait_fn=lambda t: st.wait_for(lambda: not candidates() or ctx.early_stopping_fn(), t),
        )
