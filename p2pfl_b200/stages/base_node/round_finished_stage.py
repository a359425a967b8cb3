"""Stage 5: close the round; loop or finish
(reference: p2pfl/stages/base_node/round_finished_stage.py:41-97)."""

from __future__ import annotations

from typing import Optional, Type

from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings
from p2pfl_b200.stages.common import evaluate_and_share
from p2pfl_b200.stages.stage import LearningContext, Stage
from p2pfl_b200.stages.stage_factory import StageFactory


class RoundFinishedStage(Stage):
    @staticmethod
    def name() -> str:
        return "RoundFinishedStage"

    @staticmethod
    def execute(ctx: LearningContext) -> Optional[Type[Stage]]:
        st = ctx.state
        if ctx.early_stopping_fn():
            return None
        ctx.aggregator.clear()
        st.increase_round()
        logger.round_finished(st.addr)
        logger.info(st.addr, f"round {st.round} of {st.total_rounds} finished")
        if st.round is not None and st.total_rounds is not None and st.round < st.total_rounds:
            if Settings.VOTE_EVERY_ROUND:
                st.train_set = []
                return StageFactory.get_stage("VoteTrainSetStage")
            if st.addr in st.train_set:
                return StageFactory.get_stage("TrainStage")
            return StageFactory.get_stage("WaitAggregatedModelsStage")
        # experiment over
        evaluate_and_share(ctx)
        logger.experiment_finished(st.addr)
        logger.info(st.addr, "training finished")
        st.clear()  # keeps st.learner: the trained model stays reachable (check_equal_models reads it)
        return None
