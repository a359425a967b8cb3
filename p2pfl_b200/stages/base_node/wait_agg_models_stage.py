"""Stage 3b (non-trainers): just wait for the finished aggregate
(reference: p2pfl/stages/base_node/wait_agg_models_stage.py:48)."""

from __future__ import annotations

from typing import Optional, Type

from p2pfl_b200.stages.base_node.train_stage import TrainStage
from p2pfl_b200.stages.stage import LearningContext, Stage
from p2pfl_b200.stages.stage_factory import StageFactory


class WaitAggregatedModelsStage(Stage):
    @staticmethod
    def name() -> str:
        return "WaitAggregatedModelsStage"

    @staticmethod
    def execute(ctx: LearningContext) -> Optional[Type[Stage]]:
        ctx.aggregator.set_waiting_aggregated_model(ctx.state.train_set)
        TrainStage.replay_early_models(ctx)
        TrainStage.prevote(ctx)
        return StageFactory.get_stage("GossipModelStage")
