"""Name -> stage class (reference: p2pfl/stages/stage_factory.py:30-59).
Imports are lazy because the stages refer to each other through the factory."""

from __future__ import annotations

from typing import Type


class StageFactory:
    @staticmethod
    def get_stage(stage_name: str) -> Type:
        if stage_name == "StartLearningStage":
            from p2pfl_b200.stages.base_node.start_learning_stage import StartLearningStage

            return StartLearningStage
        if stage_name == "VoteTrainSetStage":
            from p2pfl_b200.stages.base_node.vote_train_set_stage import VoteTrainSetStage

            return VoteTrainSetStage
        if stage_name == "TrainStage":
            from p2pfl_b200.stages.base_node.train_stage import TrainStage

            return TrainStage
        if stage_name == "WaitAggregatedModelsStage":
            from p2pfl_b200.stages.base_node.wait_agg_models_stage import WaitAggregatedModelsStage

            return WaitAggregatedModelsStage
        if stage_name == "GossipModelStage":
            from p2pfl_b200.stages.base_node.gossip_model_stage import GossipModelStage

            return GossipModelStage
        if stage_name == "RoundFinishedStage":
            from p2pfl_b200.stages.base_node.round_finished_stage import RoundFinishedStage

            return RoundFinishedStage
        raise ValueError(f"unknown stage: {stage_name}")
