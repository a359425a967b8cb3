"""Workflow runner (reference: p2pfl/stages/workflows.py:35-55).

A workflow object is single-use: ``Node`` builds a fresh one per experiment,
which is what makes a second ``set_start_learning`` on the same node work
(the reference keeps ``current_stage`` across runs).
"""

from __future__ import annotations

import time
from typing import List, Optional, Type

from p2pfl_b200.management import tracer
from p2pfl_b200.management.logger import logger
from p2pfl_b200.settings import Settings
from p2pfl_b200.stages.stage import LearningContext, Stage
from p2pfl_b200.stages.stage_factory import StageFactory


def _nvtx():
    """``torch.cuda.nvtx`` when ranges are wanted and a device is there, else None."""
    if not Settings.NVTX_RANGES:
        return None
    try:
        import torch

        return torch.cuda.nvtx if torch.cuda.is_available() else None
    except Exception:
        return None


class StageWorkflow:
    def __init__(self, first_stage: Type[Stage]) -> None:
        self.current_stage: Optional[Type[Stage]] = first_stage
        self.history: List[str] = []
        self.finished = False

    def run(self, ctx: LearningContext) -> None:
        while self.current_stage is not None:
            stage = self.current_stage
            self.history.append(stage.name())
            logger.debug(ctx.state.addr, f"running stage {stage.name()}")
            t0 = time.time()
            nvtx = _nvtx()
            if nvtx is not None:  # one range per stage and round, visible in nsys / ncu --nvtx
                nvtx.range_push(f"{stage.name()}[r{ctx.state.round}]")
            try:
                with tracer.span(stage.name(), ctx.state.addr, "stage", round=ctx.state.round):
                    self.current_stage = stage.execute(ctx)
            finally:
                if nvtx is not None:
                    nvtx.range_pop()
            ctx.timings["stage:" + stage.name()] = ctx.timings.get("stage:" + stage.name(), 0.0) + time.time() - t0
        self.finished = True


class LearningWorkflow(StageWorkflow):
    """StartLearning -> VoteTrainSet -> (Train | WaitAggregatedModels) -> GossipModel -> RoundFinished -> ..."""

    def __init__(self) -> None:
        super().__init__(StageFactory.get_stage("StartLearningStage"))
