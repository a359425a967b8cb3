"""Inbound message handlers (reference: p2pfl/commands/)."""

from p2pfl_b200.commands.command import Command
from p2pfl_b200.commands.control import (
    HeartbeatCommand,
    MetricsCommand,
    ModelInitializedCommand,
    ModelsAggregatedCommand,
    ModelsReadyCommand,
    StartLearningCommand,
    StopLearningCommand,
    VoteTrainSetCommand,
)
from p2pfl_b200.commands.weights import AddModelCommand, InitModelCommand

__all__ = [
    "Command",
    "HeartbeatCommand",
    "MetricsCommand",
    "ModelInitializedCommand",
    "ModelsAggregatedCommand",
    "ModelsReadyCommand",
    "StartLearningCommand",
    "StopLearningCommand",
    "VoteTrainSetCommand",
    "AddModelCommand",
    "InitModelCommand",
]
