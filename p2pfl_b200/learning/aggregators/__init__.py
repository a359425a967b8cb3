from p2pfl_b200.learning.aggregators.aggregator import Aggregator
from p2pfl_b200.learning.aggregators.fedavg import FedAvg

__all__ = ["Aggregator", "FedAvg"]
