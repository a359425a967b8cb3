"""Learning exceptions (reference: p2pfl/learning/exceptions.py:21-31)."""

from p2pfl_b200.exceptions import DecodingParamsError, ModelNotMatchingError, NoModelsToAggregateError

__all__ = ["DecodingParamsError", "ModelNotMatchingError", "NoModelsToAggregateError"]
