"""All framework exceptions in one place.

Names match the reference so ``except`` clauses port unchanged
(reference: p2pfl/exceptions.py:21-36, p2pfl/communication/exceptions.py:21-24,
p2pfl/learning/exceptions.py:21-31, p2pfl/learning/aggregators/aggregator.py:31-34).
The sub-package ``exceptions`` modules re-export from here.
"""


class NodeRunningException(Exception):
    """An operation needs the node to be running / not running and it is not."""


class LearnerNotSetException(Exception):
    """Learning was started on a node that has no learner class."""


class ZeroRoundsException(Exception):
    """``set_start_learning`` was called with ``rounds < 1``."""


class NeighborNotConnectedError(Exception):
    """A send targeted an address that is not a (direct) neighbour."""


class CommunicationError(Exception):
    """A transport-level failure while talking to a peer."""


class DecodingParamsError(Exception):
    """A received model payload could not be decoded."""


class ModelNotMatchingError(Exception):
    """A received model does not have the local model's layout."""


class NoModelsToAggregateError(Exception):
    """``aggregate`` was called with an empty model set."""


class ArenaError(Exception):
    """The device arena runtime reported a failure."""


class ExtensionMissingError(RuntimeError):
    """The in-tree CUDA extension is required on this path and is not built."""


class PayloadUnavailableError(RuntimeError):
    """A model payload that names device memory cannot be read from here (yet): the sender's arena is
    not known or not mappable.  Transient from the receiver's point of view - the message is dropped,
    the sender's gossip loop offers the model again - unlike a payload that decodes to a wrong model."""
