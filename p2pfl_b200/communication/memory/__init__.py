from p2pfl_b200.communication.memory.memory_communication_protocol import InMemoryCommunicationProtocol

__all__ = ["InMemoryCommunicationProtocol"]
