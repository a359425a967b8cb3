from p2pfl_b200.communication.grpc.grpc_communication_protocol import GrpcCommunicationProtocol

__all__ = ["GrpcCommunicationProtocol"]
