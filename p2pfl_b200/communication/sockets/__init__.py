from p2pfl_b200.communication.sockets.socket_communication_protocol import SocketCommunicationProtocol, SocketTransport

__all__ = ["SocketCommunicationProtocol", "SocketTransport"]
