from p2pfl_b200.communication.nvlink.nvlink_protocol import (
    NvlinkCommunicationProtocol,
    NvlinkMemoryProtocol,
    NvlinkSocketProtocol,
)

__all__ = ["NvlinkCommunicationProtocol", "NvlinkMemoryProtocol", "NvlinkSocketProtocol"]
