"""Message envelopes of the worker<->hub runtime.

Reference: messages/FlinkMessage.scala:8-23, ControlMessage.scala:18-23, SpokeMessage.scala:18-23,
HubMessage.scala:8-13,48-55; node ids and RPC ids are the out-of-tree
`BipartiteTopologyAPI.sites.NodeId` / `operations.RemoteCallIdentifier`
(call sites FlinkLearning.scala:70, FlinkSpoke.scala:200,222, FlinkMessage.scala:18).

`HubMessage` carries *arrays* of operations and destinations so that a broadcast payload is
held once (HubMessage.scala:8-13).  `get_size` follows "4 + sizes of present fields"
(FlinkMessage.scala:16-23); the reference's HubMessage double counting quirk
(HubMessage.scala:50,52) is deliberately not reproduced.
"""
from __future__ import annotations

import enum
from dataclasses import dataclass, field
from typing import Any, List, Optional

from .api.pojos import Request, _sizeof


class NodeType(enum.Enum):
    SPOKE = "SPOKE"
    HUB = "HUB"


@dataclass(frozen=True)
class NodeId:
    node_type: Optional[NodeType]
    node_id: int

    def get_size(self) -> int:
        return 8

    def __str__(self) -> str:
        return f"{self.node_type.value if self.node_type else 'NONE'}_{self.node_id}"


@dataclass(frozen=True)
class RemoteCallIdentifier:
    """Which remote method a message invokes on the receiving node."""
    op: str
    call_number: int = 0

    def get_size(self) -> int:
        return 4 + len(self.op)


@dataclass
class NetworkDescriptor:
    network_id: int
    n_spokes: int
    n_hubs: int


def _opt_size(o: Any) -> int:
    return 0 if o is None else _sizeof(o)


@dataclass
class FlinkMessage:
    network_id: int = -1
    operation: Optional[RemoteCallIdentifier] = None
    source: Optional[NodeId] = None
    destination: Optional[NodeId] = None
    data: Any = None
    request: Optional[Request] = None

    def get_size(self) -> int:
        return 4 + _opt_size(self.operation) + _opt_size(self.source) + _opt_size(self.destination) + \
            _opt_size(self.data) + _opt_size(self.request)


@dataclass
class ControlMessage(FlinkMessage):
    """hub -> worker deliveries and request fan-out (ControlMessage.scala:18-23)."""


@dataclass
class SpokeMessage(FlinkMessage):
    """worker -> hub (SpokeMessage.scala:18-23)."""


@dataclass
class HubMessage:
    network_id: int = -1
    operations: List[RemoteCallIdentifier] = field(default_factory=list)
    source: Optional[NodeId] = None
    destinations: List[NodeId] = field(default_factory=list)
    data: Any = None
    request: Optional[Request] = None

    def get_size(self) -> int:
        return 4 + sum(o.get_size() for o in self.operations) + _opt_size(self.source) + \
            sum(d.get_size() for d in self.destinations) + _opt_size(self.data) + _opt_size(self.request)

    def fan_out(self) -> List[ControlMessage]:
        """One ControlMessage per destination (FlinkLearning.scala:65-75)."""
        if self.network_id == -1:
            return []
        return [ControlMessage(self.network_id, op, self.source, dst, self.data, self.request)
                for op, dst in zip(self.operations, self.destinations)]
