"""The `Network` the worker / parameter-server nodes talk to, and the node wrappers.

Reference: network/FlinkNetwork.scala - `send` :242-274 (destination None = a QueryResponse or
Prediction for the job's side outputs, :246-258), `broadcast` :276-293 (one HubMessage with
arrays of operations/destinations, :282-291), `describe` :295; wrappers are the out-of-tree
`BipartiteTopologyAPI` `BufferingWrapper` / `GenericWrapper` (call sites FlinkSpoke.scala:199-219,
FlinkHub.scala:182-193).  `IntWrapper` is `mlAPI.protocols.IntWrapper` (FlinkSpoke.scala:31,346-347).
"""
from __future__ import annotations

from collections import deque
from typing import Any, Callable, Deque, Dict, Generic, List, Optional, TypeVar

from ..api.bucketing import bucket_response
from ..api.pojos import QueryResponse
from ..messages import (HubMessage, NetworkDescriptor, NodeId, NodeType, RemoteCallIdentifier, SpokeMessage)

T = TypeVar("T")


class IntWrapper:
    def __init__(self, n: int):
        self._n = int(n)

    def get_int(self) -> int:
        return self._n

    def set_int(self, n: int) -> None:
        self._n = int(n)


class DataSet(Generic[T]):
    """Bounded FIFO (`mlAPI.dataBuffers.DataSet`): append returns the EVICTED element when full
    (FlinkSpoke.scala:96-99, SpokeLogic.scala:32,35)."""

    def __init__(self, max_size: int = 100000):
        self.max_size = max_size
        self.data_buffer: Deque[T] = deque()

    def append(self, x: T) -> Optional[T]:
        self.data_buffer.append(x)
        if len(self.data_buffer) > self.max_size:
            return self.data_buffer.popleft()
        return None

    def pop(self) -> Optional[T]:
        return self.data_buffer.popleft() if self.data_buffer else None

    def is_empty(self) -> bool:
        return not self.data_buffer

    def non_empty(self) -> bool:
        return bool(self.data_buffer)

    def __len__(self) -> int:
        return len(self.data_buffer)

    @property
    def length(self) -> int:
        return len(self.data_buffer)

    def get_max_size(self) -> int:
        return self.max_size

    def clear(self) -> None:
        self.data_buffer.clear()

    def merge(self, others: List["DataSet[T]"]) -> "DataSet[T]":
        """Interleaving merge used when several old subtask states are restored into one
        (SpokeLogic.scala:37-50, CommonUtils.scala:35-48)."""
        lists = [list(self.data_buffer)] + [list(o.data_buffer) for o in others]
        merged: List[T] = []
        i = 0
        while any(i < len(l) for l in lists):
            for l in lists:
                if i < len(l):
                    merged.append(l[i])
            i += 1
        self.data_buffer = deque(merged[-self.max_size:])
        return self


class Network:
    """`BipartiteTopologyAPI.interfaces.Network`."""

    def send(self, source: NodeId, destination: Optional[NodeId], rpc: RemoteCallIdentifier, message: Any) -> None:
        raise NotImplementedError

    def broadcast(self, source: NodeId, destinations: Dict[NodeId, RemoteCallIdentifier], message: Any) -> None:
        raise NotImplementedError

    def describe(self) -> NetworkDescriptor:
        raise NotImplementedError


class OperatorNetwork(Network):
    """Adapter between nodes and the runtime's collectors (the `FlinkNetwork` role)."""

    def __init__(self, node_type: NodeType, network_id: int, n_spokes: IntWrapper, n_hubs: int,
                 collect: Callable[[Any], None], side_output: Callable[[Any], None]):
        self.node_type = node_type
        self.network_id = network_id
        self.n_spokes = n_spokes
        self.n_hubs = n_hubs
        self.collect = collect          # SpokeMessage (worker side) / HubMessage (hub side)
        self.side_output = side_output  # QueryResponse / Prediction

    def send_query_response(self, resp: QueryResponse) -> None:
        for frag in bucket_response(resp):
            self.side_output(frag)

    def send(self, source, destination, rpc, message):
        if destination is None:
            if isinstance(message, QueryResponse):
                self.send_query_response(message)
            else:
                self.side_output(message)
            return
        if self.node_type == NodeType.SPOKE:
            self.collect(SpokeMessage(self.network_id, rpc, source, destination, message, None))
        else:
            self.collect(HubMessage(self.network_id, [rpc], source, [destination], message, None))

    def broadcast(self, source, destinations, message):
        if self.node_type == NodeType.SPOKE:
            for dst, rpc in destinations.items():
                self.collect(SpokeMessage(self.network_id, rpc, source, dst, message, None))
        else:
            self.collect(HubMessage(self.network_id, list(destinations.values()), source,
                                    list(destinations.keys()), message, None))

    def describe(self):
        return NetworkDescriptor(self.network_id, self.n_spokes.get_int(), self.n_hubs)


class GenericWrapper:
    """Hub-side wrapper: owns the node and its network, forwards RPCs."""

    def __init__(self, node_id: NodeId, node, network: Network):
        self.node_id, self.node, self.network = node_id, node, network
        node.bind(node_id, network)

    def get_node(self):
        return self.node

    def get_network(self):
        return self.network

    def receive_msg(self, source, rpc, data) -> None:
        self.node.receive_msg(source, rpc, data)

    def merge(self, others) -> "GenericWrapper":
        return self


class BufferingWrapper(GenericWrapper):
    """Worker-side wrapper: buffers records while the worker waits for its hub (blocked), and
    replays them when the reply arrives.  `toggle` nudges a pipeline after another pipeline of the
    same spoke consumed a hub message (FlinkSpoke.scala:130)."""

    def __init__(self, node_id, node, network, max_buffer: int = 100000):
        super().__init__(node_id, node, network)
        self.buffer: DataSet = DataSet(max_buffer)
        self._buf_samples = 0
        self._buf_sum = 0.0

    def receive_tuple(self, args) -> None:
        point = args[0] if isinstance(args, (list, tuple)) else args
        self._buf_samples += 1
        self._buf_sum += len(self.buffer)
        if self.node.blocked:
            self.buffer.append(point)
            return
        self._drain()
        if self.node.blocked:
            self.buffer.append(point)
        else:
            self.node.receive_tuple(point)

    def _drain(self) -> None:
        while self.buffer.non_empty() and not self.node.blocked:
            self.node.receive_tuple(self.buffer.pop())

    def receive_msg(self, source, rpc, data) -> None:
        self.node.receive_msg(source, rpc, data)
        self._drain()

    def receive_query(self, response_id, payload) -> None:
        self.node.receive_query(response_id, payload)

    def toggle(self) -> None:
        self._drain()

    def get_mean_buffer_size(self) -> float:
        return self._buf_sum / self._buf_samples if self._buf_samples else 0.0
