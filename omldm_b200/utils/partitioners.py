"""Routing helpers.

* `identity_partitioner` - utils/partitioners/random_partitioner.scala:7-15: despite its name the
  key IS the destination worker id.
* `hub_message_partition` - utils/kafkaPartitioners/FlinkHubMessagePartitioner.scala:7-20:
  termination (-1) -> 0; single destination -> worker % n; broadcast -> network % n.
* `round_robin_partition` - the unused raw partitioner utils/kafkaPartitioners/HubMessagePartitioner.scala.
* `hub_key` - the "networkId_hubId" key of FlinkLearning.scala:91-94.
"""
from __future__ import annotations

import itertools

from ..messages import HubMessage


def identity_partitioner(key: int, num_partitions: int) -> int:
    assert 0 <= key < num_partitions, "destination worker id out of range"
    return key


def hub_message_partition(msg: HubMessage, n_partitions: int) -> int:
    if msg.network_id == -1:
        return 0
    if len(msg.destinations) == 1:
        return msg.destinations[0].node_id % n_partitions
    return msg.network_id % n_partitions


_rr = itertools.count()


def round_robin_partition(n_partitions: int) -> int:
    return next(_rr) % n_partitions


def hub_key(network_id: int, hub_id: int) -> str:
    return f"{network_id}_{hub_id}"
