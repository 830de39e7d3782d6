"""Hostfile handling (reference base/node_utils.hpp:15-54, .cpp:19-109): `id:hostname:port`
per line; with heartbeat_interval > 0 the node with id 1 becomes the master and leaves the
worker set."""
from __future__ import annotations

import os
from typing import List, Optional, Tuple

from ..parallel.types import Node


def parse_lines(text: str) -> List[Node]:
    nodes = []
    for raw in text.splitlines():
        line = raw.strip()
        if not line or line.startswith("#"):
            continue
        parts = line.split(":")
        if len(parts) < 3:
            raise ValueError(f"bad hostfile line {raw!r} (want id:host:port)")
        nodes.append(Node(int(parts[0]), ":".join(parts[1:-1]), int(parts[-1])))
    return nodes


def parse_file(path: str) -> List[Node]:
    from .checkpoint import _resolve
    with open(_resolve(path)) as f:
        return parse_lines(f.read())


def select_master(nodes: List[Node], heartbeat_interval: int) -> Tuple[Optional[Node], List[Node]]:
    """Returns (master or None, remaining nodes)."""
    if heartbeat_interval <= 0:
        return None, list(nodes)
    master = None
    rest = []
    for n in nodes:
        if n.id == 1 and master is None:
            master = Node(n.id, n.hostname, n.port, True)
        else:
            rest.append(n)
    return master, rest


def check_valid_node_ids(nodes: List[Node]) -> bool:
    return all(0 <= n.id < 1000 for n in nodes)


def check_unique_port(nodes: List[Node]) -> bool:
    return len({(n.hostname, n.port) for n in nodes}) == len(nodes)


def check_consecutive_ids(nodes: List[Node]) -> bool:
    return all(n.id == i for i, n in enumerate(nodes))


def has_node(nodes: List[Node], node_id: int) -> bool:
    return any(n.id == node_id for n in nodes)


def get_node_by_id(nodes: List[Node], node_id: int) -> Node:
    for n in nodes:
        if n.id == node_id:
            return n
    raise KeyError(f"Node {node_id} is not in the given node list")


def nodes_from_env() -> Tuple[Node, List[Node]]:
    """torchrun-style environment -> (my node, all nodes)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    nodes = [Node(i, "localhost", 0) for i in range(world)]
    return nodes[rank], nodes
