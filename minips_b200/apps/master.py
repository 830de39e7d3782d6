"""The master node of a fault-tolerant run (reference master/master.hpp:14-87; node id 1 when
heartbeat_interval > 0, base/node_utils.cpp:19-32):

    python -m minips_b200.apps.master --config_file=config/localnodes --heartbeat_interval=1 \
        --relaunch_cmd="python scripts/relaunch_rank.py launch.json" --store_port=29700

Records heartbeats, declares a node failed after 3 silent intervals, runs `relaunch_cmd <node id>`,
orders the global rollback when the relaunched process reports back, and exits when every node has
said goodbye.  Also hosts the store on which the ranks of every generation meet
(minips_b200/parallel/fault.py)."""
from __future__ import annotations

import sys


def main(argv=None):
    from .. import core
    from ..parallel.fault import MasterProcess
    from ..utils.flags import Context

    ctx = Context.get_instance()
    ctx.define("config_file", "", "id:host:port per line")
    ctx.define("heartbeat_interval", "1", "seconds")
    ctx.define("relaunch_cmd", "", "command that restarts a node; the failed node id is appended")
    ctx.define("store_port", "29700", "TCP port of the rendezvous store (master host)")
    ctx.define("scale_file", "", "where a scale-out node writes its id:host:port (shared path)")
    ctx.parse(argv)
    nodes = core.parse_file(ctx.get_string("config_file"))
    mp = MasterProcess(nodes, ctx.get_int32("heartbeat_interval"), ctx.get_string("relaunch_cmd"),
                       ctx.get_int32("store_port"), ctx.get_string("scale_file"))
    print(f"[master] node {mp.master_node.id} supervising {[n.id for n in mp.rank_nodes]}", flush=True)
    mp.serve()
    print(f"[master] every node quit; rollbacks ordered for {mp.rollbacks}, scale-outs for {mp.scaled}", flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
