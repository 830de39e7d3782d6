"""One node of the ForceQuit shrink test (tests/test_elastic.py), CPU / gloo: three nodes start an Engine,
node 1 "read no data" and leaves; the survivors must find themselves in a two-node cluster."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch.distributed as dist
    from minips_b200.parallel.engine import Engine
    from minips_b200.parallel.types import Node
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    ids = [0, 2, 5][:world]
    nodes = [Node(i, "localhost", 0) for i in ids]
    engine = Engine(nodes[rank], nodes)
    engine.start_everything(num_server_threads_per_node=2)
    stay = engine.force_quit(leaving=(rank == 1))
    if not stay:
        assert rank == 1 and not engine.started
        print(f"LEFT node={ids[rank]}", flush=True)
    else:
        assert [n.id for n in engine.nodes] == [0, 5], engine.nodes
        assert [n.id for n in engine.left_nodes] == [2]
        assert engine.rank_index == (0 if rank == 0 else 1)
        assert engine.control.all_gather(engine.node.id) == [0, 5]
        assert engine.id_mapper.get_all_server_threads() == [0, 1, 5000, 5001]
        r = engine.get_ranges(10)
        assert len(r) == 4 and r[0][0] == 0 and r[-1][1] == 10
        engine.control.barrier()
        # a second call with nobody leaving changes nothing
        assert engine.force_quit(False) and len(engine.nodes) == 2
        print(f"STAYED node={engine.node.id} rank={engine.rank_index} of {len(engine.nodes)}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
