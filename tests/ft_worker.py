"""CPU stand-in for a rank of a fault-tolerant run (tests/test_fault_tolerance.py): the whole control
plane — heartbeats, detection, relaunch, rollback order, new generation, resume from the checkpoint —
with a toy "model" (a vector that every iteration adds an all-reduced, iteration-dependent delta to) in
place of the device tables."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch.distributed as dist
    from minips_b200 import core
    from minips_b200.parallel.fault import FaultTolerantJob
    a = json.loads(sys.argv[1])
    relaunched = "--relaunched=true" in sys.argv
    my_id, out = int(a["my_id"]), a["out_dir"]
    nodes = core.parse_file(a["config_file"])
    scale_node = core.Node(my_id, "localhost", int(a["scale_port"])) if a.get("scale_port") else None
    job = FaultTolerantJob(nodes, my_id, a["heartbeat_interval"], a["store_port"], relaunched=relaunched,
                           pg_timeout_s=20, scale_node=scale_node, scale_file=a.get("scale_file", ""))
    job.start()
    iters, ck_every = a["num_iters"], a["checkpoint_interval"]

    def body(job):
        w = torch.zeros(8, dtype=torch.float64)
        first = 0
        ck = os.path.join(out, f"ckpt_{my_id}.pt")
        src = ck
        if job.resume and not os.path.exists(ck):
            # a node that scaled in has no checkpoint of its own: it takes the state of the others
            # (the toy state is replicated; the device backend re-shards, utils/checkpoint.py)
            others = sorted(f for f in os.listdir(out) if f.startswith("ckpt_") and f.endswith(".pt"))
            src = os.path.join(out, others[0]) if others else ck
        if job.resume and os.path.exists(src):
            st = torch.load(src)
            w, first = st["w"], st["it"]
        print(f"[node {my_id}] generation {job.generation} starts at iteration {first} as rank {job.rank} of "
              f"{job.world}", flush=True)
        for it in range(first, iters):
            job.check()
            g = torch.full((8,), float((it + 1) * (job.rank + 1)), dtype=torch.float64)
            dist.all_reduce(g)
            w = w + g * 1e-3
            if (it + 1) % ck_every == 0:
                dist.barrier()
                torch.save({"w": w, "it": it + 1}, ck + ".tmp")
                os.replace(ck + ".tmp", ck)
                dist.barrier()
            with open(os.path.join(out, f"progress_{my_id}"), "w") as f:
                f.write(str(it + 1))
            time.sleep(a.get("iter_sleep", 0.02))
        dist.barrier()
        torch.save(w, os.path.join(out, f"final_{my_id}.pt"))

    job.run(body)
    job.stop()
    print(f"[node {my_id}] done after {job.generations_run} generation(s)", flush=True)


if __name__ == "__main__":
    main()
