"""All-reduce shape sweep inside ONE multi-rank launch (8-GPU time is charged eight-fold): for each (CTA shape, CTA count)
of the own kernel, and for NCCL ring, the step time of the default model and of the scaled-width model.
    python -m torch.distributed.run --nproc-per-node N scripts/ar_sweep.py [hidden ...]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tutorial-torch-distributed-data-parallel_b200")]
import torch                      # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench                      # noqa: E402
from b200ddp import _lib          # noqa: E402


def main():
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    cores = sorted(os.sched_getaffinity(0))
    per = max(1, len(cores) // world)
    os.sched_setaffinity(0, cores[local * per:(local + 1) * per])
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    L = _lib.load()

    class A:
        batch, hidden, mode, steps, warmup = 128, 4096, "tf32", 20, 3
    hiddens = [int(h) for h in sys.argv[1:]] or [4096, 8192]

    def sync_all():
        torch.cuda.synchronize(dev)
        dist.barrier()
        torch.cuda.synchronize(dev)
    for hidden in hiddens:
        job = bench.Job(A, dev, world, rank, local, hidden=hidden)
        job.prepare_indices(4000)
        comm = job.ddp.comm
        for _ in range(5):
            job.step_resident()
        configs = [("small", c) for c in (16, 32, 64, 128)] + [("big", c) for c in (4, 8, 16)] + [("ring", 0)]
        if os.environ.get("AR_SWEEP_CONFIGS"):       # e.g. "small:16,small:32,big:4"
            configs = [(c.split(":")[0], int(c.split(":")[1])) for c in os.environ["AR_SWEEP_CONFIGS"].split(",")]
        for shape, ctas in configs:
            if shape == "ring":
                comm.set_algo("ring")
                L.b200ddp_set_sm_budget(148 - 16)
            else:
                comm.set_algo("own", ctas)
                L.b200ddp_comm_set_cta_shape(comm.handle, 1 if shape == "big" else 0)
                L.b200ddp_set_sm_budget(148 - ctas if shape == "big" else 0)
            for _ in range(3):
                job.step_resident()
            ms, _ = bench.timed_steps(job, 25, world, sync_all)
            comm.set_timing(True)
            for _ in range(2):
                job.step_resident()
            torch.cuda.synchronize()
            times = comm.bucket_times()
            comm.set_timing(False)
            coll = sum(t for t, _ in times)
            by = sum(4.0 * n for _, n in times)
            if rank == 0:
                print(json.dumps({"hidden": hidden, "shape": shape, "ctas": ctas, "ms_per_step": round(ms, 4),
                                  "collective_ms": round(coll, 4),
                                  "busbw_GBps": round(by * 2 * (world - 1) / world / (coll * 1e-3) / 1e9, 1) if coll > 0 else None}),
                      flush=True)
        job.free()
    dist.barrier()
    from b200ddp import comm as _comm
    _comm.destroy_all()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
