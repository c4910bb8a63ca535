"""Ad-hoc multi-GPU probe (torchrun): C4 tiles, frame time per exchange mode (0 none, 1 NCCL all-gather, 2 fused peer stores)."""
import os, sys, json, time
sys.path.insert(0, ".")
import torch, torch.distributed as dist
import raytracer_b200 as rtb
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
sc = rtb.Scene(6, 40, 35)
tr = rtb.Tracer(rank).setup(sc, 3840, 2160, spp, 8)
ids = [rtb.Tracer.comm_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
tr.comm_init(ids[0], world, rank)
tr.set_shard(rank, world, 8)
handles = [None] * world
dist.all_gather_object(handles, tr.peer_export())
tr.peer_open(b"".join(handles), world, rank)
for kernel in (1, 0):
    for ex in (0, 1, 2):
        tr.set_option("kernel", kernel); tr.set_option("exchange", ex)
        tr.reset(); tr.render_device(1, 999); tr.sync()
        dist.barrier(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        n = 8
        for k in range(n):
            tr.reset(); tr.render_device(1, 1000 + k); tr.sync()
        dist.barrier(); torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3 / n
        if rank == 0:
            print(json.dumps(dict(world=world, spp=spp, kernel=kernel, exchange=ex, wall_ms_per_frame=round(wall, 2), last_device_ms=round(tr.device_ms(), 2),
                                  fused=int(tr.info("exchange_fused")))), flush=True)
tr.peer_close(); tr.close()
dist.destroy_process_group()
