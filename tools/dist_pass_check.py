"""torchrun --nproc-per-node N tools/dist_pass_check.py: progressive passes over NCCL (distributed.render_pass) leave on
rank 0 the frame of the one-shot N-rank render (distributed.render_frame), bit for bit."""
import os, sys
from pathlib import Path
R = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(R))
os.environ.setdefault("B200PT_QUIET", "1")
import numpy as np, torch, torch.distributed as dist
import owl_path_tracer_b200 as pt
from owl_path_tracer_b200.distributed import render_frame, render_pass
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
sc = pt.Scene(scene="dragon", width=640, height=360, spp=4 * world, depth=8, env="color", env_intensity=0.0)
ctx = pt.Context(device=local, rank=rank, world_size=world, shard_mode=pt.SHARD_SPP); ctx.bind(sc)
render_frame(ctx)
want = ctx.framebuffer_f32().copy() if rank == 0 else None
v = []
for n in (1, 1, 2):
    info = render_pass(ctx, n)
    v.append(info["rel_variance"])
if rank == 0:
    got = ctx.framebuffer_f32()
    same = bool((got.view(np.uint32) == want.view(np.uint32)).all())
    print(f"dist_pass_check world={world}: frame after passes == one-shot frame: {same}; rel_variance per pass {v}; complete={info['complete']}", flush=True)
    assert same and info["complete"] == 1 and v[-1] > 0
dist.barrier()
dist.destroy_process_group()
