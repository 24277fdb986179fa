"""torchrun --nproc-per-node 2: sharded adaptive render over NCCL vs the single-rank render of the same seed."""
import os, sys, numpy as np
sys.path.insert(0, '.')
import torch, torch.distributed as dist
from pathtracing_b200 import capi, multi, synth
from pathtracing_b200.scene import GpuScene, render_params
rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
desc, cam = synth.cornell_scene(light_scale=40.0, light_size=0.12)
W = H = 256; spp = 4
scene = GpuScene(desc, 4, device=local)
p = render_params(W, H, spp, 5, seed=5, adaptive=1)
st, film = multi.render_sharded(scene, cam, p, dist)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); st, film = multi.render_sharded(scene, cam, p, dist); e1.record(); torch.cuda.synchronize()
if rank == 0:
    got = film.cpu().numpy().reshape(H, W, 4).copy()
    one = GpuScene(desc, 4, device=local)
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    one.render(cam, p)
    e2.record(); s1 = one.render(cam, p); e3.record(); torch.cuda.synchronize()
    ref = one.film()
    print("samples 2-rank (rank0 share)", st["samples"], "single", s1["samples"], "weights equal:", np.array_equal(got[..., 3], ref[..., 3]),
          "max rel rgb diff", float(np.abs(got[..., :3] - ref[..., :3]).max() / ref[..., :3].max()),
          f"ms 2-rank {e0.elapsed_time(e1):.1f} single {e2.elapsed_time(e3):.1f}", flush=True)
dist.barrier(); dist.destroy_process_group()
