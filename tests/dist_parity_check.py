"""Multi-GPU parity script (run under torchrun on >= 2 GPUs; not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/dist_parity_check.py

Every rank trains the same small scene: once with the distributed engine (global batch = world x 4
cameras, SH state sharded) and, on rank 0, once with the single-GPU engine on the whole global
batch.  The two must agree: same losses, Adam moments (= gradients) within 1e-3 relative."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np
import torch
import torch.distributed as dist

from clmgs import engine, engine_dist, model, scene as S

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = "cuda:%d" % int(os.environ["LOCAL_RANK"])
dist.init_process_group("nccl", device_id=torch.device(dev))
W, H, n, bl = 160, 96, 40_000, 4
B = bl * world
spec = S.SceneSpec("distcheck", n, W, H, B, "aerial", 50.0, 0)
g = S.make_gaussians(spec)
V, K, P = S.make_cameras(spec, B, g["extent"])
gts = S.make_gt_images(spec, B)
args = model.TrainArgs(bsz=B, cpu_adam_threads=2)
ctx = engine_dist.DistContext(n)
gm = model.GaussianModelCLMOffload(3, dev).create_from_arrays(g["xyz"], g["log_scale"], g["raw_quat"], g["opacity_logit"],
                                                             g["sh"][ctx.lo:ctx.hi]).training_setup(args)
cams = [model.Camera(V[i], K[i], gts[i], dev, uid=i) for i in range(rank * bl, (rank + 1) * bl)]
comm = torch.cuda.Stream()
all_losses = []
for it in range(2):
    losses, _, _ = engine_dist.clm_offload_train_one_batch_dist(gm, None, cams, gm.parameters_grad_buffer, None, None, comm,
                                                                None, ctx)
    l = torch.stack(losses)
    gathered = [torch.empty_like(l) for _ in range(world)]
    dist.all_gather(gathered, l)
    all_losses.append(torch.cat(gathered).cpu().numpy())
# gather the sharded SH table + moments on rank 0
def gather_shards(t):
    parts = [None] * world
    dist.all_gather_object(parts, t.numpy())
    return np.concatenate(parts)
sh_all, m_all = gather_shards(gm._parameters), gather_shards(gm.optimizer.cpu_adam.exp_avg)
ok = True
if rank == 0:
    ref = model.GaussianModelCLMOffload(3, dev).create_from_arrays(g["xyz"], g["log_scale"], g["raw_quat"],
                                                                  g["opacity_logit"], g["sh"]).training_setup(args)
    rcams = [model.Camera(V[i], K[i], gts[i], dev, uid=i) for i in range(B)]
    gen = torch.Generator().manual_seed(0)
    for it in range(2):
        losses, order, _ = engine.clm_offload_train_one_batch(ref, None, rcams, ref.parameters_grad_buffer, None, None, comm, gen)
        ref_l = np.empty(B, np.float32)
        ref_l[order] = torch.stack(losses).cpu().numpy()
        ok &= bool(np.allclose(all_losses[it], ref_l, rtol=1e-4, atol=1e-6))
        print("iter", it, "max loss diff", np.abs(all_losses[it] - ref_l).max())
    rel = lambda a, b: float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-20))
    for name in ("xyz", "opacity", "scaling", "rotation"):
        e = rel(gm.optimizer.gpu_adam.state[name]["exp_avg"].cpu().numpy(), ref.optimizer.gpu_adam.state[name]["exp_avg"].cpu().numpy())
        print("exp_avg", name, "rel err %.2e" % e); ok &= e < 1e-3
    e = rel(m_all, ref.optimizer.cpu_adam.exp_avg.numpy()); print("exp_avg sh rel err %.2e" % e); ok &= e < 1e-3
    well = np.abs(ref.optimizer.cpu_adam.exp_avg.numpy()) > 1e-2 * np.abs(ref.optimizer.cpu_adam.exp_avg.numpy()).max()
    d = np.abs(sh_all - ref._parameters.numpy())[well].max() / np.abs(ref._parameters.numpy() - g["sh"]).max()
    print("sh param update diff (well-conditioned) %.2e" % d); ok &= d < 5e-3
    print("DIST PARITY", "OK" if ok else "FAILED", "world", world, "peers rows", gm.last_batch_stats["rows_from_peers"])
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
