"""Multi-GPU parity script (run under torchrun on >= 2 GPUs; tests/test_gpu_dist.py launches it when the box has them):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tests/dist_parity_check.py [--sparse]

Every rank trains the same small scene: once with the distributed engine (global batch = world x 4 cameras; SH table
node-shared, gradient / moment state sharded by Gaussian id) and, on rank 0, once with the single-GPU engine on the
whole global batch.  The two must agree: same losses, same touched-row counts, Adam moments (= gradients) within
1e-3 relative, parameters of well-conditioned elements within 5e-3 of the update."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "clm-gs_b200")]
import numpy as np
import torch
import torch.distributed as dist

from clmgs import engine, engine_dist, model, scene as S

sparse = "--sparse" in sys.argv
exchange = "nccl" if "--nccl" in sys.argv else "peer"   # gradient return: peer-memory accumulate (default) / all-to-all-v
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dev = "cuda:%d" % int(os.environ["LOCAL_RANK"])
dist.init_process_group("nccl", device_id=torch.device(dev))
W, H, n, bl = 160, 96, 40_003, 4        # n not divisible by the world size: the last shard is shorter
B = bl * world
spec = S.SceneSpec("distcheck", n, W, H, B, "aerial", 50.0, 0)
g = S.make_gaussians(spec)
V, K, P = S.make_cameras(spec, B, g["extent"])
gts = S.make_gt_images(spec, B)
args = model.TrainArgs(bsz=B, cpu_adam_threads=2, sparse_adam=sparse, dist_grad_exchange=exchange)
ctx = engine_dist.DistContext(n)
table = ctx.create_shared_table()
table.array[ctx.lo:ctx.hi] = g["sh"][ctx.lo:ctx.hi]
gm = model.GaussianModelCLMOffload(3, dev).create_from_arrays(g["xyz"], g["log_scale"], g["raw_quat"], g["opacity_logit"],
                                                             table.tensor[ctx.lo:ctx.hi])
engine_dist.training_setup_dist(gm, args, ctx)
cams = [model.Camera(V[i], K[i], gts[i], dev, uid=i) for i in range(rank * bl, (rank + 1) * bl)]
comm = torch.cuda.Stream()
gen_l = torch.Generator().manual_seed(100 + rank)
all_losses = []
for it in range(2):
    losses, order, _ = engine_dist.clm_offload_train_one_batch_dist(gm, None, cams, gm.parameters_grad_buffer, None, None,
                                                                    comm, gen_l, ctx)
    l = torch.empty(bl, device=dev)
    l[torch.tensor(order, device=dev)] = torch.stack(losses)     # back to camera order
    gathered = [torch.empty_like(l) for _ in range(world)]
    dist.all_gather(gathered, l)
    all_losses.append(torch.cat(gathered).cpu().numpy())
assert not gm.parameters_grad_buffer.numpy().any(), "owner's gradient table must be left zeroed"


def gather_shards(t):
    parts = [None] * world
    dist.all_gather_object(parts, np.ascontiguousarray(t))
    return np.concatenate(parts)


m_all = gather_shards(gm.optimizer.cpu_adam.exp_avg.numpy())
gpu_m = {k: gather_shards(gm.optimizer.gpu_adam.state[k]["exp_avg"].cpu().numpy()) for k in ("xyz", "opacity", "scaling", "rotation")}
sh_all = table.array.copy()
# replicated parameters must be bit-identical on every rank after the all-gather
for name in ("_xyz", "_opacity", "_scaling", "_rotation"):
    t = getattr(gm, name)
    ref0 = t.clone()
    dist.broadcast(ref0, src=0)
    assert torch.equal(ref0, t), "replicated %s differs between ranks" % name
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
        e = rel(gpu_m[name], ref.optimizer.gpu_adam.state[name]["exp_avg"].cpu().numpy())
        print("exp_avg", name, "rel err %.2e" % e); ok &= e < 1e-3
        pr = getattr(ref, "_" + name).cpu().numpy(); pg = getattr(gm, "_" + name).cpu().numpy()
        p0 = {"xyz": g["xyz"], "opacity": g["opacity_logit"], "scaling": g["log_scale"], "rotation": g["raw_quat"]}[name].reshape(pr.shape)
        mref = np.abs(ref.optimizer.gpu_adam.state[name]["exp_avg"].cpu().numpy())
        well = mref > 1e-2 * mref.max()
        d = np.abs(pg - pr)[well].max() / max(np.abs(pr - p0).max(), 1e-20)
        print("param", name, "update diff (well-conditioned) %.2e" % d); ok &= d < 5e-3
    e = rel(m_all, ref.optimizer.cpu_adam.exp_avg.numpy()); print("exp_avg sh rel err %.2e" % e); ok &= e < 1e-3
    well = np.abs(ref.optimizer.cpu_adam.exp_avg.numpy()) > 1e-2 * np.abs(ref.optimizer.cpu_adam.exp_avg.numpy()).max()
    d = np.abs(sh_all - ref._parameters.numpy())[well].max() / np.abs(ref._parameters.numpy() - g["sh"]).max()
    print("sh param update diff (well-conditioned) %.2e" % d); ok &= d < 5e-3
    if sparse:   # untouched rows are untouched on both sides
        unt = (ref.optimizer.cpu_adam.exp_avg_sq.numpy() == 0).all(1)
        ok &= bool(np.array_equal(sh_all[unt], np.asarray(g["sh"])[unt]))
    print("DIST PARITY", "OK" if ok else "FAILED", "world", world, "sparse", sparse, "exchange",
          "peer" if ctx.symm is not None else "nccl", "rows from peers' ranges",
          gm.last_batch_stats["rows_from_peers"])
flag = torch.tensor([1 if ok else 0], device=dev)
dist.broadcast(flag, src=0)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) else 1)
