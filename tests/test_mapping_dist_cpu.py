"""N>1 path of the mapping window on CPU: two gloo ranks, keyframes sharded rank-sticky, ONE
all-reduce of the packed gradient arena, replicated optimizer step.

There is no CPU rasterizer (the product has no CPU fallback), so the renderer is a stand-in
differentiable function of the same parameters; what is under test is the host logic the GPU
path shares: shard_views, the arena layout, the single all-reduce, the rank-0-only regulariser
and the equality "N ranks == 1 rank" of the parameters after several steps."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from monogs_dev_b200 import mapping, synthetic as syn  # noqa: E402

H, W, P, VIEWS, STEPS = 6, 8, 64, 5, 4


def standin_render(cam, pc, pipe=None, bg=None):
    """differentiable toy 'renderer': every parameter group reaches the image, per-view weights"""
    k = float(cam.fx)
    n = H * W
    base = (pc.get_xyz * k).sum(1)[:n] + pc.get_opacity[:n, 0] + pc.get_scaling[:n].sum(1) \
        + pc.get_rotation[:n].sum(1) * 0.1 + pc.get_features[:n, 0].sum(1)
    img = torch.stack([base, 2 * base, 3 * base]).view(3, H, W) * torch.exp(cam.cam_rot_delta.sum())
    depth = (base * 0.5 + cam.cam_trans_delta.sum()).view(1, H, W)
    return {"render": img, "depth": depth, "radii": torch.ones(P, dtype=torch.int32),
            "visibility_filter": torch.ones(P, dtype=torch.bool), "viewspace_points": None,
            "opacity": torch.ones(1, H, W), "n_touched": torch.ones(P, dtype=torch.int32)}


def build(seed=5):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    gm = mapping.GaussianModel(r(P, 3), r(P, 1, 3), r(P, 0, 3), r(P, 1), 0.1 * r(P, 3), r(P, 4))
    # the fused optimizer is CUDA-only: same parameter groups under torch's Adam for this CPU test
    gm.optimizer = torch.optim.Adam([dict(params=[p], lr=gr["lr"], name=gr["name"])
                                     for gr in gm.optimizer.param_groups for p in gr["params"]],
                                    lr=0.0, eps=1e-15)
    kfs = []
    for v in range(VIEWS):
        cam = syn.Camera(W, H, 10.0 + v, 10.0 + v, W / 2, H / 2, device="cpu")
        kfs.append(dict(cam=cam, gt_image=r(3, H, W).abs(), gt_depth=r(1, H, W).abs()))
    return gm, kfs


def run_steps(gm, kfs, group=None):
    win = mapping.MappingWindow(gm, kfs, torch.zeros(3), render_fn=standin_render, adam_cls=torch.optim.Adam,
                                loss_fn=mapping.get_loss_mapping_rgbd_torch, group=group, densify_stats=False)
    losses = [float(win.step()) for _ in range(STEPS)]
    return win, losses


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        gm, kfs = build()
        win, losses = run_steps(gm, kfs)
        assert win.world == world and win.rank == rank
        assert win.local == [v for v in range(VIEWS) if v % world == rank]
        poses = {v: (kfs[v]["cam"].T.clone(), kfs[v]["cam"].exposure_a.detach().clone()) for v in win.local}
        # densification statistics are per-rank (own views) until densify sums / maxes them
        acc, den, rad = gm._stats()
        acc.fill_(rank + 1.0); den.fill_(float(len(win.local))); rad.fill_(10.0 * (rank + 1))
        gm.sync_densification_stats()
        assert float(acc[0]) == sum(r + 1.0 for r in range(world)) and float(den[0]) == VIEWS
        assert float(rad[0]) == 10.0 * world
        torch.save(dict(flat=gm.flat.detach().clone(), losses=losses, local=win.local, poses=poses),
                   os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_views_partition():
    for n in (1, 5, 8, 10, 16):
        for world in (1, 2, 3, 8):
            parts = [mapping.shard_views(n, r, world) for r in range(world)]
            assert sorted(v for p in parts for v in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    assert mapping.shard_views(8, 3, 8) == [3]            # one keyframe per GPU at cfg3 / 8 GPUs
    assert mapping.shard_views(16, 1, 8) == [1, 9]        # two per GPU at cfg5


def test_two_ranks_equal_one_rank(tmp_path):
    gm1, kfs1 = build()
    win1, losses1 = run_steps(gm1, kfs1)
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    outs = [torch.load(os.path.join(tmp_path, f"rank{r}.pt")) for r in range(world)]
    # replicated state: both ranks hold the same parameters, equal to the 1-rank run
    assert torch.allclose(outs[0]["flat"], outs[1]["flat"], rtol=0, atol=0), "replicas diverged"
    assert torch.allclose(outs[0]["flat"], gm1.flat.detach(), rtol=1e-4, atol=1e-5)
    # the job's loss is the sum of the ranks' partial losses (regulariser on rank 0 only)
    for s in range(STEPS):
        assert abs(outs[0]["losses"][s] + outs[1]["losses"][s] - losses1[s]) <= 1e-4 * abs(losses1[s])
    # per-keyframe pose / exposure are optimised by the rank that owns the keyframe
    for o in outs:
        for v, (T, ea) in o["poses"].items():
            assert torch.allclose(T, kfs1[v]["cam"].T, atol=1e-5)
            assert torch.allclose(ea, kfs1[v]["cam"].exposure_a.detach(), atol=1e-5)
