"""The mapping schedule with the window sharded over two gloo ranks (monogs_dev_b200/backend.py, SURVEY.md
§8e items 3-5): both ranks take the same decisions on identical replicas — random views, opacity
reset, densify, prune — exchange the per-view n_touched masks and the optimised poses, and end where
the one-rank run ends."""
import copy
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import test_backend_cpu as T  # noqa: E402   (stand-in renderer, scene builder, config)

WINDOW = [4, 2, 1]


def schedule(be):
    """plain x3, opacity reset at 4 | plain, densify at 6 | plain | a prune-mode call"""
    out = []
    be.set_keyframe_optimizers(WINDOW)
    out.append(be.map(WINDOW, iters=4))          # ends on the reset iteration: gaussian_split is the LAST one's
    out.append(be.map(WINDOW, iters=2))          # ends on the densify iteration
    out.append(be.map(WINDOW, iters=1))
    out.append(be.map(WINDOW, prune=True))
    return out


def config():
    cfg = copy.deepcopy(T.CONFIG)
    cfg["Training"].update(gaussian_update_every=6, gaussian_update_offset=0, gaussian_reset=4, gaussian_th=0.0,
                           gaussian_extent=100.0, size_threshold=10 ** 6, monocular=True)
    return cfg


def run(rank=None):
    gm, views = T.build()
    be = T.make_backend(gm, views, config())
    be.initialized = True
    # force some clones / splits at the densify iteration from statistics both ranks share
    flags = schedule(be)
    return be, gm, views, flags


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        be, gm, views, flags = run(rank)
        torch.save(dict(flat=gm.flat.detach().clone(), flags=flags, P=gm._xyz.shape[0],
                        occ={k: v.clone() for k, v in be.occ_aware_visibility.items()},
                        n_obs=gm.n_obs.clone(), kf=gm.unique_kfIDs.clone(),
                        poses={k: (views[k]["cam"].R.clone(), views[k]["cam"].T.clone(),
                                   views[k]["cam"].exposure_a.detach().clone()) for k in views}),
                   os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_two_rank_schedule_equals_one_rank(tmp_path):
    be1, gm1, views1, flags1 = run()
    assert flags1 == [True, True, False, False]
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    outs = [torch.load(os.path.join(tmp_path, f"rank{r}.pt")) for r in range(world)]
    a, b = outs
    assert a["flags"] == b["flags"] == flags1
    assert a["P"] == b["P"] == gm1._xyz.shape[0] != T.P
    assert torch.equal(a["flat"], b["flat"]), "replicas diverged"
    assert torch.allclose(a["flat"], gm1.flat.detach(), rtol=1e-4, atol=1e-5)
    assert torch.equal(a["kf"], b["kf"]) and torch.equal(a["kf"].cpu(), gm1.unique_kfIDs.cpu())
    assert torch.equal(a["n_obs"], b["n_obs"]) and torch.equal(a["n_obs"].cpu(), gm1.n_obs.cpu())
    for k in WINDOW:                              # every rank holds every window view's mask
        assert torch.equal(a["occ"][k], b["occ"][k]) and torch.equal(a["occ"][k], be1.occ_aware_visibility[k])
    for k in views1:                              # poses / exposures optimised by the owner, known to all
        for o in outs:
            R, Tt, ea = o["poses"][k]
            assert torch.allclose(R, views1[k]["cam"].R, atol=1e-5), k
            assert torch.allclose(Tt, views1[k]["cam"].T, atol=1e-5), k
            assert torch.allclose(ea, views1[k]["cam"].exposure_a.detach(), atol=1e-5), k
