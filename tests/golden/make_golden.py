#!/usr/bin/env python
"""Generates tests/golden/*.pt ON A B200 BOX by running the UNMODIFIED reference CUDA
extension (oracle/_ref, built by oracle/build_ref.py) on the seeded synthetic cases.
Sub-sampled (pixels, Gaussians) so the fixtures stay small.

    gpurun -- python tests/golden/make_golden.py gpurun_out/golden
then copy gpurun_out/golden/*.pt into tests/golden/.
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import build_case  # noqa: E402
from oracle import ref_ext  # noqa: E402

CASES = [("tiny", "sh", (0.2, 0.4, 0.1), {}), ("tiny", "precomp", (0.0, 0.0, 0.0), {}),
         ("tiny_sh3", "sh", (0.1, 0.2, 0.3), {}), ("cfg1_plumbing", "sh", (0.0, 0.0, 0.0), {})]


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    assert ref_ext.available(), "reference extension not built / no GPU"
    for name, mode, bg, kw in CASES:
        sc, rs, inp, gc, gd = build_case(name, device="cuda", mode=mode, bg=bg, **kw)
        fwd = ref_ext.forward(rs, **inp)
        bw = ref_ext.backward(rs, fwd, gc, gd, inp["means3D"], shs=inp.get("shs"),
                              colors_precomp=inp.get("colors_precomp"), scales=inp.get("scales"),
                              rotations=inp.get("rotations"), cov3D_precomp=inp.get("cov3D_precomp"))
        torch.cuda.synchronize()
        geom = ref_ext.decode_geom(fwd, sc.P)
        g = torch.Generator().manual_seed(1)
        gi = torch.randperm(sc.P, generator=g)[:2000].sort().values
        pi = torch.randperm(sc.W * sc.H, generator=g)[:4000].sort().values
        rec = dict(case=name, mode=mode, bg=list(bg), kw=kw, gauss_idx=gi, pix_idx=pi,
                   num_rendered=int(fwd["num_rendered"]), radii=fwd["radii"].cpu()[gi],
                   depths=geom["depths"].cpu()[gi], means2D=geom["means2D"].cpu()[gi],
                   n_touched=fwd["n_touched"].cpu()[gi], dL_dtau=bw["dL_dtau"].cpu())
        vis = rec["radii"] > 0
        rec["depths"][~vis] = 0
        rec["means2D"][~vis] = 0
        for k in ("color", "depth", "opacity"):
            rec[k] = fwd[k].cpu().reshape(fwd[k].shape[0], -1)[:, pi]
        keys = ["dL_dmeans3D", "dL_dmeans2D", "dL_dopacity"]
        keys += ["dL_dsh", "dL_dscales", "dL_drotations"] if mode == "sh" else ["dL_dcolors", "dL_dcov3D"]
        for k in keys:
            t = bw[k].cpu()
            rec[k] = t[gi].clone()
            rec[k + "_scale"] = float(t.abs().max())
        path = os.path.join(outdir, f"ref_{name}_{mode}.pt")
        torch.save(rec, path)
        print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
