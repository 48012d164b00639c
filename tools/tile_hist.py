"""Per-tile list-length histogram of the cfg3 views — dev tool."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench, helpers
from monogs_dev_b200 import synthetic as syn, mapping
dev = torch.device("cuda", 0)
sc = syn.make_scene(bench.WORKLOAD, device=dev, n_views=bench.N_VIEWS)
kfs, host = bench.build_window(dev, mapping.render_fused, mapping.GaussianModel, None, sc)
gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation)
for i, cam in enumerate([k["cam"] for k in kfs[:3]]):
    pkg = mapping.render(cam, gm, None, sc.bg)
    st = pkg["render"].grad_fn.state
    import math
    from monogs_dev_b200.rasterizer import GaussianRasterizationSettings
    rs = GaussianRasterizationSettings(
        image_height=int(cam.image_height), image_width=int(cam.image_width),
        tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), bg=sc.bg,
        scale_modifier=1.0, viewmatrix=cam.world_view_transform,
        projmatrix=cam.full_proj_transform, projmatrix_raw=cam.projection_matrix,
        sh_degree=gm.active_sh_degree, campos=cam.camera_center, prefiltered=False, debug=False)
    inp = dict(means3D=gm.get_xyz.detach(), shs=gm.get_features.detach(), opacities=gm.get_opacity.detach(),
               scales=gm.get_scaling.detach(), rotations=gm.get_rotation.detach())
    ex = helpers.export_native(rs, inp, st)
    r = ex["ranges"].long()
    n = (r[:, 1] - r[:, 0]).float()
    q = torch.quantile(n, torch.tensor([0.1, 0.5, 0.9, 0.99], device=dev))
    print(f"view {i}: R={st.num_rendered} tiles={n.numel()} mean={n.mean():.0f} max={n.max():.0f} "
          f"q10/50/90/99={[int(v) for v in q]} >2048:{int((n > 2048).sum())} >4096:{int((n > 4096).sum())}")
    R = st.num_rendered
    m = st.binning.view(torch.int32)[:R] & 0xff          # raw list words: id << 8 | block mask
    v84 = sum(((m >> k) & 1).sum() for k in range(8))
    # 8x8 blocks: (col, R) = 8x4 blocks w = 2*(2R)+col and 2*(2R+1)+col
    v88 = sum((((m >> (4 * Rr + c)) | (m >> (4 * Rr + 2 + c))) & 1).sum() for Rr in range(2) for c in range(2))
    v168 = sum((((m >> (2 * r)) | (m >> (2 * r + 1))) & 1).sum() for r in range(4))   # 16x4 rows
    print(f"   visits 8x4: {int(v84)} ({int(v84)/R:.2f}/instance)  8x8: {int(v88)} ({int(v88)/int(v84):.3f} of 8x4)  16x4: {int(v168)} ({int(v168)/int(v84):.3f})  any: {int((m != 0).sum())}")
