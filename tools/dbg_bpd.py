import sys; sys.path.insert(0, '.')
import numpy as np, torch
from foundationpose_b200 import synthetic, predict_pose_refine as pr, utils as U
from oracle import ops as oo
from tests.helpers import centered, hypo_poses, mesh_np
sc = synthetic.make_scene(1, n_points=600000)
mesh, _ = centered(sc)
poses, d = hypo_poses(sc, n=6, jitter_seed=5)
xyz = oo.depth2xyzmap(d, sc['K'])
rgb_t, xyz_t = torch.as_tensor(sc['rgb']).cuda(), torch.as_tensor(xyz).cuda()
mt = U.make_mesh_tensors(mesh)
S = 160
pd = pr.make_crop_data_batch([S, S], torch.as_tensor(poses).cuda(), mesh, rgb_t, None, sc['K'], 1.2, xyz_t, mesh_diameter=sc['diameter'], mesh_tensors=mt, bf16=False)
cp = oo.crop_params(poses, sc['K'], sc['diameter'], 1.2, S, 480, 640)
color, oxyz, _ = oo.render(poses, cp, mesh_np(mesh), S)
A = oo.make_A(color, oxyz, poses, sc['diameter'], 0.001)
g = pd.rgbAs.permute(0, 2, 3, 1).cpu().numpy()
print('rgbAs diff max', np.abs(g - A[..., :3]).max(), 'n diff', (g != A[..., :3]).sum(), 'nonzero', (A[..., :3] != 0).sum(), (g != 0).sum())
params = U.crop_params(torch.as_tensor(poses).cuda(), sc['K'], sc['diameter'], 1.2, S, 480, 640)
A2, _ = U.render_crops(mt, torch.as_tensor(poses).cuda(), params, S, sc['diameter'], 0.001, bf16=False)
print('dense render vs oracle', (A2[..., :6].cpu().numpy() != A).sum(), 'padded vs dense', (pd.AB[:6, :, 4:164, :6] != A2[..., :6]).sum().item())
print('tf equal', np.array_equal(params[:, :9].cpu().numpy(), cp['tf'].reshape(-1, 9)))
